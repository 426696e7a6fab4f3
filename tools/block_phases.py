"""Wall time per phase (synchronised) of one end-to-end GPTQ / Awq block calibration."""
import os
import sys
import time
from collections import defaultdict

import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import block_bench as bb  # noqa: E402
from lightcompress_b200.algo import blockwise as bw  # noqa: E402

acc = defaultdict(float)


def timed(cls, name):
    f = getattr(cls, name)

    def g(*a, **k):
        torch.cuda.synchronize()
        t = time.perf_counter()
        r = f(*a, **k)
        torch.cuda.synchronize()
        acc[name] += time.perf_counter() - t
        return r
    setattr(cls, name, g)


method = sys.argv[1] if len(sys.argv) > 1 else 'gptq'
dev = torch.device('cuda:0')
from lightcompress_b200.algo import algos  # noqa: E402
from lightcompress_b200.algo import gptq as lcgptq  # noqa: E402
for c, n in ((bw.BaseBlockwiseQuantization, 'block_forward'), (algos.GPTQ, 'subset_transform'),
             (algos.Awq, 'subset_transform'), (bw.BaseBlockwiseQuantization, 'rehook_next_subset'),
             (bw.BaseBlockwiseQuantization, '_forward_fake_quant'), (lcgptq.HessianState, 'factor'),
             (lcgptq.HessianState, 'finalize'), (algos.GPTQ, 'w_qdq')):
    timed(c, n)
if method == 'awq':
    from lightcompress_b200.algo import autoclip
    timed(autoclip, 'auto_clip_block')
if method == 'gptq':
    sp = {'actorder': True, 'static_groups': False, 'percdamp': 0.01, 'blocksize': 128, 'true_sequential': True}
    bb.run('GPTQ', dev, 4, 2048, sp, True)
    acc.clear()
    t = bb.run('GPTQ', dev, 128, 2048, sp, True)
else:
    sp = {'trans': True, 'trans_version': 'v2', 'weight_clip': True, 'clip_sym': False}
    bb.run('Awq', dev, 2, 512, sp, bs=-1)
    acc.clear()
    t = bb.run('Awq', dev, 128, 512, sp, bs=-1)
print('wall per block', [round(v, 4) for v in t], '(phase sums cover BOTH blocks)', {k: round(v, 4) for k, v in acc.items()})
# nested: block_forward is inside rehook_next_subset / _forward_fake_quant; factor / finalize / w_qdq inside the others
