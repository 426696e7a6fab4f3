"""torch.profiler kernel table of one end-to-end block calibration (tools/block_bench.py): where the wall time goes."""
import os
import sys

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import block_bench as bb  # noqa: E402

method = sys.argv[1] if len(sys.argv) > 1 else 'gptq'
dev = torch.device('cuda:0')
if method == 'gptq':
    sp = {'actorder': False, 'static_groups': False, 'percdamp': 0.01, 'blocksize': 128, 'true_sequential': True}
    args = ('GPTQ', dev, 128, 2048, sp)
else:
    sp = {'trans': True, 'trans_version': 'v2', 'weight_clip': True, 'clip_sym': False}
    args = ('Awq', dev, 128, 512, sp)
bb.run(args[0], dev, 2, args[3], sp)
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    t = bb.run(*args)
print('wall', t)
print(prof.key_averages().table(sort_by='cuda_time_total', row_limit=30, max_name_column_width=70))
