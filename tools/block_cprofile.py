"""cProfile of one end-to-end block calibration: host-side cost of the block loop."""
import cProfile
import os
import pstats
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import block_bench as bb  # noqa: E402

method = sys.argv[1] if len(sys.argv) > 1 else 'gptq'
dev = torch.device('cuda:0')
if method == 'gptq':
    sp = {'actorder': False, 'static_groups': False, 'percdamp': 0.01, 'blocksize': 128, 'true_sequential': True}
    a = ('GPTQ', dev, 128, 2048, sp)
else:
    sp = {'trans': True, 'trans_version': 'v2', 'weight_clip': True, 'clip_sym': False}
    a = ('Awq', dev, 128, 512, sp)
bb.run(a[0], dev, 2, a[3], sp)
pr = cProfile.Profile()
pr.enable()
bb.run(*a)
pr.disable()
pstats.Stats(pr).sort_stats('cumulative').print_stats(45)
