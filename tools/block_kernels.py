"""Per-kernel device time of ONE steady-state block calibration through the plug-ins (torch.profiler, CUDA activities):
which kernels the end-to-end s/block of tools/block_bench.py is made of — the plug-in's own kernels against the library
kernels of the hooked HF block forwards.

    python tools/block_kernels.py gptq|awq [--samples N] > profiles/r2_<method>_block_kernels.txt
"""
import os
import sys

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import block_bench as bb  # noqa: E402
from lightcompress_b200 import algo  # noqa: E402

method = sys.argv[1] if len(sys.argv) > 1 else 'gptq'
samples = int(sys.argv[sys.argv.index('--samples') + 1]) if '--samples' in sys.argv else 128
dev = torch.device('cuda:0')
if method == 'gptq':
    name, seq, qo = 'GPTQ', 2048, True
    sp = {'actorder': True, 'static_groups': False, 'percdamp': 0.01, 'blocksize': 128, 'true_sequential': True}
else:
    name, seq, qo = 'Awq', 512, False
    sp = {'trans': True, 'trans_version': 'v2', 'weight_clip': True, 'clip_sym': False}
model = bb.one_block_llama(dev, n_blocks=2)
g = torch.Generator().manual_seed(1)
bs = samples if method == 'awq' else 1   # awq_w_only.yml: bs -1 (one batch); gptq_w_only.yml: bs 1
data = [{'input_ids': torch.randint(0, 1024, (bs, seq), generator=g).to(dev)} for _ in range(samples // bs)]
ad = algo.adapters.LlamaAdapter(model)
inp = ad.collect_first_block_input(data)
a = algo.ALGO_REGISTRY[name](ad, {'weight': dict(bb.W4), 'special': sp, 'quant_out': qo}, inp, None, {'calib': {'seq_len': seq}})
a.block_idx = 0
a.block_opt(a.blocks[0])           # warm: allocator, handles, attributes
torch.cuda.synchronize()
a.block_idx = 1
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    a.block_opt(a.blocks[1])
    torch.cuda.synchronize()
ev = [e for e in prof.key_averages() if e.device_time_total > 0 and e.device_type.name == 'CUDA'] or \
     [e for e in prof.key_averages() if e.device_time_total > 0]
ev.sort(key=lambda e: -e.device_time_total)
total = sum(e.device_time_total for e in ev)
print(f'{name}: one steady-state Llama-3-8B-sized block, {samples} x {seq} tokens; total device time {total / 1e6:.4f} s')
print(f'{"device ms":>10} {"calls":>7} {"share":>6}  kernel')
for e in ev[:40]:
    print(f'{e.device_time_total / 1e3:10.2f} {e.count:7d} {e.device_time_total / total:6.1%}  {e.key[:150]}')
