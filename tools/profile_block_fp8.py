"""Driver for ncu captures of the per_block FP8 path (DeepSeek-style 128 x 128 weight blocks): fake-quant and real-quant
of a 14336 x 4096 bf16 weight, 3 launches each."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lightcompress_b200.quant import FloatQuantizer  # noqa: E402

dev = 'cuda:0'
w = (torch.randn(14336, 4096, device=dev) * 0.02).to(torch.bfloat16)
q = FloatQuantizer('e4m3', True, 'per_block', block_size=128, use_qtorch=True)
for _ in range(3):
    y = q.fake_quant_weight_dynamic(w)
for _ in range(3):
    c, s, _ = q.real_quant_weight_dynamic(w)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for name, fn, by in (('fake', lambda: q.fake_quant_weight_dynamic(w), 4), ('real', lambda: q.real_quant_weight_dynamic(w), 3)):
    e0.record()
    for _ in range(10):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(name, round(ms, 4), 'ms', round(w.numel() * by / ms / 1e6, 1), 'GB/s')

# MoE experts (DeepSeek-V2-Lite: 66 experts x gate/up [1408, 2048] + down [2048, 1408]): per-tensor calls vs 16 per launch
ex = []
for _ in range(66):
    ex += [(torch.randn(1408, 2048, device=dev) * 0.02).to(torch.bfloat16), (torch.randn(1408, 2048, device=dev) * 0.02).to(torch.bfloat16),
           (torch.randn(2048, 1408, device=dev) * 0.02).to(torch.bfloat16)]
nel = sum(w.numel() for w in ex)
for name, fn, by in (('experts fake per-tensor', lambda: [q.fake_quant_weight_dynamic(w) for w in ex], 4),
                     ('experts fake multi', lambda: q.fake_quant_weights_dynamic(ex), 4),
                     ('experts real per-tensor', lambda: [q.real_quant_weight_dynamic(w) for w in ex], 3),
                     ('experts real multi', lambda: q.real_quant_weights_dynamic(ex), 3)):
    fn()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(3):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    print(name, round(ms, 3), 'ms', round(nel * by / ms / 1e6, 1), 'GB/s')
