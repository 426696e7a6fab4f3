"""FloatQuantizer without qtorch (per-element scales, csrc/floatq.cu) at Llama-3-8B call sizes: the product kernel vs the
reference's op chain (the same arithmetic as torch ops on the GPU: quant.py:1005-1027, 1061-1076).

    python tools/floatq_bench.py
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lightcompress_b200.quant import FloatQuantizer  # noqa: E402

dev = torch.device('cuda:0')


def torch_chain(x, e_bits, m_bits):
    """the reference's ops, verbatim order (test_floatq_gpu.py pins the product to the real class; this is the timing twin)"""
    max_val, min_val = x.amax(dim=-1, keepdim=True), x.amin(dim=-1, keepdim=True)
    maxval = torch.max(max_val, -min_val)
    e = torch.tensor(e_bits, dtype=torch.float32, device=x.device)
    m = torch.tensor(m_bits, dtype=torch.float32, device=x.device)
    if e >= 5:
        maxval = maxval.to(dtype=torch.float32)
    bias = 2 ** e - torch.log2(maxval) + torch.log2(2 - 2 ** (-m)) - 1
    xc = torch.min(torch.max(x, -maxval), maxval)
    ls = torch.clamp(torch.floor(torch.log2(torch.abs(xc)) + bias), 1.0)
    s = 2.0 ** (ls - m - bias)
    s[s == 0] = 1
    return ((torch.round(xc / s + 0) - 0) * s).to(x.dtype)


def ev(fn, rep=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(rep):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e-3 / rep


g = torch.Generator(device=dev).manual_seed(0)
for bit in ('e4m3', 'e5m2'):
    for name, shape, gran in (('act 8x2048x4096 per_token', (8, 2048, 4096), 'per_token'),
                              ('gate_proj 14336x4096 per_channel', (14336, 4096), 'per_channel'),
                              ('down_proj 4096x14336 per_channel', (4096, 14336), 'per_channel')):
        x = (torch.randn(*shape, device=dev, generator=g) * 0.05).to(torch.bfloat16)
        q = FloatQuantizer(bit, True, gran)
        fn = q.fake_quant_act_dynamic if gran == 'per_token' else q.fake_quant_weight_dynamic
        got, ref = fn(x), torch_chain(x, q.e_bits, q.m_bits)
        same = bool(torch.equal(got.view(torch.int16), ref.view(torch.int16)))
        t, tr = ev(lambda: fn(x)), ev(lambda: torch_chain(x, q.e_bits, q.m_bits), rep=3)
        nb = x.numel() * 4
        print(json.dumps({'bit': bit, 'what': name, 'ms': round(t * 1e3, 4), 'GBps': round(nb / t / 1e9, 1),
                          'torch_ops_ms': round(tr * 1e3, 3), 'speedup': round(tr / t, 1), 'bit_equal_to_torch_ops': same}),
              flush=True)
