"""AutoClip v1 search time at Llama-3-8B shapes (512 sampled tokens, w4 g128) and agreement with the pinned oracle.

    python tools/autoclip_bench.py
"""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lightcompress_b200 import ops  # noqa: E402
from lightcompress_b200.quant import IntegerQuantizer  # noqa: E402


def ev(fn, rep):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(rep):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / rep


def reference_like(W, x, wq, n_grid=20, steps=10):
    """The reference's arithmetic on the GPU in torch (auto_clip.py:84-191), for a time comparison only."""
    oc, C = W.shape
    gs = wq.group_size
    w_all = W.reshape(oc, 1, -1, gs)
    xx = x.reshape(1, x.shape[0], -1, gs)
    for i_b in range(oc // 256):
        w = w_all[i_b * 256:(i_b + 1) * 256]
        org_max, org_min = w.amax(-1, keepdim=True), w.amin(-1, keepdim=True)
        min_errs = torch.ones_like(org_max) * 1e9
        bmax, bmin = org_max.clone(), org_min.clone()
        org_out = (xx * w).sum(-1)
        for i_s in range(steps):
            mx, mn = org_max * (1 - i_s / n_grid), org_min * (1 - i_s / n_grid)
            q = wq.fake_quant_weight_dynamic(torch.clamp(w, mn, mx).reshape(256, C)).reshape(w.shape)
            err = ((xx * q).sum(-1) - org_out).pow(2).mean(1).view(min_errs.shape)
            idx = err < min_errs
            min_errs[idx], bmax[idx], bmin[idx] = err[idx], mx[idx], mn[idx]


def main():
    dev = torch.device('cuda:0')
    wq = IntegerQuantizer(4, False, 'per_group', group_size=128)
    g = torch.Generator(device=dev).manual_seed(0)
    for name, R, C in (('o_proj', 4096, 4096), ('gate/up_proj', 14336, 4096), ('down_proj', 4096, 14336)):
        W = (torch.randn(R, C, device=dev, generator=g) * 0.02).to(torch.bfloat16)
        x = torch.randn(512, C, device=dev, generator=g).to(torch.bfloat16)
        ms = ev(lambda: ops.autoclip_search(W, x, wq, False), 3)
        row = {'linear': name, 'R': R, 'C': C, 'tokens': 512, 'autoclip_ms': round(ms, 3),
               'gemm_equiv_TFLOPs': round(11 * 2.0 * R * C * 512 / (ms * 1e-3) / 1e12, 1)}
        if name == 'o_proj':
            t0 = time.perf_counter()
            reference_like(W, x, wq)
            torch.cuda.synchronize()
            row['torch_reference_arithmetic_ms'] = round((time.perf_counter() - t0) * 1e3, 1)
        print(json.dumps(row), flush=True)


if __name__ == '__main__':
    main()
