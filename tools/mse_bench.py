"""calib_algo='mse' range search: kernel time against the reference's torch loop on the same GPU."""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lightcompress_b200.quant import IntegerQuantizer  # noqa: E402


def torch_mse_range(t, sym, qmin, qmax, maxshrink=0.8, grid=100, norm=2.4):
    """The reference arithmetic (quant.py:145-203) as torch ops on the GPU — the baseline being replaced."""
    t = t.float()
    mn, mx = t.amin(-1, keepdim=True), t.amax(-1, keepdim=True)
    best = torch.full([t.shape[0]], float('inf'), device=t.device)
    for i in range(int(maxshrink * grid)):
        p = 1 - i / grid
        xmin, xmax = p * mn, p * mx
        if sym:
            s = torch.max(xmax.abs(), xmin.abs()).clamp(min=1e-5) / qmax
            z = 0
        else:
            s = (xmax - xmin).clamp(min=1e-5) / (qmax - qmin)
            z = (qmin - torch.round(xmin / s)).clamp(qmin, qmax)
        q = (torch.clamp(torch.round(t / s) + z, qmin, qmax) - z) * s
        q -= t
        q.abs_()
        q.pow_(norm)
        err = torch.sum(q, 1)
        tmp = err < best
        if torch.any(tmp):
            best[tmp] = err[tmp]
            mn[tmp] = xmin[tmp]
            mx[tmp] = xmax[tmp]
    return mn, mx


def timeit(f, n=3):
    f()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        f()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3


def main():
    dev = torch.device('cuda:0')
    torch.manual_seed(0)
    for name, shape, gran, kw, sym in [
        ('w4 g128 gate_proj 14336x4096', (14336, 4096), 'per_group', {'group_size': 128}, False),
        ('w4 per_channel q_proj 4096x4096', (4096, 4096), 'per_channel', {}, True),
        ('w4 per_channel down_proj 4096x14336', (4096, 14336), 'per_channel', {}, True),
    ]:
        w = (torch.randn(shape, device=dev) * 0.02).to(torch.bfloat16)
        q = IntegerQuantizer(4, sym, gran, calib_algo='mse', **kw)
        L = kw.get('group_size', shape[1])
        qmin, qmax = (-8.0, 7.0) if sym else (0.0, 15.0)
        ours = timeit(lambda: q.get_mse_range(w))
        ref = timeit(lambda: torch_mse_range(w.reshape(-1, L), sym, qmin, qmax), n=1)
        a, b = q.get_mse_range(w)
        c, d = torch_mse_range(w.reshape(-1, L), sym, qmin, qmax)
        agree = float(((a == c) & (b == d)).float().mean())
        print(json.dumps({'case': name, 'lcb_mse_range_ms': round(ours, 3), 'torch_loop_ms': round(ref, 2),
                          'speedup': round(ref / ours, 1), 'units_equal_to_torch_cuda': round(agree, 4),
                          'elem_steps_per_s': round(w.numel() * 80 / ours / 1e6, 1)}), flush=True)


if __name__ == '__main__':
    main()
