"""Accuracy and time of lcb_cholinv_upper against the reference's cuSOLVER chain (gptq.py:165-174).

    python tools/cholinv_bench.py [n ...]         (default: 4096 14336; Llama-3-8B Hessian sizes)

Truth is the same chain in float64. err(X) = ||X - U64||_F / ||U64||_F.
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lightcompress_b200 import ops  # noqa: E402


def make_h(n, dev, tokens=8192, seed=0):
    g = torch.Generator(device=dev).manual_seed(seed)
    gain = torch.exp(0.7 * torch.randn(n, device=dev, generator=g))
    gain[torch.randint(0, n, (8,), device=dev, generator=g)] *= 30.0           # outlier channels
    mix = torch.randn(64, n, device=dev, generator=g) / 8.0                      # low-rank correlated part
    S = torch.zeros(n, n, device=dev)
    for _ in range(max(1, tokens // 4096)):
        x = torch.randn(4096, n, device=dev, generator=g) + torch.randn(4096, 64, device=dev, generator=g) @ mix
        ops.hessian_accum(S, (x * gain).to(torch.bfloat16))
    return ops.hessian_finalize(S, max(1, tokens // 4096))


def chain(H, percdamp, dtype):
    Hd = H.to(dtype).clone()
    d = torch.arange(H.shape[0], device=H.device)
    Hd[d, d] += percdamp * torch.mean(torch.diag(Hd))
    L = torch.linalg.cholesky(Hd)
    return torch.linalg.cholesky(torch.cholesky_inverse(L), upper=True)


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


def main():
    dev = torch.device('cuda:0')
    sizes = [int(a) for a in sys.argv[1:] if a.isdigit()] or [4096, 14336]
    for n in sizes:
        H = make_h(n, dev)
        U64 = chain(H, 0.01, torch.float64)
        nrm = torch.linalg.norm(U64)
        U32 = chain(H, 0.01, torch.float32)
        row = {'n': n, 'cond_est': float(torch.linalg.cond(H.double() + 0.01 * H.diag().double().mean() * torch.eye(n, device=dev, dtype=torch.float64))) if n <= 4096 else None,
               'err_torch_fp32': float(torch.linalg.norm(U32.double() - U64) / nrm)}
        del U32
        for splits in (2, 3):
            U = ops.cholinv_upper(H.clone(), 0.01, splits=splits)
            row[f'err_ours_splits{splits}'] = float(torch.linalg.norm(U.double() - U64) / nrm)
            row[f'lower_zero_splits{splits}'] = bool((torch.tril(U, -1) == 0).all())
            del U
        Hc = H.clone()
        row['ms_torch_chain_fp32'] = ev(lambda: chain(H, 0.01, torch.float32), 2)
        for splits in (2, 3):
            row[f'ms_ours_splits{splits}'] = ev(lambda: ops.cholinv_upper(Hc.copy_(H), 0.01, splits=splits, check=False), 3)
        row['ms_copy'] = ev(lambda: Hc.copy_(H), 3)
        print(json.dumps(row), flush=True)
        del H, U64, Hc
        torch.cuda.empty_cache()


if __name__ == '__main__':
    main()
