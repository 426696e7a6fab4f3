"""Time the tcgen05 Hessian kernel at Llama-3-8B shapes against the reference's own op (fp32 matmul, gptq.py:288-290)
and a bf16 cuBLAS matmul of the same shape.

FLOP convention: 2*T*C^2 per call (full GEMM). The kernel is a SYRK (upper blocks only), so its
hardware FLOPs are about half of that; both figures are printed.
"""
import json
import math
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lightcompress_b200 import ops  # noqa: E402


def timeit(fn, rep):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(rep):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e-3 / rep


def main():
    dev = torch.device('cuda:0')
    quick = '--quick' in sys.argv
    shapes = [(4096, 2048), (4096, 8192), (4096, 32768), (14336, 2048), (14336, 8192)]
    if not quick:
        shapes += [(8192, 8192), (14336, 32768)]
    for C, T in shapes:
        x = torch.randn(T, C, device=dev, dtype=torch.bfloat16)
        S = torch.zeros(C, C, device=dev)
        rep = max(3, min(50, int(2e13 / (2.0 * T * C * C))))
        t = timeit(lambda: ops.hessian_accum(S, x), rep)
        flops = 2.0 * T * C * C
        out = {'C': C, 'T': T, 'ms': round(t * 1e3, 4), 'gemm_equiv_TFLOPs': round(flops / t / 1e12, 1),
               'syrk_hw_TFLOPs': round(flops / t / 1e12 * (C / 128 + 1) / (2 * C / 128), 1)}
        if C <= 8192:
            H = torch.zeros(C, C, device=dev)
            xt = x.t()

            def ref():
                c = math.sqrt(2 / 3) * xt.float()
                H.add_(c.matmul(c.t()))
            tr = timeit(ref, max(2, rep // 10))
            out['ref_fp32_matmul_ms'] = round(tr * 1e3, 3)
            out['speedup_vs_ref'] = round(tr / t, 1)
        xb = x.t().contiguous()
        Hb = torch.empty(C, C, device=dev, dtype=torch.bfloat16)
        tb = timeit(lambda: torch.matmul(xb, xb.t(), out=Hb), rep)
        out['cublas_bf16_gemm_ms'] = round(tb * 1e3, 4)
        out['cublas_bf16_TFLOPs'] = round(flops / tb / 1e12, 1)
        print(json.dumps(out), flush=True)
        del x, S, xb, Hb
        torch.cuda.empty_cache()


if __name__ == '__main__':
    main()
