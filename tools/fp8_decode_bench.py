"""lcb_fp8_gemm at decode shapes (M <= 128 rows: the call streams the e4m3 weight once): one-pass kernel (N / 256 CTAs)
vs the split-K path (lcb_fp8_gemm_ws), with cuBLAS bf16 on the dequantised operands and cuBLASLt fp8 per-tensor beside.
Reported: us per call and weight-streaming GB/s = (N K + M K + 2 M N) / time. Calls rotate through weight copies that
together exceed the L2 and are replayed from one CUDA graph (device time only).

    python tools/fp8_decode_bench.py [--out file]
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lightcompress_b200 import _cabi  # noqa: E402
from lightcompress_b200 import fp8_infer as fi  # noqa: E402

dev = 'cuda:0'


def timeit(fns, rep=5):
    """fns: one callable per weight copy (together > L2). All calls are recorded into ONE CUDA graph (no host time between
    launches), replayed `rep` times; returns ms per call."""
    for f in fns:
        f()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for f in fns:
            f()
    g.replay()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(rep):
        g.replay()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / (rep * len(fns))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--out', default=None)
    args = ap.parse_args()
    lib = _cabi.load()
    lines = []
    for (N, K) in [(7168, 2048), (2048, 7168), (4096, 4096), (1536, 7168), (24576, 1536)]:
        copies = max(2, (160 << 20) // (N * K) + 1)          # the weight copies of one pass exceed the 126 MB L2
        bs = [(torch.randn(N, K, device=dev) * 50).clamp(-448, 448).to(torch.float8_e4m3fn) for _ in range(copies)]
        b_s = torch.rand((N + 127) // 128, (K + 127) // 128, device=dev) * 0.01 + 1e-3
        xbs = [b.bfloat16() for b in bs[:max(2, copies // 2)]]
        for M in (1, 16, 64, 128):
            a = (torch.randn(M, K, device=dev) * 50).clamp(-448, 448).to(torch.float8_e4m3fn)
            a_s = torch.rand(M, (K + 127) // 128, device=dev) * 0.01 + 1e-3
            c = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
            ws = torch.zeros(max(16, int(lib.lcb_fp8_gemm_workspace_bytes(M, N, K))), dtype=torch.uint8, device=dev)

            def one_pass(b):
                def f():
                    assert lib.lcb_fp8_gemm(a.data_ptr(), a_s.data_ptr(), b.data_ptr(), b_s.data_ptr(), c.data_ptr(), None, M, N,
                                            K, 2, 0, torch.cuda.current_stream().cuda_stream) == 0
                return f

            def split_k(b):
                def f():
                    assert lib.lcb_fp8_gemm_ws(a.data_ptr(), a_s.data_ptr(), b.data_ptr(), b_s.data_ptr(), c.data_ptr(), None, M,
                                               N, K, 2, 0, ws.data_ptr(), torch.cuda.current_stream().cuda_stream) == 0
                return f
            ms1 = timeit([one_pass(b) for b in bs])
            ms2 = timeit([split_k(b) for b in bs])
            xa = a.bfloat16()
            ms_cublas = timeit([(lambda xb=xb: torch.nn.functional.linear(xa, xb)) for xb in xbs])
            ms_lt = None
            if M % 16 == 0:
                one = torch.ones((), device=dev)
                try:
                    ms_lt = timeit([(lambda b=b: torch._scaled_mm(a, b.t(), one, one, out_dtype=torch.bfloat16)) for b in bs])
                except Exception:
                    ms_lt = None
            by = N * K + M * K + 2 * M * N
            lines.append({'M': M, 'N': N, 'K': K, 'one_pass_us': round(ms1 * 1e3, 2), 'split_k_us': round(ms2 * 1e3, 2),
                          'speedup': round(ms1 / ms2, 2), 'split_k_GBps': round(by / ms2 / 1e6, 1),
                          'cublas_bf16_us': round(ms_cublas * 1e3, 2),
                          'cublaslt_fp8_per_tensor_us': None if ms_lt is None else round(ms_lt * 1e3, 2)})
            print(json.dumps(lines[-1]), flush=True)
        del bs, xbs
        torch.cuda.empty_cache()
    if args.out:
        with open(args.out, 'w') as f:
            for ln in lines:
                f.write(json.dumps(ln) + '\n')


if __name__ == '__main__':
    main()
