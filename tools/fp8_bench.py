"""Bench of the block-wise FP8 inference kernels (csrc/fp8_infer.cu): act_quant GB/s, fp8_gemm TFLOP/s, the
LlmcFp8Linear forward; cuBLAS bf16 matmul of the same shape timed beside it. One JSON line per case.
  python tools/fp8_bench.py [--out profiles/rN_fp8_bench.jsonl]"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lightcompress_b200 import fp8_infer as fi  # noqa: E402


def timeit(fn, iters=20, warm=5, flush=None):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush is not None:
            flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--out', default=None)
    ap.add_argument('--quick', action='store_true')
    args = ap.parse_args()
    dev = 'cuda:0'
    peaks = json.load(open(os.path.join(os.path.dirname(__file__), '..', 'MEASURED_PEAKS.json')))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    lines = []
    for (T, C) in [(8192, 7168), (32768, 4096), (4096, 2048)]:
        x = torch.randn(T, C, device=dev, dtype=torch.bfloat16)
        ms = timeit(lambda: fi.act_quant(x), flush=flush)
        by = x.numel() * 3 + x.numel() // 128 * 4
        lines.append({'kernel': 'act_quant', 'shape': [T, C], 'ms': ms, 'GBps': by / ms / 1e6,
                      'frac_of_hbm_peak': by / ms / 1e6 / peaks['hbm_gbs']})
    shapes = [(4096, 4096, 4096), (8192, 7168, 2048), (8192, 2048, 7168), (4096, 1408, 2048), (4096, 2048, 1408),
              (8192, 8192, 8192), (128, 7168, 2048), (16, 4096, 4096)]
    if args.quick:
        shapes = shapes[:2]
    for (M, N, K) in shapes:
        a = (torch.randn(M, K, device=dev) * 50).clamp(-448, 448).to(torch.float8_e4m3fn)
        b = (torch.randn(N, K, device=dev) * 50).clamp(-448, 448).to(torch.float8_e4m3fn)
        a_s = torch.rand(M, (K + 127) // 128, device=dev) * 0.01 + 1e-3
        b_s = torch.rand((N + 127) // 128, (K + 127) // 128, device=dev) * 0.01 + 1e-3
        ms = timeit(lambda: fi.fp8_gemm(a, a_s, b, b_s, out_dtype=torch.bfloat16), flush=flush)
        xa, xb = a.bfloat16(), b.bfloat16()
        ms_cublas = timeit(lambda: torch.nn.functional.linear(xa, xb), flush=flush)
        try:   # cuBLASLt fp8 with per-tensor scales: the library's fp8 tensor-pipe rate on this chip (not the same op)
            one = torch.ones((), device=dev)
            ms_lt = timeit(lambda: torch._scaled_mm(a, b.t(), one, one, out_dtype=torch.bfloat16), flush=flush)
        except Exception:
            ms_lt = None
        x = torch.randn(M, K, device=dev, dtype=torch.bfloat16)
        ms_fwd = timeit(lambda: fi.block_wise_fp8_forward_func(x, b, b_s, 128, None), flush=flush)
        fl = 2.0 * M * N * K
        lines.append({'kernel': 'fp8_gemm', 'M': M, 'N': N, 'K': K, 'ms': ms, 'TFLOPs': fl / ms / 1e9,
                      'frac_of_fp8_nominal_4500': fl / ms / 1e9 / 4500.0, 'cublas_bf16_ms': ms_cublas,
                      'cublas_bf16_TFLOPs': fl / ms_cublas / 1e9, 'forward_func_ms': ms_fwd,
                      'cublaslt_fp8_per_tensor_TFLOPs': None if ms_lt is None else fl / ms_lt / 1e9,
                      'weight_bytes_GBps': (N * K + M * K + 2 * M * N) / ms / 1e6})
    for ln in lines:
        print(json.dumps(ln))
    if args.out:
        with open(args.out, 'w') as f:
            for ln in lines:
                f.write(json.dumps(ln) + '\n')


if __name__ == '__main__':
    main()
