"""Small driver for ncu captures of the FP8 inference kernels: 3 act_quant launches (8192 x 7168 bf16) and 3 fp8_gemm
launches (8192 x 2048 x 7168), nothing else on the stream."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lightcompress_b200 import fp8_infer as fi  # noqa: E402

dev = 'cuda:0'
x = torch.randn(8192, 7168, device=dev, dtype=torch.bfloat16)
b = (torch.randn(2048, 7168, device=dev) * 50).clamp(-448, 448).to(torch.float8_e4m3fn)
b_s = torch.rand(16, 56, device=dev) * 0.01 + 1e-3
for _ in range(3):
    a, a_s = fi.act_quant(x)
for _ in range(3):
    c = fi.fp8_gemm(a, a_s, b, b_s, out_dtype=torch.bfloat16)
torch.cuda.synchronize()
print('ok', float(c.float().abs().mean()))
