"""Small fixed workload for ncu: two fused AWQ candidates (scale, fake-quant, tcgen05 linear + MSE epilogue), one lcb_linear
(the projection GEMM of the multi-linear inspect modules) and one AutoClip search, at a Llama-3-8B o_proj shape."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lightcompress_b200 import ops  # noqa: E402
from lightcompress_b200.quant import IntegerQuantizer  # noqa: E402

dev = torch.device('cuda:0')
R = C = 4096
T = 8192
g = torch.Generator(device=dev).manual_seed(0)
x = torch.randn(T, C, device=dev, generator=g).to(torch.bfloat16)
W = (torch.randn(R, C, device=dev, generator=g) * 0.02).to(torch.bfloat16)
wq = IntegerQuantizer(4, False, 'per_group', group_size=128)
org = ops.linear(x, W)
cand = (torch.rand(2, C, device=dev, generator=g) + 0.5).to(torch.bfloat16)
ops.awq_candidates(W, x, org, cand, wq)
ops.linear(x, W)
ops.autoclip_search(W, x[:512], wq, False)
torch.cuda.synchronize()
print('ok')
