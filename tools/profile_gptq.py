"""Small fixed workload for ncu: 2 Hessian batches + one blocked GPTQ update at a Llama-3-8B q_proj shape."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lightcompress_b200 import ops  # noqa: E402

dev = torch.device('cuda:0')
R = C = 4096
g = torch.Generator(device=dev).manual_seed(0)
xs = [(torch.randn(2048, C, device=dev, generator=g)).to(torch.bfloat16) for _ in range(2)]
S = torch.zeros(C, C, device=dev)
for x in xs:
    ops.hessian_accum(S, x)
ops.hessian_accum(S, torch.cat(xs + xs, 0))
ops.hessian_finalize(S, 6)
d = torch.arange(C, device=dev)
S[d, d] += 0.01 * torch.mean(torch.diag(S))
Hinv = torch.linalg.cholesky(torch.cholesky_inverse(torch.linalg.cholesky(S)), upper=True).contiguous()
W = torch.randn(R, C, device=dev, generator=g) * 0.02
ops.gptq_weight_transform(W, Hinv, 128, 128, 0.0, 15.0, False)
torch.cuda.synchronize()
print('ok')
