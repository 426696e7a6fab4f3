"""Where does lcb_gptq_weight_transform spend its time: host enqueue vs GPU execution (q_proj shape)."""
import os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lightcompress_b200 import ops
dev = torch.device('cuda:0')
R = C = 4096
g = torch.Generator(device=dev).manual_seed(0)
x = torch.randn(8192, C, device=dev, generator=g).to(torch.bfloat16)
S = torch.zeros(C, C, device=dev)
ops.hessian_accum(S, x); ops.hessian_finalize(S, 4)
d = torch.arange(C, device=dev); S[d, d] += 0.01 * torch.mean(torch.diag(S))
Hinv = torch.linalg.cholesky(torch.cholesky_inverse(torch.linalg.cholesky(S)), upper=True).contiguous()
W = torch.randn(R, C, device=dev, generator=g) * 0.02
ops.gptq_weight_transform(W.clone(), Hinv, 128, 128, 0.0, 15.0, False)
for _ in range(3):
    Wc = W.clone(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    ops.gptq_weight_transform(Wc, Hinv, 128, 128, 0.0, 15.0, False)
    t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
    print(f'enqueue {1e3*(t1-t0):.3f} ms, total {1e3*(t2-t0):.3f} ms')
