import sys, time, torch
sys.path.insert(0, '.')
from lightcompress_b200 import ops
for n in (4096, 14336):
    g = torch.Generator(device='cuda').manual_seed(0)
    x = torch.randn(8192, n, device='cuda', generator=g)
    S = torch.zeros(n, n, device='cuda'); ops.hessian_accum(S, x.to(torch.bfloat16)); H = ops.hessian_finalize(S, 1)
    Hc = H.clone()
    for splits in (2, 3):
        ops.cholinv_upper(Hc.copy_(H), 0.01, splits=splits, check=False)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ops.cholinv_upper(Hc, 0.01, splits=splits, check=False)
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        print(n, splits, 'host_ms', round((t1 - t0) * 1e3, 3), 'total_ms', round((t2 - t0) * 1e3, 3), flush=True)
