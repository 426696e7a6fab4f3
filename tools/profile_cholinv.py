import sys, torch
sys.path.insert(0, '.')
from lightcompress_b200 import ops
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
g = torch.Generator(device='cuda').manual_seed(0)
x = torch.randn(8192, n, device='cuda', generator=g)
H = (x.t() @ x) / 4096
ops.cholinv_upper(H.clone(), 0.01, splits=3)
torch.cuda.synchronize()
ops.cholinv_upper(H.clone(), 0.01, splits=3)
torch.cuda.synchronize()
