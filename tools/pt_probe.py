import os, sys, torch, json
sys.path.insert(0, '/root/repo')
from lightcompress_b200.quant import IntegerQuantizer
dev='cuda:0'
q = IntegerQuantizer(8, True, 'per_token')
for cols in (4096, 2048, 1024):
    rows = (1<<28)//cols
    xs=[torch.randn(rows, cols, device=dev, dtype=torch.bfloat16) for _ in range(2)]
    for _ in range(3): q.fake_quant_act_dynamic(xs[0])
    torch.cuda.synchronize()
    a,b=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(10): q.fake_quant_act_dynamic(xs[i%2])
    b.record(); torch.cuda.synchronize()
    t=a.elapsed_time(b)/10*1e-3
    print(json.dumps({'env':os.environ.get('LCB_GROUP_MAX_NVEC'),'cols':cols,'GBps':round(rows*cols*4/t/1e9)}))
