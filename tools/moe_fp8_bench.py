"""FP8 E4M3 per-block real quant of the 198 expert matrices of a DeepSeek-V2-Lite MoE block (BASELINE configs[4]):
multi-tensor entry (real_quant_weights_dynamic) vs one reference-API call per matrix, plus the host-only share
(wall time of the Python call before any synchronisation).

    python tools/moe_fp8_bench.py
"""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lightcompress_b200.quant import FloatQuantizer  # noqa: E402

dev = torch.device('cuda:0')
gen = torch.Generator(device=dev).manual_seed(0)
f8 = FloatQuantizer('e4m3', True, 'per_block', block_size=128, use_qtorch=True)
ws = []
for _ in range(66):
    for r, c in ((1408, 2048), (1408, 2048), (2048, 1408)):
        ws.append((torch.randn(r, c, device=dev, generator=gen, dtype=torch.float32) * 0.02).to(torch.bfloat16))
nbytes = sum(w.numel() for w in ws) * 3 + sum(-(-w.shape[0] // 128) * -(-w.shape[1] // 128) for w in ws) * 4


def ev(fn, rep=5):
    outs = [fn() for _ in range(rep)]
    del outs
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    a.record()
    outs = [fn() for _ in range(rep)]
    b.record()
    host = (time.perf_counter() - t0) / rep
    torch.cuda.synchronize()
    del outs
    return a.elapsed_time(b) * 1e-3 / rep, host


for name, fn in (('multi_launch', lambda: f8.real_quant_weights_dynamic(ws)),
                 ('per_tensor_calls', lambda: [f8.real_quant_weight_dynamic(w) for w in ws]),
                 ('multi_launch_fake', lambda: f8.fake_quant_weights_dynamic(ws))):
    t, host = ev(fn)
    nb = nbytes if 'fake' not in name else sum(w.numel() for w in ws) * 4
    print(json.dumps({'what': name, 'matrices': len(ws), 'ms': round(t * 1e3, 4), 'host_ms': round(host * 1e3, 4),
                      'GBps': round(nb / t / 1e9, 1)}), flush=True)
