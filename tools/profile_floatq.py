"""A few FloatQuantizer (no qtorch) fake-quant calls at Llama-3-8B sizes, for ncu:
    ncu --set full --clock-control none --import-source on -k regex:floatq -c 4 python tools/profile_floatq.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lightcompress_b200.quant import FloatQuantizer  # noqa: E402

dev = torch.device('cuda:0')
g = torch.Generator(device=dev).manual_seed(0)
a = (torch.randn(8, 2048, 4096, device=dev, generator=g) * 0.05).to(torch.bfloat16)
w = (torch.randn(4096, 14336, device=dev, generator=g) * 0.02).to(torch.bfloat16)
for bit in ('e4m3', 'e5m2'):
    FloatQuantizer(bit, True, 'per_token').fake_quant_act_dynamic(a)
    FloatQuantizer(bit, True, 'per_channel').fake_quant_weight_dynamic(w)
torch.cuda.synchronize()
