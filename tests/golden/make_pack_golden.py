"""Golden vectors of the real-quant packers from the UNMODIFIED reference classmethods on CPU.

Run in the build container only (needs /root/reference):
    PYTHONPATH=/root/reference python tests/golden/make_pack_golden.py
Writes tests/golden/pack_golden.pt.gz. Environment shims, none touching reference code: `Tensor.cuda` is the identity and
`Tensor.to(..., device='cuda')` drops the device (the classmethods hard-code CUDA: module_utils.py:953, 1121, 1141, 1125).
"""
import gzip
import io
import os
import sys

import torch
import torch.nn as nn

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_golden import make_input  # noqa: E402

OUT = os.path.join(HERE, 'pack_golden.pt.gz')
VLLM = [dict(bit=4, shape=(24, 512), gs=128, dtype='bfloat16', seed=401), dict(bit=8, shape=(16, 264), gs=0, dtype='float16', seed=402),
        dict(bit=4, shape=(8, 260), gs=0, dtype='bfloat16', seed=403)]
AWQ = [dict(shape=(64, 256), gs=128, dtype='float16', seed=411), dict(shape=(96, 384), gs=64, dtype='bfloat16', seed=412)]


def main():
    sys.path.insert(0, '/root/reference')
    torch.Tensor.cuda = lambda self, *a, **k: self
    _to = torch.Tensor.to

    def to_no_cuda(self, *a, **k):
        if k.get('device') == 'cuda':
            k.pop('device')
        return _to(self, *a, **k)
    torch.Tensor.to = to_no_cuda
    _zeros = torch.zeros
    torch.zeros = lambda *a, **k: _zeros(*a, **{kk: vv for kk, vv in k.items() if not (kk == 'device' and str(vv).startswith('cuda'))})
    from llmc.compression.quantization import quant as refq
    from llmc.compression.quantization.module_utils import AutoawqRealQuantLinear, VllmRealQuantLinear
    recs = {'vllm': [], 'awq': []}
    with torch.no_grad():
        for c in VLLM:
            x = make_input(c['shape'], getattr(torch, c['dtype']), c['seed'])
            q = (refq.IntegerQuantizer(c['bit'], True, 'per_group', group_size=c['gs']) if c['gs']
                 else refq.IntegerQuantizer(c['bit'], True, 'per_channel'))
            codes, scales, _ = q.real_quant_weight_dynamic(x.clone())
            cfg = {'weight': {'bit': c['bit']}}
            packed, s16 = VllmRealQuantLinear.pack(codes, scales, cfg)
            recs['vllm'].append(dict(case=c, codes=codes, scales=scales, packed=packed, scales16=s16))
        for c in AWQ:
            x = make_input(c['shape'], getattr(torch, c['dtype']), c['seed'])
            q = refq.IntegerQuantizer(4, False, 'per_group', group_size=c['gs'])
            _, scales, zeros = q.real_quant_weight_dynamic(x.clone())
            m = nn.Linear(c['shape'][1], c['shape'][0], bias=False)
            m.weight.data = x.clone()
            cfg = {'weight': {'bit': 4, 'group_size': c['gs']}}
            qw, s16, qz = AutoawqRealQuantLinear.gemm_pack(m, m.weight.data, scales, zeros, cfg)
            recs['awq'].append(dict(case=c, scales=scales, zeros=zeros, qweight=qw, scales16=s16, qzeros=qz))
    buf = io.BytesIO()
    torch.save(recs, buf)
    with gzip.open(OUT, 'wb') as f:
        f.write(buf.getvalue())
    print('wrote', OUT, os.path.getsize(OUT), 'bytes')


if __name__ == '__main__':
    main()
