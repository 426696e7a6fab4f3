"""Golden vectors for FloatQuantizer WITHOUT qtorch (per-element scales, quant.py:1005-1027, 1061-1076), from the UNMODIFIED
reference class. That path moves its 0-dim operands with `.cuda()`, so the reference can only produce these on a CUDA
device: run on the GPU box, where the reference copy under oracle/_ref travels with the snapshot:

    gpurun -- 'python tests/golden/make_floatq_golden.py gpurun_out/floatq_golden.pt.gz'
    cp gpurun_out/floatq_golden.pt.gz tests/golden/

Each record: the case, the reference's fake-quant output, and (xc, scales) of get_tensor_qparams. Inputs are regenerated
from the seed by `make_input` below (CPU generator: identical on every machine).
"""
import gzip
import io
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))

CASES = [dict(bit=b, gran=g, kwargs=kw, shape=s, dtype=d, seed=900 + i)
         for i, (b, g, kw, s, d) in enumerate([
             ('e4m3', 'per_channel', {}, (48, 1024), 'bfloat16'), ('e4m3', 'per_token', {}, (2, 24, 768), 'bfloat16'),
             ('e4m3', 'per_group', {'group_size': 128}, (32, 512), 'float16'), ('e5m2', 'per_channel', {}, (48, 1024), 'bfloat16'),
             ('e5m2', 'per_token', {}, (2, 24, 768), 'float16'), ('e2m1', 'per_group', {'group_size': 128}, (32, 512), 'bfloat16'),
             ('e3m2', 'per_channel', {}, (16, 4100), 'bfloat16'), ('e4m3', 'per_channel', {}, (24, 640), 'float32'),
             ('e4m3', 'per_group', {'group_size': 32}, (16, 96), 'bfloat16'), ('e5m2', 'per_group', {'group_size': 128}, (32, 512), 'float32')])]


def make_input(shape, dtype, seed):
    g = torch.Generator().manual_seed(seed)
    w = torch.randn(*shape, generator=g) * torch.exp(torch.randn(*shape[:-1], 1, generator=g) - 3.0)
    flat = w.view(-1, shape[-1])
    flat[1, 3] *= 50.0
    flat[2, :] = 0.0          # an all-zero unit: NaN in the reference (log2(0) + inf)
    flat[3, 7] = 0.0
    return w.to(dtype)


def main(out):
    sys.path.insert(0, ROOT)
    from oracle import ref_harness as rh
    ref = rh.activate('cuda')['quant']
    recs = []
    for c in CASES:
        x = make_input(c['shape'], getattr(torch, c['dtype']), c['seed']).cuda()
        q = ref.FloatQuantizer(c['bit'], True, c['gran'], **c['kwargs'])
        act = c['gran'] in ('per_token', 'per_head')
        y = (q.fake_quant_act_dynamic if act else q.fake_quant_weight_dynamic)(x.clone())
        xc, sc, z, qmax, qmin = q.get_tensor_qparams(x.clone())
        assert qmax is None and qmin is None and int(z) == 0
        recs.append({'case': c, 'fake': y.cpu(), 'xc': xc.cpu(), 'scales': sc.cpu()})
    buf = io.BytesIO()
    torch.save({'cases': recs, 'torch': torch.__version__, 'device': torch.cuda.get_device_name(0)}, buf)
    with gzip.open(out, 'wb') as f:
        f.write(buf.getvalue())
    print('wrote', out, len(recs), 'cases')


if __name__ == '__main__':
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(HERE, 'floatq_golden.pt.gz'))
