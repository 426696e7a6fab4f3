"""Golden vectors for the range / qparams / dequant entry points, from the UNMODIFIED reference quantizer on CPU.

Run in the build container only (needs /root/reference):
    PYTHONPATH=/root/reference python tests/golden/make_range_golden.py
Writes tests/golden/range_golden.pt.gz: get_minmax_range (quant.py:132-143), get_qparams on scaled ranges (:545-559),
fake_quant_weight_dynamic with calib_algo='learnable' and bound factors (:205-219, base_bq.py:46-65), quant + dequant
(:699-712). Inputs are regenerated from the seed by tests/golden/make_golden.make_input.
"""
import gzip
import io
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_golden import make_input  # noqa: E402

OUT = os.path.join(HERE, 'range_golden.pt.gz')

CASES = [
    dict(bit=4, sym=False, gran='per_group', kwargs={'group_size': 128}, shape=(16, 512), dtype='bfloat16', seed=301),
    dict(bit=4, sym=True, gran='per_group', kwargs={'group_size': 64}, shape=(8, 256), dtype='float16', seed=302),
    dict(bit=8, sym=False, gran='per_channel', kwargs={}, shape=(24, 768), dtype='bfloat16', seed=303),
    dict(bit=8, sym=True, gran='per_tensor', kwargs={}, shape=(32, 160), dtype='float32', seed=304),
    dict(bit=8, sym=False, gran='per_tensor', kwargs={}, shape=(32, 160), dtype='bfloat16', seed=305),
    dict(bit=8, sym=True, gran='per_block', kwargs={'block_size': 128}, shape=(200, 384), dtype='bfloat16', seed=306),
    dict(bit=4, sym=False, gran='per_head', kwargs={'head_num': 8}, shape=(2, 8, 96), dtype='bfloat16', seed=307),
]


def factors(n_units, seed, dtype):
    g = torch.Generator().manual_seed(seed)
    lo = (torch.rand(n_units, 1, generator=g) * 4 + 1).to(dtype)
    up = (torch.rand(n_units, 1, generator=g) * 4 + 1).to(dtype)
    return lo, up


def run(c):
    from llmc.compression.quantization import quant as refq
    dtype = getattr(torch, c['dtype'])
    x = make_input(c['shape'], dtype, c['seed'])
    q = refq.IntegerQuantizer(c['bit'], c['sym'], c['gran'], **c['kwargs'])
    rec = dict(case=c)
    with torch.no_grad():
        t = q.reshape_tensor(x.clone())
        mn, mx = q.get_minmax_range(t)
        rec['min'], rec['max'] = mn, mx
        s, z, qmax, qmin = q.get_qparams((mn * 0.75, mx * 0.9), torch.device('cpu'))
        rec['scaled_scales'], rec['scaled_zeros'] = s, z
        if c['gran'] in ('per_group', 'per_channel'):
            codes = q.quant(t, *q.get_tensor_qparams(x.clone())[1:])
            _, s0, z0, _, _ = q.get_tensor_qparams(x.clone())
            rec['codes'], rec['dequant'] = codes, q.dequant(codes, s0, z0)
            ql = refq.IntegerQuantizer(c['bit'], c['sym'], c['gran'], calib_algo='learnable', **c['kwargs'])
            for fdt in ('float32', c['dtype']):
                lo, up = factors(mn.numel(), c['seed'], getattr(torch, fdt))
                rec[f'learnable_{fdt}'] = ql.fake_quant_weight_dynamic(x.clone(), {'lowbound_factor': lo, 'upbound_factor': up})
    return rec


STATIC = [dict(bit=8, sym=True, algo='static_minmax', dtype='bfloat16', seed=321),
          dict(bit=8, sym=False, algo='static_minmax', dtype='float16', seed=322),
          dict(bit=8, sym=False, algo='static_moving_minmax', dtype='bfloat16', seed=323)]


def static_acts(c):
    return [make_input((1, 48, 256), getattr(torch, c['dtype']), c['seed'] + i, 'act') for i in range(3)]


def run_static(c):
    from llmc.compression.quantization import quant as refq
    q = refq.IntegerQuantizer(c['bit'], c['sym'], 'per_tensor', calib_algo=c['algo'])
    with torch.no_grad():
        sc, zp, qmin, qmax = q.get_batch_tensors_qparams(static_acts(c), alpha=0.01)
        x = static_acts(c)[0]
        fake = q.fake_quant_act_static(x.clone(), {'scales': sc[0], 'zeros': zp[0], 'qmax': qmax[0], 'qmin': qmin[0]})
    return dict(case=c, scales=sc, zeros=zp, fake=fake)


def main():
    sys.path.insert(0, '/root/reference')
    recs = [run(c) for c in CASES] + [run_static(c) for c in STATIC]
    buf = io.BytesIO()
    torch.save(recs, buf)
    with gzip.open(OUT, 'wb') as f:
        f.write(buf.getvalue())
    print('wrote', OUT, os.path.getsize(OUT), 'bytes')


if __name__ == '__main__':
    main()
