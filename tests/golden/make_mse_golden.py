"""Golden vectors for calib_algo='mse' (get_mse_range, quant.py:145-203) from the UNMODIFIED reference quantizer on CPU.

Run in the build container only (needs /root/reference):
    PYTHONPATH=/root/reference python tests/golden/make_mse_golden.py
Writes tests/golden/mse_golden.pt.gz: per case the searched (min, max) of every unit, the qparams and the fake-quant
output. Inputs are regenerated from the seed by tests/golden/make_golden.make_input.
"""
import gzip
import io
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_golden import fp8_cast_quantize, make_input  # noqa: E402

OUT = os.path.join(HERE, 'mse_golden.pt.gz')

CASES = [
    dict(cls='int', bit=4, sym=False, gran='per_group', kwargs={'group_size': 128}, shape=(16, 512), dtype='bfloat16', seed=401),
    dict(cls='int', bit=4, sym=True, gran='per_group', kwargs={'group_size': 64}, shape=(8, 256), dtype='float16', seed=402),
    dict(cls='int', bit=4, sym=True, gran='per_channel', kwargs={}, shape=(24, 768), dtype='bfloat16', seed=403),
    dict(cls='int', bit=8, sym=False, gran='per_channel', kwargs={'mse_b_num': 2}, shape=(12, 1536), dtype='float32', seed=404),
    dict(cls='int', bit=3, sym=False, gran='per_group', kwargs={'group_size': 32, 'maxshrink': 0.5, 'mse_grid': 40},
         shape=(8, 128), dtype='bfloat16', seed=405),
    dict(cls='int', bit=4, sym=False, gran='per_token', kwargs={'round_zp': False}, shape=(12, 320), dtype='bfloat16', seed=406),
    dict(cls='float', bit='e4m3', sym=True, gran='per_channel', kwargs={'use_qtorch': True}, shape=(16, 256), dtype='bfloat16',
         seed=407),
]


def run(c):
    from llmc.compression.quantization import quant as refq
    dtype = getattr(torch, c['dtype'])
    x = make_input(c['shape'], dtype, c['seed'])
    cls = refq.IntegerQuantizer if c['cls'] == 'int' else refq.FloatQuantizer
    if c['cls'] == 'float':   # qtorch is not installable offline: the same substitution as make_golden.py (fp8 cast)
        refq.float_quantize = fp8_cast_quantize
    q = cls(c['bit'], c['sym'], c['gran'], calib_algo='mse', **c['kwargs'])
    rec = dict(case=c)
    with torch.no_grad():
        t = q.reshape_tensor(x.clone())
        mn, mx = q.get_tensor_range(t)
        rec['min'], rec['max'] = mn.clone(), mx.clone()
        _, s, z, _, _ = q.get_tensor_qparams(x.clone())
        rec['scales'], rec['zeros'] = s, torch.as_tensor(z)
        rec['fake'] = q.fake_quant_weight_dynamic(x.clone())
    return rec


def main():
    recs = [run(c) for c in CASES]
    buf = io.BytesIO()
    torch.save(recs, buf)
    with gzip.open(OUT, 'wb', compresslevel=9) as f:
        f.write(buf.getvalue())
    for r in recs:
        print(r['case']['seed'], tuple(r['min'].shape), r['min'].dtype, r['fake'].dtype, tuple(r['fake'].shape))
    print('wrote', OUT, os.path.getsize(OUT), 'bytes')


if __name__ == '__main__':
    main()
