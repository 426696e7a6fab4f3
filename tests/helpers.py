"""Shared helpers of the parity tests: golden loading, case -> oracle call, case -> product call."""
import gzip
import io
import os

import torch

from oracle import quant_oracle as qo
from tests.golden.make_golden import make_input  # noqa: F401  (inputs are regenerated from the seed)

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden', 'quant_golden.pt.gz')


def load_golden():
    with gzip.open(GOLDEN, 'rb') as f:
        return torch.load(io.BytesIO(f.read()), weights_only=False)


def case_id(rec):
    c = rec['case']
    kw = ','.join(f'{k}={v}' for k, v in c.get('kwargs', {}).items())
    return f"{c['cls']}-{c['bit']}-{'sym' if c.get('sym', True) else 'asym'}-{c['gran']}-{kw}-{c['dtype']}-{'x'.join(map(str, c['shape']))}"


def case_input(c):
    return make_input(c['shape'], getattr(torch, c['dtype']), c['seed'], c.get('input', 'weight'))


def qrange(c):
    if c['cls'] == 'fp':
        return (-448.0, 448.0) if c['bit'] == 'e4m3' else (-57344.0, 57344.0)
    b = c['bit']
    return (-(2.0 ** (b - 1)), 2.0 ** (b - 1) - 1) if c['sym'] else (0.0, 2.0 ** b - 1)


def units(c, shape):
    """(rows, cols, ur, uc) of the quantization-unit grid, from the reference's reshape_tensor (quant.py:612-642)."""
    last = shape[-1]
    numel = 1
    for d in shape:
        numel *= d
    rows = numel // last
    g = c['gran']
    if g == 'per_group' and last >= c['kwargs']['group_size']:
        return rows, last, 1, c['kwargs']['group_size']
    if g == 'per_block':
        return rows, last, c['kwargs']['block_size'], c['kwargs']['block_size']
    if g == 'per_tensor':
        return rows, last, rows, last
    return rows, last, 1, last


def oracle_case(c, x, out_mode, **over):
    rows, cols, ur, uc = units(c, tuple(x.shape))
    qmin, qmax = qrange(c)
    kind = 0 if c['cls'] == 'int' else (1 if c['bit'] == 'e4m3' else 2)
    md = torch.float32 if c['gran'] == 'per_block' else x.dtype
    qp = None
    if c['gran'] == 'per_tensor' and (c['cls'] == 'fp' or not c['sym']):
        qp = torch.float32  # 0-dim range / 0-dim fp32 divisor (qmax - qmin with qmin = tensor(0.0)) -> fp32 qparams
    kw = dict(qp_math=qp, kind=kind, sym=c.get('sym', True), round_zp=c.get('kwargs', {}).get('round_zp', True), qmin=qmin,
              qmax=qmax, out_mode=out_mode, math_dtype=md)
    kw.update(over)
    return qo.fakequant(x, rows, cols, ur, uc, **kw)


def code_mode(c):
    if c['cls'] == 'fp':
        return qo.OUT_FP8
    if c['bit'] == 8:
        return qo.OUT_I8 if c['sym'] else qo.OUT_U8
    return qo.OUT_I32


def bits_equal(a, b):
    """bit-exact comparison that treats tensors as raw storage (NaN == NaN, -0 != +0)."""
    if a.shape != b.shape or a.dtype != b.dtype:
        return False
    if a.dtype in (torch.float32,):
        return torch.equal(a.contiguous().view(torch.int32), b.contiguous().view(torch.int32))
    if a.dtype in (torch.float16, torch.bfloat16):
        return torch.equal(a.contiguous().view(torch.int16), b.contiguous().view(torch.int16))
    return torch.equal(a, b)


def equal_nan_aware(a, b):
    """bit-exact on every non-NaN element, NaN exactly where the other has NaN (payload bits are unspecified)."""
    if a.shape != b.shape or a.dtype != b.dtype:
        return False
    na, nb = torch.isnan(a), torch.isnan(b)
    if not torch.equal(na, nb):
        return False
    z = torch.zeros((), dtype=a.dtype)
    return bits_equal(torch.where(na, z, a), torch.where(nb, z, b))


def make_quantizer(mod, c):
    """Instantiate the product (or reference) quantizer class of module `mod` for a golden case."""
    kw = dict(c.get('kwargs', {}))
    if c['cls'] == 'int':
        return mod.IntegerQuantizer(c['bit'], c['sym'], c['gran'], **kw)
    return mod.FloatQuantizer(c['bit'], True, c['gran'], use_qtorch=True, **kw)
