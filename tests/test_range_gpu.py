"""GPU parity of get_minmax_range / get_qparams / dequant / calib_algo='learnable' (lcb_minmax, lcb_qparams_from_range,
LCB_OUT_DEQUANT) against outputs of the UNMODIFIED reference quantizer (tests/golden/range_golden.pt.gz). Bit-exact."""
import gzip
import io
import os

import pytest
import torch

from lightcompress_b200 import quant as Q
from tests.golden.make_golden import make_input
from tests.golden.make_range_golden import factors, static_acts

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'
PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden', 'range_golden.pt.gz')
with gzip.open(PATH, 'rb') as f:
    GOLD = torch.load(io.BytesIO(f.read()), weights_only=False)
STATIC = [r for r in GOLD if 'algo' in r['case']]
GOLD = [r for r in GOLD if 'algo' not in r['case']]
IDS = [f"{r['case']['gran']}-b{r['case']['bit']}-{'sym' if r['case']['sym'] else 'asym'}-{r['case']['dtype']}" for r in GOLD]


def _bits(t):
    t = t.detach().cpu().contiguous()
    return t.view({2: torch.int16, 4: torch.int32}[t.element_size()])


def _same(a, b):
    a, b = a.detach().cpu(), b.detach().cpu()
    return a.dtype == b.dtype and a.numel() == b.numel() and torch.equal(_bits(a).reshape(-1), _bits(b).reshape(-1))


@pytest.mark.parametrize('rec', GOLD, ids=IDS)
def test_minmax_and_qparams_match_reference(rec):
    c = rec['case']
    x = make_input(c['shape'], getattr(torch, c['dtype']), c['seed']).to(DEV)
    q = Q.IntegerQuantizer(c['bit'], c['sym'], c['gran'], **c['kwargs'])
    mn, mx = q.get_minmax_range(x)
    assert _same(mn, rec['min']) and _same(mx, rec['max'])
    assert tuple(mn.shape) == tuple(rec['min'].shape)
    s, z, qmax, qmin = q.get_qparams((mn * 0.75, mx * 0.9), x.device)
    assert _same(s, rec['scaled_scales']), 'scales of a modified range differ'
    if c['sym']:
        assert float(z) == 0.0
    else:
        assert _same(z, rec['scaled_zeros'])
    assert float(qmax) == float(q.qmax) and float(qmin) == float(q.qmin)


@pytest.mark.parametrize('rec', [r for r in GOLD if 'codes' in r], ids=[i for i, r in zip(IDS, GOLD) if 'codes' in r])
def test_quant_dequant_and_learnable_match_reference(rec):
    c = rec['case']
    dtype = getattr(torch, c['dtype'])
    x = make_input(c['shape'], dtype, c['seed']).to(DEV)
    q = Q.IntegerQuantizer(c['bit'], c['sym'], c['gran'], **c['kwargs'])
    t, s, z, qmax, qmin = q.get_tensor_qparams(x)
    codes = q.quant(t, s, z, qmax, qmin)
    assert _same(codes, rec['codes'])
    assert _same(q.dequant(codes, s, z), rec['dequant'])
    ql = Q.IntegerQuantizer(c['bit'], c['sym'], c['gran'], calib_algo='learnable', **c['kwargs'])
    for fdt in ('float32', c['dtype']):
        lo, up = factors(rec['min'].numel(), c['seed'], getattr(torch, fdt))
        out = ql.fake_quant_weight_dynamic(x, {'lowbound_factor': lo.to(DEV), 'upbound_factor': up.to(DEV)})
        assert out.dtype == dtype
        assert _same(out, rec[f'learnable_{fdt}']), f'learnable bounds ({fdt} factors) differ'


@pytest.mark.parametrize('rec', STATIC, ids=[f"{r['case']['algo']}-{'sym' if r['case']['sym'] else 'asym'}" for r in STATIC])
def test_static_activation_calibration_matches_reference(rec):
    c = rec['case']
    q = Q.IntegerQuantizer(c['bit'], c['sym'], 'per_tensor', calib_algo=c['algo'])
    acts = [a.to(DEV) for a in static_acts(c)]
    sc, zp, qmin, qmax = q.get_batch_tensors_qparams(acts, alpha=0.01)
    assert len(sc) == len(rec['scales']) == 1
    assert _same(sc[0], rec['scales'][0]), (sc[0], rec['scales'][0])
    if not c['sym']:
        assert _same(zp[0], rec['zeros'][0])
    fake = q.fake_quant_act_static(acts[0], {'scales': sc[0], 'zeros': zp[0], 'qmax': qmax[0], 'qmin': qmin[0]})
    assert _same(fake, rec['fake'])
