"""FloatQuantizer WITHOUT qtorch (csrc/floatq.cu) against tests/golden/floatq_golden.pt.gz: outputs of the UNMODIFIED
reference class on a B200 (tests/golden/make_floatq_golden.py). Bit-exact, and independent of oracle/_ref."""
import pytest
import torch

from lightcompress_b200 import quant as Q
from tests.test_floatq_gpu import _same

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'


def test_product_matches_committed_reference_golden():
    """tests/golden/floatq_golden.pt.gz: outputs of the unmodified reference class on a B200 (make_floatq_golden.py). The
    same bar as against the live reference — bit-exact — without needing oracle/_ref."""
    import gzip
    import io
    import os
    from tests.golden.make_floatq_golden import make_input
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden', 'floatq_golden.pt.gz')
    with gzip.open(path, 'rb') as f:
        recs = torch.load(io.BytesIO(f.read()), weights_only=False)['cases']
    assert len(recs) >= 10
    for rec in recs:
        c = rec['case']
        x = make_input(c['shape'], getattr(torch, c['dtype']), c['seed']).to(DEV)
        pq = Q.FloatQuantizer(c['bit'], True, c['gran'], **c['kwargs'])
        act = c['gran'] in ('per_token', 'per_head')
        got = (pq.fake_quant_act_dynamic if act else pq.fake_quant_weight_dynamic)(x.clone())
        what = f"{c['bit']} {c['gran']} {c['dtype']}"
        _same(got.cpu(), rec['fake'], what + ' fake (golden)')
        xc, sc, _, _, _ = pq.get_tensor_qparams(x.clone())
        _same(xc.cpu(), rec['xc'], what + ' xc (golden)')
        _same(sc.cpu(), rec['scales'], what + ' scales (golden)')
