"""CPU: the C oracle reproduces the reference's own outputs (tests/golden/quant_golden.pt.gz) bit for bit."""
import pytest
import torch

from oracle import quant_oracle as qo
from tests import helpers as H

RECS = H.load_golden()


@pytest.mark.parametrize('rec', RECS, ids=[H.case_id(r) for r in RECS])
def test_oracle_matches_reference(rec):
    c = rec['case']
    x = H.case_input(c)
    if rec.get('raises') == 'ValueError':
        with pytest.raises(ValueError):
            H.oracle_case(c, x, qo.OUT_FAKE)
        return
    assert float(x.double().sum()) == rec['x_sum'], 'input regeneration drifted'
    fake, _, _ = H.oracle_case(c, x, qo.OUT_FAKE)
    assert H.bits_equal(fake, rec['fake']), 'fake-quant values differ'
    _, scales, zeros = H.oracle_case(c, x, qo.OUT_NONE)  # get_tensor_qparams
    assert H.bits_equal(scales.view(rec['qp_scales'].shape), rec['qp_scales']), 'scales differ'
    if zeros is not None:
        assert H.bits_equal(zeros.view(rec['qp_zeros'].shape), rec['qp_zeros']), 'zeros differ'
    if 'codes' in rec:
        codes, s2, z2 = H.oracle_case(c, x, H.code_mode(c))
        ref = rec['codes']
        got = codes if ref.dtype == torch.uint8 and codes.dtype == torch.uint8 else codes.to(torch.int16)
        assert torch.equal(got, ref.view(got.shape)), 'codes differ'
        assert H.bits_equal(s2.view(rec['scales'].shape), rec['scales'])
        if rec['zeros'] is not None:
            assert torch.equal(z2.view(rec['zeros'].shape).to(torch.int16), rec['zeros'].to(torch.int16))
    if 'static_fake' in rec:
        out, _, _ = H.oracle_case(c, x, qo.OUT_FAKE, dynamic=False, math_dtype=torch.float32,
                                  scales=rec['static_scales'].reshape(-1),
                                  zeros=None if c['sym'] else rec['static_zeros'].reshape(-1))
        assert H.bits_equal(out, rec['static_fake']), 'static (fp32 qparams) differs'
        out, _, _ = H.oracle_case(c, x, qo.OUT_FAKE, dynamic=False, scales=rec['qp_scales'].reshape(-1),
                                  zeros=None if c['sym'] else rec['qp_zeros'].reshape(-1))
        assert H.bits_equal(out, rec['static_fake_T']), 'static (T qparams) differs'
