"""Pin the AutoClip oracle (oracle/autoclip_oracle.py + the clip mode of oracle/lc_oracle.c) to outputs of the UNMODIFIED
reference AutoClipper (tests/golden/autoclip_golden.pt.gz, made by tests/golden/make_autoclip_golden.py). Bit-exact."""
import gzip
import io
import os

import pytest
import torch

from oracle import autoclip_oracle as aco
from oracle import quant_oracle as qo
from tests.golden.make_autoclip_golden import inputs

PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden', 'autoclip_golden.pt.gz')
with gzip.open(PATH, 'rb') as f:
    GOLD = torch.load(io.BytesIO(f.read()), weights_only=False)


def spec_of(c):
    qmin, qmax = (-(2.0 ** (c['bit'] - 1)), 2.0 ** (c['bit'] - 1) - 1) if c['sym'] else (0.0, 2.0 ** c['bit'] - 1)
    return dict(kind=0, sym=c['sym'], qmin=qmin, qmax=qmax)


IDS = [f"oc{r['case']['oc']}-C{r['case']['C']}-{r['case']['dtype']}-sym{int(r['case']['clip_sym'])}" for r in GOLD]


@pytest.mark.parametrize('rec', GOLD, ids=IDS)
def test_clamped_fake_quant_matches_reference(rec):
    c = rec['case']
    W, _ = inputs(c)
    for i_s, ref in rec['q_w'].items():
        ratio = 1 - i_s / 20
        got, _, _ = qo.fakequant(W, c['oc'], c['C'], 1, c['gs'], clip=(-ratio if c['clip_sym'] else ratio), **spec_of(c))
        assert torch.equal(got.view(torch.int16), ref.view(torch.int16)), f'i_s={i_s}'


@pytest.mark.parametrize('rec', GOLD, ids=IDS)
def test_autoclip_oracle_matches_reference(rec):
    c = rec['case']
    W, x = inputs(c)
    bmax, bmin, _ = aco.auto_clip_layer(W, x, spec_of(c), c['gs'], c['clip_sym'], n_sample_token=c['tok'])
    assert torch.equal(bmax.view(torch.int16), rec['best_max'].view(torch.int16))
    assert torch.equal(bmin.view(torch.int16), rec['best_min'].view(torch.int16))
