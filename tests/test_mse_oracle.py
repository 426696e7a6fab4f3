"""The mse-range oracle against outputs of the unmodified reference (tests/golden/mse_golden.pt.gz). CPU only."""
import gzip
import io
import os
import sys

import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, 'golden'))
from make_golden import make_input  # noqa: E402
from oracle import mse_oracle as mo  # noqa: E402


def load_golden():
    with gzip.open(os.path.join(HERE, 'golden', 'mse_golden.pt.gz'), 'rb') as f:
        return torch.load(io.BytesIO(f.read()), weights_only=False)


GOLD = load_golden()


def spec_of(c):
    """(kind, sym, qmin, qmax, round_zp, search kwargs, unit_len) of a golden case."""
    kw = dict(c['kwargs'])
    search = {k: kw[k] for k in ('maxshrink', 'mse_grid') if k in kw}
    if c['cls'] == 'float':
        return 1, True, -448.0, 448.0, True, search
    b = c['bit']
    qmin, qmax = (-2.0 ** (b - 1), 2.0 ** (b - 1) - 1) if c['sym'] else (0.0, 2.0 ** b - 1)
    return 0, c['sym'], qmin, qmax, kw.get('round_zp', True), search


def unit_len(c):
    return c['kwargs'].get('group_size', c['shape'][-1])


@pytest.mark.parametrize('rec', GOLD, ids=[str(r['case']['seed']) for r in GOLD])
def test_mse_oracle_matches_reference(rec):
    c = rec['case']
    x = make_input(c['shape'], getattr(torch, c['dtype']), c['seed'])
    kind, sym, qmin, qmax, rzp, search = spec_of(c)
    t = x.reshape(-1, unit_len(c))
    fake, s, z, mn, mx = mo.fake_quant(t, kind, sym, qmin, qmax, rzp, **search)
    assert torch.equal(mn, rec['min']) and torch.equal(mx, rec['max'])
    assert torch.equal(s, rec['scales'])
    if not sym and kind == 0:
        assert torch.equal(z, rec['zeros'])
    assert torch.equal(fake.reshape(c['shape']), rec['fake'])


def test_mse_range_shrinks_and_aliases():
    # the kept range is a product of grid factors (the aliasing), never wider than min/max
    g = torch.Generator().manual_seed(3)
    t = torch.randn(32, 128, generator=g)
    t[:, 0] = 9.0      # an outlier makes shrinking pay
    mn, mx = mo.get_mse_range(t, 0, False, 0.0, 15.0)
    assert (mx <= t.amax(1, keepdim=True)).all() and (mn >= t.amin(1, keepdim=True)).all()
    assert (mx < t.amax(1, keepdim=True)).any()
