"""GPU parity of lcb_gptq_weight_transform against the oracle restatement of GPTQ.weight_transform (gptq.py:199-244).

The in-block column loop is bit-identical to the reference given identical inputs (checked on the first block, whose
inputs are identical by construction). Later blocks see W after tensor-core lazy updates (bf16 hi/lo split, ~1e-5
relative), so they are compared in the north star's tolerance: 1e-3 relative (Frobenius) on the updated weights."""
import pytest
import torch

from lightcompress_b200 import ops
from oracle import gptq_oracle as go

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'


def _problem(R, C, T, seed):
    g = torch.Generator().manual_seed(seed)
    W = torch.randn(R, C, generator=g) * 0.02
    gain = torch.exp(0.5 * torch.randn(C, generator=g))
    X = (torch.randn(T, C, generator=g) * gain).to(torch.bfloat16)
    H = torch.zeros(C, C)
    go.add_batch(H, 0, X.unsqueeze(0))
    Wf, Hinv = go.process_hessian(H, W, 0.01)
    return Wf, Hinv.contiguous()


def _rel(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm())


@pytest.mark.parametrize('R,C,bs,gs,sym,bit', [
    (64, 256, 128, 128, False, 4), (96, 384, 128, 128, True, 4), (40, 512, 128, 64, False, 3),
    (64, 512, 128, 256, False, 4), (33, 256, 64, 64, False, 4), (64, 1152, 128, 128, False, 4),
])
def test_weight_transform_matches_oracle(R, C, bs, gs, sym, bit):
    W, Hinv = _problem(R, C, 2 * C, seed=R + C)
    qmin, qmax = (-(2.0 ** (bit - 1)), 2.0 ** (bit - 1) - 1) if sym else (0.0, 2.0 ** bit - 1)
    spec = dict(kind=0, sym=sym, qmin=qmin, qmax=qmax)
    Wo = W.clone()
    tmp_o, loss_o, sc_o, zp_o = go.weight_transform(Wo, Hinv, bs, gs, spec)
    Wg = W.clone().to(DEV)
    tmp, losses, sc, zp = ops.gptq_weight_transform(Wg, Hinv.to(DEV), bs, gs, qmin, qmax, sym)
    tmp, losses, sc = tmp.cpu(), losses.cpu(), sc.cpu()
    # first block: identical inputs -> identical bits
    assert torch.equal(tmp[:, :bs], tmp_o[:, :bs])
    assert torch.equal(losses[:, :bs], loss_o[:, :bs])
    ng0 = max(1, bs // gs)
    assert torch.equal(sc[:ng0], sc_o[:ng0])
    if not sym:
        assert torch.equal(zp.cpu()[:ng0], zp_o[:ng0])
    # whole matrix: 1e-3 relative
    assert _rel(tmp, tmp_o) < 1e-3
    assert _rel(sc, sc_o) < 1e-3
    assert abs(float(losses.sum()) - float(loss_o.sum())) <= 2e-2 * float(loss_o.sum())
    # the quantization grid decisions agree almost everywhere
    frac = float(((tmp - tmp_o).abs() > 1e-4 * tmp_o.abs().max()).float().mean())
    assert frac < 0.02, frac


def test_weight_transform_static_per_channel():
    R, C, bs = 48, 384, 128
    W, Hinv = _problem(R, C, 2 * C, seed=5)
    spec = dict(kind=0, sym=True, qmin=-8.0, qmax=7.0)
    tmp_o, loss_o, sc_o, _ = go.weight_transform(W.clone(), Hinv, bs, 0, spec)
    tmp, losses, _, _ = ops.gptq_weight_transform(W.clone().to(DEV), Hinv.to(DEV), bs, 0, -8.0, 7.0, True,
                                                  static_scales=sc_o.to(DEV))
    assert torch.equal(tmp.cpu()[:, :bs], tmp_o[:, :bs])
    assert _rel(tmp.cpu(), tmp_o) < 1e-3


def test_weight_transform_outside_exactness_window_takes_literal_pass():
    # operands below 1e-30 leave the window of the speculative (Markstein) pass: the warp re-runs the literal pass, and the
    # first block must still be bit-identical to the reference arithmetic
    R, C, bs = 64, 256, 128
    W, Hinv = _problem(R, C, 2 * C, seed=9)
    W[3, 5] = 1e-33
    W[10, 40] = -3e-35
    W[20, :128] = 0.0          # an all-zero group (scale clamps to 1e-5 / 15)
    spec = dict(kind=0, sym=False, qmin=0.0, qmax=15.0)
    tmp_o, loss_o, sc_o, zp_o = go.weight_transform(W.clone(), Hinv, bs, 128, spec)
    tmp, losses, sc, zp = ops.gptq_weight_transform(W.clone().to(DEV), Hinv.to(DEV), bs, 128, 0.0, 15.0, False)
    assert torch.equal(tmp.cpu()[:, :bs], tmp_o[:, :bs])
    assert torch.equal(losses.cpu()[:, :bs], loss_o[:, :bs])
    assert torch.equal(sc.cpu()[:1], sc_o[:1]) and torch.equal(zp.cpu()[:1], zp_o[:1])
    assert _rel(tmp.cpu(), tmp_o) < 1e-3
