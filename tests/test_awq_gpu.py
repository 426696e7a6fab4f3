"""GPU parity of the AWQ scale-search kernels against the oracle restatement of awq.py:202-248.
Bar (north star): the same selected candidate index; losses are also compared (GEMM summation order differs, so the
16-bit rounding of `out` may flip on isolated elements: 1e-3 relative on the loss)."""
import pytest
import torch

from lightcompress_b200 import ops
from lightcompress_b200 import quant as Q
from oracle import awq_oracle as ao

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'


def _case(T, C, R, dtype, seed):
    g = torch.Generator().manual_seed(seed)
    gain = torch.exp(torch.randn(C, generator=g))
    x = (torch.randn(T, C, generator=g) * gain).to(dtype)
    W = (torch.randn(R, C, generator=g) * 0.02).to(dtype)
    return x, W


@pytest.mark.parametrize('dtype', [torch.bfloat16, torch.float16])
def test_scale_cols_bit_exact(dtype):
    x, _ = _case(300, 520, 8, dtype, 0)
    s = (torch.rand(520) + 0.5).to(dtype)
    for div in (False, True):
        ref = (x / s.view(1, -1)) if div else (x * s.view(1, -1))
        got = ops.scale_cols(x.to(DEV), s.to(DEV), div=div).cpu()
        assert torch.equal(got.view(torch.int16), ref.view(torch.int16))


@pytest.mark.parametrize('T,C,R', [(256, 512, 384), (200, 320, 136), (1024, 1024, 1024)])
@pytest.mark.parametrize('dtype', [torch.bfloat16, torch.float16])
def test_linear_mse_matches_torch(T, C, R, dtype):
    x, W = _case(T, C, R, dtype, 1)
    org = torch.nn.functional.linear(x, W)
    W2 = (W.float() * (1 + 0.05 * torch.randn(R, C))).to(dtype)
    out = torch.nn.functional.linear(x, W2)
    ref = float((org - out).float().pow(2).sum())
    got = float(ops.linear_mse(x.to(DEV), W2.to(DEV), org.to(DEV)).item())
    assert abs(got - ref) <= 2e-3 * ref


@pytest.mark.parametrize('dtype,bit,sym,gs', [(torch.bfloat16, 4, False, 128), (torch.float16, 4, False, 128),
                                              (torch.bfloat16, 3, True, 64), (torch.bfloat16, 8, True, 0)])
def test_awq_candidates_same_argmin(dtype, bit, sym, gs):
    T, C, R = 512, 512, 384
    x, W = _case(T, C, R, dtype, 7)
    w_max = ao.get_weight_scale([W], gs)
    cands = torch.stack([ao.get_scales(x, w_max, n / 20) for n in range(20)])
    qmin, qmax = (-(2.0 ** (bit - 1)), 2.0 ** (bit - 1) - 1) if sym else (0.0, 2.0 ** bit - 1)
    losses, best_i, org = ao.search_losses(W, x, cands, gs, dict(kind=0, sym=sym, qmin=qmin, qmax=qmax))
    wq = Q.IntegerQuantizer(bit, sym, 'per_group' if gs > 0 else 'per_channel', **({'group_size': gs} if gs > 0 else {}))
    got = ops.awq_candidates(W.to(DEV), x.to(DEV), org.to(DEV), cands.to(DEV), wq).cpu()
    ref = torch.tensor(losses, dtype=torch.float64)
    assert float(((got - ref).abs() / ref).max()) < 2e-3
    assert int(got.argmin()) == best_i


def test_awq_search_unit_sharding_equals_single_rank(monkeypatch):
    """(candidate, row tile) work units (SURVEY 8e, 8 ranks -> 2 row tiles x 20 candidates): the partial sums of the fake
    ranks add up to the single-rank losses, and the selected candidate is the same."""
    from lightcompress_b200.algo import awq as lcawq
    from lightcompress_b200.algo import dist as lcdist
    x, W = _case(512, 512, 384, torch.bfloat16, 9)
    x, W = x.to(DEV), W.to(DEV)
    wq = Q.IntegerQuantizer(4, False, 'per_group', group_size=128)
    sc1, best1, losses1 = lcawq.awq_search_scale(x, W, wq, 'v2')
    world = 8
    parts = []

    def fake_reduce(t):
        parts.append(t.clone())
        return t
    monkeypatch.setattr(lcdist, 'reduce_sum', fake_reduce)
    for r in range(world):
        monkeypatch.setattr(lcdist, 'world', lambda r=r: (r, world))
        nt, units = lcdist.shard_units(20)
        assert nt == 2 and len(units) == 5
        try:
            lcawq.awq_search_scale(x, W, wq, 'v2')
        except RuntimeError:
            pass        # a single fake rank sees inf for the candidates it does not own
    total = torch.stack(parts).sum(0) / float(512 * 384)
    assert float(((total.cpu() - losses1) .abs() / losses1).max()) < 1e-6
    assert int(total.argmin()) == best1


@pytest.mark.parametrize('dtype', [torch.bfloat16, torch.float16])
@pytest.mark.parametrize('T,C,R,bias', [(300, 512, 384, False), (1024, 1024, 520, True), (77, 256, 8, True),
                                        (4096, 4096, 1024, False)])
def test_linear_matches_torch(dtype, T, C, R, bias):
    """lcb_linear = F.linear on tcgen05: fp32 accumulation, ONE rounding to the tensor dtype (also after the bias)."""
    x, W = _case(T, C, R, dtype, 3)
    b = (torch.randn(R, generator=torch.Generator().manual_seed(4)) * 0.1).to(dtype) if bias else None
    ref32 = x.double() @ W.double().t() + (b.double() if bias else 0.0)
    got = ops.linear(x.to(DEV), W.to(DEV), None if b is None else b.to(DEV)).cpu()
    assert got.dtype == dtype and got.shape == (T, R)
    ulp = 2.0 ** -8 if dtype == torch.bfloat16 else 2.0 ** -11
    err = (got.double() - ref32).abs()
    # one rounding of an fp32-accumulated sum: half an ulp of the result + the fp32 accumulation error
    assert bool((err <= ulp * ref32.abs() * 1.01 + 1e-4 * ref32.abs().max()).all())
    lib = torch.nn.functional.linear(x.to(DEV), W.to(DEV), None if b is None else b.to(DEV)).cpu()
    assert float((got != lib).float().mean()) < 0.02      # the library rounds the same sums: isolated one-ulp flips only
    x3 = x.to(DEV).reshape(1, T, C)
    assert ops.linear(x3, W.to(DEV)).shape == (1, T, R)


def test_own_linears_swaps_and_restores():
    from lightcompress_b200.algo import awq as lcawq
    mlp = torch.nn.Sequential(torch.nn.Linear(256, 512, bias=True), torch.nn.SiLU(), torch.nn.Linear(512, 256, bias=False))
    mlp = mlp.to(torch.bfloat16).to(DEV)
    x = torch.randn(64, 256, device=DEV).to(torch.bfloat16)
    ref = mlp(x)
    with lcawq.own_linears(mlp) as o:
        assert len(o.mods) == 2 and all('forward' in m.__dict__ for m in o.mods)
        got = mlp(x)
    assert all('forward' not in m.__dict__ for m in mlp.modules())
    assert float((got.float() - ref.float()).abs().max()) <= 2e-2 * float(ref.float().abs().max())


@pytest.mark.parametrize('dtype', [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize('shape', [(64, 512, 1024), (7, 33), (1000, 4099)])
def test_sqdiff_mean_is_the_reference_loss_in_one_pass(dtype, shape):
    """ops.sqdiff_mean == (a - b).float().pow(2).mean() (calculate_loss, awq.py:134-145): same element roundings (difference
    in the tensor dtype, fp32 squares); only the summation order differs, bounded here against an fp64 sum of the same terms."""
    g = torch.Generator(device=DEV).manual_seed(shape[0])
    a = torch.randn(*shape, device=DEV, generator=g).to(dtype)
    b = (a.float() + 0.05 * torch.randn(*shape, device=DEV, generator=g)).to(dtype)
    got = float(ops.sqdiff_mean(a, b))
    terms = (a - b).float().pow(2)
    exact = float(terms.double().sum() / terms.numel())
    ref = float(terms.mean())
    assert abs(got - exact) <= 2e-6 * exact
    assert abs(got - ref) <= 1e-5 * exact
    assert float(ops.sqdiff_mean(a, a)) == 0.0
