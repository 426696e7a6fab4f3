"""FloatQuantizer WITHOUT qtorch (per-element scales: quant.py:1005-1027, 1061-1076; csrc/floatq.cu) against the UNMODIFIED
reference class run on the same B200 (oracle/_ref). The reference path only exists on CUDA (`get_float_qparams` moves its
0-dim operands with `.cuda()`), so the GPU box is where it can be pinned. Bar: bit-exact (NaNs in the same places — an
all-zero unit is NaN in the reference: log2(0) + inf)."""
import pytest
import torch

from lightcompress_b200 import quant as Q
from oracle import ref_harness as rh

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not rh.available(), reason='oracle/_ref missing (python oracle/build_ref.py)')]
DEV = 'cuda:0'


@pytest.fixture(scope='module')
def ref():
    return rh.activate('cuda')['quant']


def _same(a, b, what):
    assert a.dtype == b.dtype and a.shape == b.shape, (what, a.dtype, b.dtype, a.shape, b.shape)
    an, bn = torch.isnan(a), torch.isnan(b)
    assert torch.equal(an, bn), f'{what}: NaN pattern differs'
    bits = {2: torch.int16, 4: torch.int32}[a.element_size()]
    av, bv = a.contiguous().view(bits), b.contiguous().view(bits)
    ne = (av != bv) & ~an
    n = int(ne.sum())
    if n:
        i = int(ne.flatten().nonzero()[0])
        raise AssertionError(f'{what}: {n} of {a.numel()} values differ, first at {i}: '
                             f'{a.flatten()[i].item()!r} vs {b.flatten()[i].item()!r}')


def _weights(shape, dtype, seed):
    g = torch.Generator().manual_seed(seed)
    w = torch.randn(*shape, generator=g) * torch.exp(torch.randn(*shape[:-1], 1, generator=g) - 3.0)
    w.view(-1, shape[-1])[1, 3] *= 50.0
    w.view(-1, shape[-1])[2, :] = 0.0         # an all-zero unit (per_channel / per_token): NaN in the reference
    w.view(-1, shape[-1])[3, 7] = 0.0
    return w.to(dtype).to(DEV)


CASES = [('per_channel', {}, (96, 1024)), ('per_group', {'group_size': 128}, (64, 512)), ('per_token', {}, (2, 40, 768)),
         ('per_group', {'group_size': 32}, (16, 96)), ('per_channel', {}, (8, 4100)), ('per_head', {'head_num': 8}, (4, 8, 64))]


@pytest.mark.parametrize('dtype', [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize('bit', ['e4m3', 'e5m2', 'e2m1', 'e3m2'])
@pytest.mark.parametrize('gran,kw,shape', CASES)
def test_fake_quant_dynamic_matches_reference(ref, bit, dtype, gran, kw, shape):
    w = _weights(shape, dtype, len(shape) * 100 + shape[-1])
    rq = ref.FloatQuantizer(bit, True, gran, **kw)
    pq = Q.FloatQuantizer(bit, True, gran, **kw)
    is_act = gran in ('per_token', 'per_head')
    want = (rq.fake_quant_act_dynamic if is_act else rq.fake_quant_weight_dynamic)(w.clone())
    got = (pq.fake_quant_act_dynamic if is_act else pq.fake_quant_weight_dynamic)(w.clone())
    _same(got, want, f'{bit} {gran} {dtype} fake')


@pytest.mark.parametrize('dtype', [torch.bfloat16, torch.float16])
@pytest.mark.parametrize('bit', ['e4m3', 'e5m2', 'e2m1'])
def test_qparams_static_and_ic_match_reference(ref, bit, dtype):
    w = _weights((128, 512), dtype, 5)
    for gran, kw in (('per_channel', {}), ('per_group', {'group_size': 128})):
        rq = ref.FloatQuantizer(bit, True, gran, **kw)
        pq = Q.FloatQuantizer(bit, True, gran, **kw)
        t_r, s_r, z_r, qmax_r, qmin_r = rq.get_tensor_qparams(w.clone())
        t_p, s_p, z_p, qmax_p, qmin_p = pq.get_tensor_qparams(w.clone())
        assert qmax_r is None and qmax_p is None and qmin_r is None and qmin_p is None
        assert int(z_r) == 0 and int(z_p) == 0
        _same(t_p, t_r, f'{bit} {gran} clamped tensor')
        _same(s_p, s_r, f'{bit} {gran} scales')
        # static: the scales of get_tensor_qparams fed back (finite rows only: NaN scales stay NaN in both)
        args_r = {'scales': s_r.clone(), 'zeros': z_r, 'qmax': None, 'qmin': None}
        args_p = {'scales': s_p.clone(), 'zeros': z_p, 'qmax': None, 'qmin': None}
        _same(pq.fake_quant_weight_static(w.clone(), args_p), rq.fake_quant_weight_static(w.clone(), args_r),
              f'{bit} {gran} static')
        _same(pq.quant_dequant(t_p, s_p.clone(), z_p, None, None), rq.quant_dequant(t_r, s_r.clone(), z_r, None, None),
              f'{bit} {gran} quant_dequant')
    rq = ref.FloatQuantizer(bit, True, 'per_channel')
    pq = Q.FloatQuantizer(bit, True, 'per_channel')
    _same(pq.fake_quant_weight_dynamic(w.clone(), {'dim': 'ic'}), rq.fake_quant_weight_dynamic(w.clone(), {'dim': 'ic'}),
          f'{bit} ic')


def test_llama_sized_rows_and_error_behaviour(ref):
    """A [2048, 4096] per-token activation and a [4096, 14336]-wide row at the shipped recipe's formats; what the
    reference cannot run raises the same exception type."""
    for bit in ('e4m3', 'e5m2'):
        rq = ref.FloatQuantizer(bit, True, 'per_token')
        pq = Q.FloatQuantizer(bit, True, 'per_token')
        g = torch.Generator(device=DEV).manual_seed(1)
        a = (torch.randn(1, 2048, 4096, device=DEV, generator=g) * 3).to(torch.bfloat16)
        _same(pq.fake_quant_act_dynamic(a), rq.fake_quant_act_dynamic(a), f'{bit} act')
        w = (torch.randn(256, 14336, device=DEV, generator=g) * 0.02).to(torch.bfloat16)
        rq = ref.FloatQuantizer(bit, True, 'per_channel')
        pq = Q.FloatQuantizer(bit, True, 'per_channel')
        _same(pq.fake_quant_weight_dynamic(w), rq.fake_quant_weight_dynamic(w), f'{bit} wide rows')
    w = torch.randn(8, 64, device=DEV).to(torch.bfloat16)
    with pytest.raises(IndexError):
        ref.FloatQuantizer('e4m3', True, 'per_tensor').fake_quant_weight_dynamic(w)
    with pytest.raises(IndexError):
        Q.FloatQuantizer('e4m3', True, 'per_tensor').fake_quant_weight_dynamic(w)
    with pytest.raises(NotImplementedError):
        Q.FloatQuantizer('e4m3', True, 'per_channel').real_quant_weight_dynamic(w)

