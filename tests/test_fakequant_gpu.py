"""GPU parity: the CUDA path, called through the quantizer API -> C ABI, against
  (1) the committed reference golden vectors (bit-exact),
  (2) the C oracle on larger seeded inputs (bit-exact),
  (3) size-independent properties at Llama-3-8B sizes (idempotence, codes/scales consistency).
"""
import pytest
import torch

from lightcompress_b200 import quant as Q
from oracle import quant_oracle as qo
from tests import helpers as H

pytestmark = pytest.mark.gpu
RECS = H.load_golden()
DEV = 'cuda:0'


@pytest.mark.parametrize('rec', RECS, ids=[H.case_id(r) for r in RECS])
def test_cuda_matches_reference_golden(rec):
    c = rec['case']
    x = H.case_input(c).to(DEV)
    if rec.get('raises') == 'ValueError':
        with pytest.raises(ValueError):
            H.make_quantizer(Q, c).fake_quant_weight_dynamic(x)
        return
    q = H.make_quantizer(Q, c)
    x0 = x.clone()
    fake = q.fake_quant_act_dynamic(x) if c.get('api') == 'act' else q.fake_quant_weight_dynamic(x)
    assert torch.equal(x, x0), 'input was mutated'
    assert fake.shape == x.shape and fake.dtype == x.dtype
    assert H.bits_equal(fake.cpu(), rec['fake']), 'fake-quant values differ from the reference'
    _, s, z, qmax, qmin = q.get_tensor_qparams(x)
    assert s.shape == rec['qp_scales'].shape and s.dtype == rec['qp_scales'].dtype
    assert H.bits_equal(s.cpu(), rec['qp_scales']), 'scales differ'
    if rec['qp_zeros'].numel() > 1 or not c.get('sym', True):
        assert H.bits_equal(z.cpu().reshape(rec['qp_zeros'].shape), rec['qp_zeros']), 'zeros differ'
    assert float(qmax) == float(rec['qmax']) and float(qmin) == float(rec['qmin'])
    if 'codes' in rec:
        codes, s2, z2 = q.real_quant_weight_dynamic(x)
        assert str(codes.dtype) == rec['codes_dtype']
        got = codes.view(torch.uint8).cpu() if rec['codes'].dtype == torch.uint8 and codes.element_size() == 1 \
            and 'float8' in rec['codes_dtype'] else codes.cpu().to(torch.int16)
        assert torch.equal(got, rec['codes']), 'codes differ'
        assert s2.shape == rec['scales'].shape and H.bits_equal(s2.cpu(), rec['scales'])
        if rec['zeros'] is None:
            assert z2 is None
        else:
            assert z2.dtype == rec['zeros'].dtype and torch.equal(z2.cpu(), rec['zeros'])
    if 'static_fake' in rec:
        args = {'scales': rec['static_scales'].to(DEV), 'zeros': rec['static_zeros'].to(DEV), 'qmax': qmax, 'qmin': qmin}
        assert H.bits_equal(q.fake_quant_weight_static(x, args).cpu(), rec['static_fake'])
        args = {'scales': rec['qp_scales'].to(DEV), 'zeros': rec['qp_zeros'].to(DEV), 'qmax': qmax, 'qmin': qmin}
        assert H.bits_equal(q.fake_quant_weight_static(x, args).cpu(), rec['static_fake_T'])


BIG = [
    dict(cls='int', bit=4, sym=False, gran='per_group', kwargs={'group_size': 128}, shape=(1024, 4096)),
    dict(cls='int', bit=4, sym=True, gran='per_group', kwargs={'group_size': 64}, shape=(512, 4096)),
    dict(cls='int', bit=8, sym=True, gran='per_token', shape=(4, 256, 4096), input='act'),
    dict(cls='int', bit=8, sym=False, gran='per_token', shape=(2, 64, 14336), input='act'),
    dict(cls='int', bit=8, sym=True, gran='per_channel', shape=(1024, 4096)),
    dict(cls='int', bit=8, sym=True, gran='per_channel', shape=(16, 40000)),      # streaming row kernel
    dict(cls='int', bit=8, sym=False, gran='per_tensor', shape=(1024, 2048)),
    dict(cls='int', bit=3, sym=False, gran='per_group', kwargs={'group_size': 32}, shape=(256, 1024)),
    dict(cls='int', bit=4, sym=False, gran='per_channel', shape=(333, 1000)),       # 1000 % 8 == 0, odd row count
    dict(cls='int', bit=4, sym=False, gran='per_channel', shape=(64, 1001)),        # generic kernel
    dict(cls='fp', bit='e4m3', gran='per_block', kwargs={'block_size': 128}, shape=(1408, 2048)),
    dict(cls='fp', bit='e4m3', gran='per_block', kwargs={'block_size': 128}, shape=(576, 2048)),
    dict(cls='fp', bit='e5m2', gran='per_group', kwargs={'group_size': 128}, shape=(256, 2048)),
    dict(cls='fp', bit='e4m3', gran='per_block', kwargs={'block_size': 128}, shape=(300, 264)),   # ragged both ways (TMA zero fill)
    dict(cls='fp', bit='e5m2', gran='per_block', kwargs={'block_size': 128}, shape=(130, 72)),
    dict(cls='fp', bit='e4m3', gran='per_block', kwargs={'block_size': 128}, shape=(64, 100)),    # cols % 8 != 0: register kernel
    dict(cls='int', bit=8, sym=True, gran='per_block', kwargs={'block_size': 128}, shape=(300, 520)),
]


@pytest.mark.parametrize('dtype', ['bfloat16', 'float16', 'float32'])
@pytest.mark.parametrize('c', BIG, ids=[f"{c['cls']}{c['bit']}-{c['gran']}-{'x'.join(map(str, c['shape']))}" for c in BIG])
def test_cuda_matches_oracle(c, dtype):
    c = dict(c, dtype=dtype, seed=7)
    x = H.case_input(c)
    x.view(-1)[::997] = 0
    q = H.make_quantizer(Q, c)
    xd = x.to(DEV)
    fake = q.fake_quant_act_dynamic(xd) if c.get('input') == 'act' else q.fake_quant_weight_dynamic(xd)
    ref, _, _ = H.oracle_case(c, x, qo.OUT_FAKE)
    assert H.bits_equal(fake.cpu(), ref)
    if x.dim() == 2:
        codes, s, z = q.real_quant_weight_dynamic(xd)
        rc, rs, rz = H.oracle_case(c, x, H.code_mode(c))
        assert torch.equal(codes.view(torch.uint8).cpu() if codes.element_size() == 1 else codes.cpu(),
                           rc.view(torch.uint8) if rc.element_size() == 1 else rc)
        assert H.bits_equal(s.cpu().reshape(-1), rs.reshape(-1))
        if rz is not None:
            assert torch.equal(z.cpu().reshape(-1).float(), rz.reshape(-1).float())


def test_nonfinite_inputs_take_the_literal_chain():
    c = dict(cls='int', bit=4, sym=False, gran='per_group', kwargs={'group_size': 128}, shape=(8, 512), dtype='bfloat16', seed=3)
    x = H.case_input(c)
    x[0, 5] = float('inf')
    x[1, 7] = float('nan')
    x[2, 300] = float('-inf')
    q = H.make_quantizer(Q, c)
    fake = q.fake_quant_weight_dynamic(x.to(DEV))
    ref, _, _ = H.oracle_case(c, x, qo.OUT_FAKE)
    assert H.equal_nan_aware(fake.cpu(), ref)
    assert int(torch.isnan(ref).sum()) >= 256  # the inf / nan groups really are poisoned in the reference model


def test_full_size_properties_llama3_8b():
    """At BASELINE sizes: fake-quant is idempotent on its own grid and equals dequantised codes."""
    torch.manual_seed(0)
    for shape in [(4096, 4096), (14336, 4096), (4096, 14336)]:
        w = (torch.randn(shape, device=DEV) * 0.02).to(torch.bfloat16)
        q = Q.IntegerQuantizer(4, False, 'per_group', group_size=128)
        fq = q.fake_quant_weight_dynamic(w)
        codes, s, z = q.real_quant_weight_dynamic(w)
        assert codes.dtype == torch.int32 and int(codes.min()) >= 0 and int(codes.max()) <= 15
        deq = ((codes.view(-1, 128).to(torch.bfloat16) - z.view(-1, 1).to(torch.bfloat16)) * s.view(-1, 1)).view(shape)
        assert torch.equal(deq, fq)
        # every group reaches code 0 (min -> 0); the max lands on 15 or, when round(mn/s) and round(mx/s) both round
        # towards the range, on 14 (the reference itself does this for ~2.5 % of the groups of such a tensor)
        g = codes.view(-1, 128)
        assert bool((g.min(dim=1).values == 0).all()) and bool((g.max(dim=1).values >= 14).all())
        # quantising the fake-quant tensor with the same (static) qparams is the identity
        _, s0, z0, qmax, qmin = q.get_tensor_qparams(w)
        again = q.fake_quant_weight_static(fq, {'scales': s0, 'zeros': z0, 'qmax': qmax, 'qmin': qmin})
        assert torch.equal(again, fq)


def test_act_per_token_full_size():
    torch.manual_seed(1)
    x = torch.randn(8, 2048, 4096, device=DEV, dtype=torch.bfloat16)
    q = Q.IntegerQuantizer(8, True, 'per_token')
    y = q.fake_quant_act_dynamic(x)
    amax = x.float().abs().amax(-1, keepdim=True)
    # error model of the reference's bf16 chain: rd(x/s) has a step of 0.5 at |code| >= 64 (<= 0.25 code), rint adds
    # <= 0.5 code, s itself and the dequant product carry 2^-9 relative each
    assert bool(((y.float() - x.float()).abs() <= amax / 127 * 0.76 + 6e-3 * amax).all())
    _, s, _, _, _ = q.get_tensor_qparams(x)
    assert s.shape == (8, 2048, 1)


def test_fp8_block_roundtrip():
    torch.manual_seed(2)
    w = (torch.randn(1408, 2048, device=DEV) * 0.02).to(torch.bfloat16)
    codes, scales = Q.weight_cast_to_fp8(w, 128)
    assert codes.dtype == torch.float8_e4m3fn and scales.shape == (11, 16) and scales.dtype == torch.float32
    back = Q.weight_cast_to_bf16(codes, scales, 128)
    q = Q.FloatQuantizer('e4m3', True, 'per_block', block_size=128, use_qtorch=True)
    assert torch.equal(back, q.fake_quant_weight_dynamic(w))
    # re-quantising the bf16 round trip: the block scale moves by at most the bf16 rounding of 448*s, so every code
    # moves by at most one fp8 step (it is NOT exactly idempotent in the reference either: `back` is rounded to bf16)
    codes2, scales2 = Q.weight_cast_to_fp8(back, 128)
    assert torch.allclose(scales2, scales, rtol=2 ** -8)
    d = (codes2.float() - codes.float()).abs()
    assert bool((d <= codes.float().abs() * 0.13 + 2 ** -9).all())


def test_multi_tensor_launch_equals_single_launches():
    torch.manual_seed(5)
    for dtype, bit, sym, gs in [(torch.bfloat16, 4, False, 128), (torch.bfloat16, 8, True, 64), (torch.float16, 4, True, 128),
                                (torch.float32, 4, False, 128), (torch.bfloat16, 3, False, 32)]:
        shapes = [(64, 256), (8, 1024), (130, 384), (1, 128), (300, 128)]
        ws = [(torch.randn(s, device=DEV) * 0.02).to(dtype) for s in shapes]
        q = Q.IntegerQuantizer(bit, sym, 'per_group', group_size=gs)
        multi = q.fake_quant_weights_dynamic(ws)
        for w, m in zip(ws, multi):
            assert H.bits_equal(m.cpu(), q.fake_quant_weight_dynamic(w).cpu())
    # 20 tensors -> two launches; ineligible layout -> per-tensor path
    ws = [(torch.randn(16, 256, device=DEV) * 0.02).to(torch.bfloat16) for _ in range(20)]
    q = Q.IntegerQuantizer(4, False, 'per_group', group_size=128)
    for w, m in zip(ws, q.fake_quant_weights_dynamic(ws)):
        assert H.bits_equal(m.cpu(), q.fake_quant_weight_dynamic(w).cpu())
    qc = Q.IntegerQuantizer(8, True, 'per_channel')
    for w, m in zip(ws, qc.fake_quant_weights_dynamic(ws)):
        assert H.bits_equal(m.cpu(), qc.fake_quant_weight_dynamic(w).cpu())


def test_per_block_fp8_nonfinite_tile_takes_the_literal_chain():
    """A tile holding inf / NaN goes through the out-of-line literal chain of the TMA-staged kernel; other tiles are untouched."""
    c = dict(cls='fp', bit='e4m3', gran='per_block', kwargs={'block_size': 128}, shape=(256, 384), dtype='bfloat16', seed=11)
    x = H.case_input(c)
    x[3, 5] = float('inf')
    x[200, 300] = float('nan')
    q = H.make_quantizer(Q, c)
    fake = q.fake_quant_weight_dynamic(x.to(DEV)).cpu()
    ref, _, _ = H.oracle_case(c, x, qo.OUT_FAKE)
    assert H.equal_nan_aware(fake, ref)


@pytest.mark.parametrize('bit', ['e4m3', 'e5m2'])
@pytest.mark.parametrize('dtype', [torch.bfloat16, torch.float16])
def test_block_fp8_multi_launch_equals_single_launches(bit, dtype):
    """lcb_fakequant_block_fp8_multi (64 weights per launch: MoE experts; outputs of equal shape share one allocation) ==
    the per-tensor calls, bit for bit."""
    torch.manual_seed(9)
    shapes = [(1408, 2048), (2048, 1408), (576, 2048), (130, 72), (128, 128), (1, 8), (300, 264)] * 11    # 77 -> two launches (64 + 13)
    ws = [(torch.randn(s, device=DEV) * (0.02 * (i % 9 + 1))).to(dtype) for i, s in enumerate(shapes)]
    ws[4][5, 7] = float('inf')     # one tile on the literal chain
    q = Q.FloatQuantizer(bit, True, 'per_block', block_size=128, use_qtorch=True)
    for w, m in zip(ws, q.fake_quant_weights_dynamic(ws)):
        assert H.equal_nan_aware(m.cpu(), q.fake_quant_weight_dynamic(w).cpu())
    for w, (c, s, z) in zip(ws, q.real_quant_weights_dynamic(ws)):
        c1, s1, z1 = q.real_quant_weight_dynamic(w)
        assert c.dtype == c1.dtype and z is None and z1 is None and s.shape == s1.shape and s.dtype == s1.dtype
        assert torch.equal(c.view(torch.uint8), c1.view(torch.uint8))
        assert H.equal_nan_aware(s.cpu(), s1.cpu())
    # outside the envelope (fp32 weights, odd columns): the per-tensor path, same results
    odd = [torch.randn(64, 100, device=DEV).to(dtype), torch.randn(256, 384, device=DEV)]
    for w in odd:
        assert H.equal_nan_aware(q.fake_quant_weights_dynamic([w])[0].cpu(), q.fake_quant_weight_dynamic(w).cpu())
        assert torch.equal(q.real_quant_weights_dynamic([w])[0][0].view(torch.uint8), q.real_quant_weight_dynamic(w)[0].view(torch.uint8))


@pytest.mark.parametrize('shape', [(1408, 2048), (200, 272), (130, 72), (1, 16)])
@pytest.mark.parametrize('fp8', [torch.float8_e4m3fn, torch.float8_e5m2])
def test_weight_cast_to_bf16_matches_torch(shape, fp8):
    """lcb_fp8_block_dequant (vector path for cols % 16 == 0, element path otherwise) == fp8.float() * scale[block] -> bf16
    (quant.py:18-31 / kernel.py:86-138), bit for bit."""
    g = torch.Generator().manual_seed(shape[0] + shape[1])
    codes = (torch.randn(shape, generator=g) * 40).to(fp8)
    nb = ((shape[0] + 127) // 128, (shape[1] + 127) // 128)
    scales = torch.rand(nb, generator=g) * 0.01 + 1e-4
    ref = (codes.float() * scales.repeat_interleave(128, 0)[:shape[0]].repeat_interleave(128, 1)[:, :shape[1]]).to(torch.bfloat16)
    got = Q.weight_cast_to_bf16(codes.to(DEV), scales.to(DEV), 128).cpu()
    assert H.bits_equal(got, ref)
