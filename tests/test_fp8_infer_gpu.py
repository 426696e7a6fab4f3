"""GPU parity of the block-wise FP8 inference kernels (csrc/fp8_infer.cu through the C ABI) against the CPU oracle
(oracle/fp8_infer_oracle.py = kernel.py:7-54, 140-242; module_utils.py:41-46).
Bars: act_quant — the kernel divides with PTX div.full.f32 like the reference's Triton kernel (bit-exact against THAT in
tests/test_ref_kernels_gpu.py); this oracle divides with IEEE rounding, so here scales agree to one fp32 ulp and codes
everywhere except on e4m3 rounding ties (each within one code step); fp8_gemm — floating point: the tensor pipe's accumulation inside a
128-deep block is not IEEE-ordered, so the fp32 result is held to 2e-6 of the row/column scale product (|a_s| |b_s|
448^2 * 128 bounds a block sum) and the bf16 / fp16 output to one ulp of the output dtype on every element (north star:
1e-3 relative for floating-point outputs)."""
import pytest
import torch

from lightcompress_b200 import fp8_infer as fi
from oracle import fp8_infer_oracle as fo

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'
F8 = torch.float8_e4m3fn


def _acts(shape, dtype, seed):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(*shape, generator=g) * torch.exp(torch.randn(*shape[:-1], 1, generator=g))
    x[..., 3] *= 30.0     # an outlier channel, as activations have
    return x.to(dtype)


def _codes_equal(a, b, max_frac=3e-3):
    """Same NaN pattern; codes equal except on rounding ties of x / s (div.full.f32 vs IEEE division), one step apart."""
    a, b = a.view(torch.uint8), b.view(torch.uint8)
    nan_a, nan_b = (a & 0x7f) == 0x7f, (b & 0x7f) == 0x7f
    if not bool(torch.equal(nan_a, nan_b)):
        return False
    x, y = a[~nan_a].to(torch.int16), b[~nan_b].to(torch.int16)
    ne = x != y
    if not bool(ne.any()):
        return True
    return float(ne.float().mean()) <= max_frac and int((x[ne] - y[ne]).abs().max()) <= 1


def _scales_close(s, s_ref):
    """fp32 scales within one ulp (a * rcp(448) against the IEEE quotient), NaN where the oracle has NaN."""
    na, nb = s.isnan(), s_ref.isnan()
    if not bool(torch.equal(na, nb)):
        return False
    a = s[~na].contiguous().view(torch.int32).long()
    b = s_ref[~nb].contiguous().view(torch.int32).long()
    return a.numel() == 0 or int((a - b).abs().max()) <= 1


@pytest.mark.parametrize('dtype', [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize('shape,bs', [((3, 7, 512), 128), ((1000, 128), 128), ((64, 2048), 128), ((33, 768), 64),
                                      ((5, 512), 256), ((17, 96), 32), ((9, 80), 16), ((4, 40), 8), ((1, 128), 128)])
def test_act_quant_vs_ieee_oracle(dtype, shape, bs):
    x = _acts(shape, dtype, 0)
    y_ref, s_ref = fo.act_quant(x, bs)
    y, s = fi.act_quant(x.to(DEV), bs)
    assert y.dtype == F8 and y.shape == x.shape and s.dtype == torch.float32 and s.shape == s_ref.shape
    assert _scales_close(s.cpu(), s_ref)
    assert _codes_equal(y.cpu(), y_ref, max_frac=3e-3 if x.numel() >= 4096 else 0.05)


def test_act_quant_edge_cases():
    x = torch.zeros(4, 256, dtype=torch.bfloat16)
    x[1, 0] = 1e-30            # denormal-scale group: IEEE-division path
    x[2, 130] = -3e38          # huge
    x[3, 7] = float('inf')     # inf / inf = NaN for that element, 0 elsewhere
    y_ref, s_ref = fo.act_quant(x, 128)
    y, s = fi.act_quant(x.to(DEV), 128)
    assert _scales_close(s.cpu().flatten(), s_ref.flatten())
    assert _codes_equal(y.cpu(), y_ref, max_frac=0.05)
    # empty input and the reference's assertions
    e, es = fi.act_quant(torch.empty(0, 128, dtype=torch.bfloat16, device=DEV))
    assert e.shape == (0, 128) and es.shape == (0, 1)
    with pytest.raises(AssertionError):
        fi.act_quant(torch.zeros(4, 100, device=DEV))
    with pytest.raises(AssertionError):
        fi.act_quant(torch.zeros(4, 256, device=DEV).t())
    with pytest.raises(RuntimeError):
        fi.act_quant(torch.zeros(4, 128))


def test_act_quant_large_roundtrip_property():
    """Full-size activation batch (8192 x 7168 bf16): scales exact against torch on the GPU, dequantisation error bound."""
    g = torch.Generator(device=DEV).manual_seed(3)
    x = torch.randn(8192, 7168, generator=g, device=DEV, dtype=torch.bfloat16)
    y, s = fi.act_quant(x)
    xf = x.float().view(8192, 56, 128)
    amax = xf.abs().amax(-1)
    assert _scales_close(s.flatten(), (amax / torch.full_like(amax, 448.0)).flatten())
    err = (y.float().view(8192, 56, 128) * s[..., None] - xf).abs()
    bound = torch.maximum(xf.abs() * 2.0 ** -4, s[..., None] * 2.0 ** -10) * 1.0001
    assert bool((err <= bound).all())


def _gemm_case(M, N, K, seed, lead=None):
    g = torch.Generator().manual_seed(seed)
    nkb, nb = (K + 127) // 128, (N + 127) // 128
    a = (torch.randn(M, K, generator=g) * 100).clamp(-448, 448).to(F8)
    b = (torch.randn(N, K, generator=g) * 100).clamp(-448, 448).to(F8)
    a_s = torch.rand(M, nkb, generator=g) * 0.02 + 1e-3
    b_s = torch.rand(nb, nkb, generator=g) * 0.002 + 1e-4
    if lead is not None:
        a = a.reshape(*lead, K)
        a_s = a_s.reshape(*lead, nkb)
    return a, a_s, b, b_s


def _check_gemm(a, a_s, b, b_s, out_dtype, bias=None):
    ref32 = fo.fp8_gemm(a, a_s, b, b_s, out_dtype=torch.float32)
    got = fi.fp8_gemm(a.to(DEV), a_s.to(DEV), b.to(DEV), b_s.to(DEV), out_dtype=out_dtype,
                      bias=None if bias is None else bias.to(DEV)).cpu()
    assert got.dtype == out_dtype and got.shape == ref32.shape
    M = a.numel() // a.size(-1)
    # magnitude scale of a block sum for (row, col): sum_kb a_s * b_s * (448^2 * 128) is a hard bound; observed error is
    # held to 2e-6 of the root-sum-square analogue
    mag = (a.float().reshape(M, -1).norm(dim=1, keepdim=True) * float(a_s.max())) * \
          (b.float().norm(dim=1)[None, :] * float(b_s.max()))
    mag = mag.reshape(ref32.shape)
    ref = ref32.to(out_dtype)
    pre = ref.float().abs()       # a one-ulp flip of the pre-bias value survives the bias add at that magnitude
    if bias is not None:
        ref = (ref + bias.to(out_dtype)).to(out_dtype) if bias.dtype == out_dtype else (ref.float() + bias.float()).to(out_dtype)
    if out_dtype == torch.float32:
        assert bool(((got - ref).abs() <= 2e-6 * mag + 1e-30).all()), float(((got - ref).abs() / mag).max())
    else:
        ulp = {torch.bfloat16: 2.0 ** -7, torch.float16: 2.0 ** -10}[out_dtype]
        d = (got.float() - ref.float()).abs()
        assert bool((d <= ulp * (pre + ref.float().abs()) + 2e-6 * mag).all())
        assert float((d > 0).float().mean()) < 0.02      # a different rounding only where the fp32 value sits on a tie


@pytest.mark.parametrize('M,N,K', [(128, 128, 128), (256, 384, 512), (200, 136, 384), (1, 128, 256), (77, 520, 1040),
                                   (130, 72, 144), (512, 1408, 2048), (4096, 256, 7168)])
def test_fp8_gemm_fp32_out(M, N, K):
    _check_gemm(*_gemm_case(M, N, K, M + N + K), torch.float32)


@pytest.mark.parametrize('out_dtype', [torch.bfloat16, torch.float16])
@pytest.mark.parametrize('M,N,K', [(300, 264, 640), (64, 2048, 1408)])
def test_fp8_gemm_16bit_out_and_bias(out_dtype, M, N, K):
    a, a_s, b, b_s = _gemm_case(M, N, K, 7)
    _check_gemm(a, a_s, b, b_s, out_dtype)
    g = torch.Generator().manual_seed(8)
    _check_gemm(a, a_s, b, b_s, out_dtype, bias=torch.randn(N, generator=g).to(out_dtype))
    _check_gemm(a, a_s, b, b_s, out_dtype, bias=torch.randn(N, generator=g))


def test_fp8_gemm_leading_dims_default_dtype_and_errors():
    a, a_s, b, b_s = _gemm_case(6 * 50, 256, 256, 9, lead=(6, 50))
    _check_gemm(a, a_s, b, b_s, torch.bfloat16)
    old = torch.get_default_dtype()
    try:
        torch.set_default_dtype(torch.bfloat16)
        c = fi.fp8_gemm(a.to(DEV), a_s.to(DEV), b.to(DEV), b_s.to(DEV))
        assert c.dtype == torch.bfloat16 and c.shape == (6, 50, 256)
    finally:
        torch.set_default_dtype(old)
    with pytest.raises(AssertionError):
        fi.fp8_gemm(a.to(DEV).transpose(0, 1), a_s.to(DEV), b.to(DEV), b_s.to(DEV))
    with pytest.raises(TypeError):
        fi.fp8_gemm(a.to(DEV).float(), a_s.to(DEV), b.to(DEV), b_s.to(DEV))
    with pytest.raises(RuntimeError):
        fi.fp8_gemm(a, a_s, b, b_s)
    e = fi.fp8_gemm(torch.empty(0, 256, dtype=F8, device=DEV), torch.empty(0, 2, device=DEV), b.to(DEV), b_s.to(DEV),
                    out_dtype=torch.bfloat16)
    assert e.shape == (0, 256)


def test_linearity_in_the_scales_full_size():
    """Size-independent property at a DeepSeek-V3-sized projection (4096 x 7168 -> 2048): doubling a_s doubles the fp32
    output exactly (power-of-two scaling commutes with every rounding), and b_s likewise."""
    a, a_s, b, b_s = [t.to(DEV) for t in _gemm_case(4096, 2048, 7168, 11)]
    c1 = fi.fp8_gemm(a, a_s, b, b_s, out_dtype=torch.float32)
    c2 = fi.fp8_gemm(a, a_s * 2, b, b_s, out_dtype=torch.float32)
    c3 = fi.fp8_gemm(a, a_s, b, b_s * 4, out_dtype=torch.float32)
    assert torch.equal(c2, c1 * 2) and torch.equal(c3, c1 * 4)
    # against cuBLAS on the dequantised operands
    a_d = (a.float().view(4096, 56, 128) * a_s[..., None]).view(4096, 7168)
    b_d = (b.float().view(2048, 56, 128) * b_s.repeat_interleave(128, 0)[..., None]).view(2048, 7168)
    ref = a_d.double() @ b_d.double().t()
    assert float((c1.double() - ref).norm() / ref.norm()) < 1e-6


@pytest.mark.parametrize('bias', [False, True])
def test_block_wise_forward_and_module(bias):
    g = torch.Generator().manual_seed(12)
    lin = torch.nn.Linear(1024, 640, bias=bias).to(torch.bfloat16)
    x = _acts((3, 40, 1024), torch.bfloat16, 13)
    m = fi.LlmcFp8Linear.new(lin.to(DEV), 128)
    assert m.weight.dtype == F8 and m.weight_scale_inv.shape == (5, 8)
    y = m(x.to(DEV)).cpu()
    ref = fo.block_wise_fp8_forward_func(x, m.weight.cpu(), m.weight_scale_inv.cpu(), 128,
                                         None if not bias else m.bias.cpu())
    assert y.dtype == torch.bfloat16 and y.shape == (3, 40, 640)
    # the CPU oracle quantizes the activation with IEEE divisions, the kernel with div.full.f32 like the reference's Triton
    # kernel (bit-exact against THAT composition in tests/test_ref_kernels_gpu.py): a few activation codes sit one e4m3
    # step apart, so the outputs agree to fp8 accuracy, not to an output ulp
    d = (y.float() - ref.float()).abs()
    assert float(d.norm() / ref.float().norm()) < 5e-3
    assert bool((d <= 0.05 * float(ref.float().abs().max())).all())
    # and the quantised linear tracks the float one to fp8 accuracy
    full = lin.float().cpu()(x.float())
    assert float((y.float() - full).norm() / full.norm()) < 0.06
    del g


def test_c_abi_status_codes_surface_as_exceptions():
    """K % 16 != 0 (TMA row pitch) and unsupported group sizes are refused by the C ABI, not silently mis-computed."""
    a = torch.zeros(4, 136, device=DEV).to(F8)
    b = torch.zeros(128, 136, device=DEV).to(F8)
    with pytest.raises(Exception, match='lcb_fp8_gemm'):
        fi.fp8_gemm(a, torch.ones(4, 2, device=DEV), b, torch.ones(1, 2, device=DEV), out_dtype=torch.bfloat16)
    with pytest.raises(Exception, match='lcb_act_quant_fp8'):
        fi.act_quant(torch.zeros(4, 96, device=DEV, dtype=torch.bfloat16), 48)


@pytest.mark.parametrize('out_dtype', [torch.float32, torch.bfloat16])
@pytest.mark.parametrize('M,N,K', [(1, 2048, 7168), (16, 7168, 2048), (100, 520, 4096), (128, 4096, 4096), (7, 136, 4096)])
def test_fp8_gemm_split_k_decode_shapes(out_dtype, M, N, K):
    """Few output tiles: lcb_fp8_gemm_ws splits the K blocks of a tile over the idle SMs (fp32 reductions into a
    workspace + a finishing launch). Same bar as the one-pass kernel, with and without bias, and against it."""
    from lightcompress_b200 import _cabi
    lib = _cabi.load()
    assert lib.lcb_fp8_gemm_workspace_bytes(M, N, K) >= M * N * 4, 'this shape is expected to take the split-K path'
    assert lib.lcb_fp8_gemm_workspace_bytes(4096, 4096, 4096) == 0 and lib.lcb_fp8_gemm_workspace_bytes(128, 7168, 2048) == 0
    a, a_s, b, b_s = _gemm_case(M, N, K, M + K)
    _check_gemm(a, a_s, b, b_s, out_dtype)
    g = torch.Generator().manual_seed(3)
    _check_gemm(a, a_s, b, b_s, out_dtype, bias=torch.randn(N, generator=g).to(torch.bfloat16 if out_dtype != torch.float32
                                                                                 else torch.float32))
    # against the one-pass entry (no workspace) on the same operands: fp32 sums in another order
    ad, asd, bd, bsd = a.to(DEV), a_s.to(DEV), b.to(DEV), b_s.to(DEV)
    one = torch.empty(M, N, dtype=torch.float32, device=DEV)
    st = lib.lcb_fp8_gemm(ad.data_ptr(), asd.data_ptr(), bd.data_ptr(), bsd.data_ptr(), one.data_ptr(), None, M, N, K, 0, 0,
                          torch.cuda.current_stream().cuda_stream)
    assert st == 0
    scale = float(one.abs().max())
    for _ in range(3):      # the workspace is left zero: back-to-back calls on one buffer agree
        two = fi.fp8_gemm(ad, asd, bd, bsd, out_dtype=torch.float32)
        assert float((one - two).abs().max()) <= 2e-6 * scale
    ws = fi._zero_workspace(torch.device(DEV), 16)
    assert int(ws.count_nonzero()) == 0, 'split-K calls must leave their workspace zero-filled'
