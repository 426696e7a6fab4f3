"""CPU: properties of the block-wise FP8 inference oracle (oracle/fp8_infer_oracle.py = kernel.py:7-54, 140-242).
The Triton reference cannot run here, so the restatement is held to properties that follow from the kernel source:
exact scales, the e4m3 grid, the dequantisation error bound, and fp8_gemm == float matmul of the dequantised operands."""
import torch

from oracle import fp8_infer_oracle as fo


def test_act_quant_scale_codes_and_roundtrip():
    g = torch.Generator().manual_seed(0)
    x = (torch.randn(37, 512, generator=g) * torch.exp(torch.randn(37, 1, generator=g))).bfloat16()
    y, s = fo.act_quant(x, 128)
    assert y.dtype == torch.float8_e4m3fn and y.shape == x.shape and s.shape == (37, 4) and s.dtype == torch.float32
    amax = x.float().reshape(37, 4, 128).abs().amax(-1)
    assert torch.equal(s, amax / 448.0)
    yf = y.float().reshape(37, 4, 128)
    assert float(yf.abs().max()) <= 448.0
    # the group's extreme element maps to +-448 (or its neighbour after the fp32 division): |y| >= 416 there
    assert bool((yf.abs().amax(-1) >= 416).all())
    # e4m3 has 3 mantissa bits: relative error <= 2^-4 for normal values, absolute <= 2^-10 * s below 2^-6
    deq = yf * s[..., None]
    err = (deq - x.float().reshape(37, 4, 128)).abs()
    bound = torch.maximum(x.float().reshape(37, 4, 128).abs() * 2.0 ** -4, s[..., None] * 2.0 ** -10) * 1.0001
    assert bool((err <= bound).all())


def test_act_quant_zero_group_is_nan_like_the_reference():
    x = torch.zeros(2, 128)
    x[1, 5] = 3.0
    y, s = fo.act_quant(x, 128)
    assert float(s[0, 0]) == 0.0 and bool(y[0].float().isnan().all())      # 0 / 0 -> NaN code (kernel.py:24-26)
    assert float(y[1, 5].float()) == 448.0 and float(y[1].float().abs().sum()) == 448.0


def test_fp8_gemm_equals_matmul_of_dequantised_operands():
    g = torch.Generator().manual_seed(1)
    M, N, K = 70, 200, 384
    x = torch.randn(M, K, generator=g).bfloat16()
    w = (torch.randn(N, K, generator=g) * 0.05)
    a, a_s = fo.act_quant(x, 128)
    # per-block weight scales: [ceil(N/128), K/128]
    nb = (N + 127) // 128
    wp = torch.zeros(nb * 128, K)
    wp[:N] = w
    b_s = wp.reshape(nb, 128, K // 128, 128).abs().amax(dim=(1, 3)) / 448.0
    b = (wp.reshape(nb, 128, K // 128, 128) / b_s[:, None, :, None]).reshape(nb * 128, K)[:N].to(torch.float8_e4m3fn)
    c = fo.fp8_gemm(a, a_s, b, b_s, out_dtype=torch.float32)
    a_deq = (a.float().reshape(M, K // 128, 128) * a_s[..., None]).reshape(M, K).double()
    b_deq = (b.float().reshape(N, K // 128, 128) * b_s.repeat_interleave(128, 0)[:N][..., None]).reshape(N, K).double()
    ref = (a_deq @ b_deq.t())
    assert torch.allclose(c.double(), ref, rtol=1e-5, atol=1e-5 * float(ref.abs().max()))
    # and it approximates the float linear to fp8 accuracy
    full = x.double() @ w.double().t()
    assert float((c.double() - full).norm() / full.norm()) < 0.05
    # accumulation-order independent variant agrees with the plain fp32 one to fp32 round-off
    c2 = fo.fp8_gemm(a, a_s, b, b_s, out_dtype=torch.float32, exact_dot=False)
    assert torch.allclose(c, c2, rtol=1e-5, atol=1e-6 * float(ref.abs().max()))


def test_forward_func_bias_is_added_after_the_bf16_cast():
    g = torch.Generator().manual_seed(2)
    x = torch.randn(5, 256, generator=g).bfloat16()
    w = torch.randn(128, 256, generator=g).to(torch.float8_e4m3fn)
    ws = torch.rand(1, 2, generator=g) * 0.01
    bias = torch.randn(128, generator=g).bfloat16()
    y0 = fo.block_wise_fp8_forward_func(x, w, ws, 128, None)
    y1 = fo.block_wise_fp8_forward_func(x, w, ws, 128, bias)
    assert y0.dtype == torch.bfloat16 and torch.equal(y1, y0 + bias)


def test_markstein_division_by_448_is_correctly_rounded():
    """The kernel computes s = amax / 448 as q0 = a * r, s = fma(fma(-448, q0, a), r, q0) with r = RN(1 / 448): equal to
    the IEEE quotient for every bf16 value and 2e7 random fp32 values of the range the fast path accepts."""
    import numpy as np
    rng = np.random.default_rng(0)
    a = np.abs(rng.standard_normal(20_000_000).astype(np.float32)) * np.exp(rng.standard_normal(20_000_000) * 8).astype(np.float32)
    b = (np.arange(0, 0x7f80, dtype=np.uint32) << 16).view(np.float32)
    r = np.float32(1.0) / np.float32(448.0)
    for arr in (a, b):
        arr = arr[(arr >= np.float32(1e-17)) & (arr <= np.float32(1e22))]
        q0 = (arr * r).astype(np.float32)
        e = arr.astype(np.float64) - 448.0 * q0.astype(np.float64)          # exact in fp64
        assert (e.astype(np.float32).astype(np.float64) == e).all()           # and representable in fp32 (= the fma)
        s = (e * np.float64(r) + q0.astype(np.float64)).astype(np.float32)
        assert (s == (arr / np.float32(448.0)).astype(np.float32)).all()
