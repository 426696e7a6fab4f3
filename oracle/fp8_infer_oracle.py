"""TEST INFRASTRUCTURE — CPU restatement of the reference's block-wise FP8 inference kernels
(llmc/compression/quantization/kernel.py). Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import it.

The reference functions are Triton kernels and cannot run without a GPU. Pinned by tests/golden/fp8_infer_golden.pt.gz:
outputs of the reference's own kernels produced on the B200 box (tests/golden/make_fp8_infer_golden.py) — against this
restatement's IEEE divisions the Triton scales (div.full.f32) sit at most one fp32 ulp apart and < 0.1 % of the codes
(exact e4m3 rounding ties) one code step apart; fp8_gemm on identical operands agrees to 7e-8; weight_cast_to_bf16 is
bit-exact (tests/test_fp8_infer_golden.py). The pieces that the reference ALSO states in torch (weight_cast_to_fp8 /
weight_cast_to_bf16, quant.py:18-43) are pinned by tests/golden/quant_golden.pt.gz as well.
"""
import torch

F8 = torch.float8_e4m3fn


def act_quant(x, block_size=128):
    """kernel.py:7-29: per program (= group of block_size consecutive elements):
    x = load.to(fp32); s = max(|x|) / 448.0; y = (x / s).to(e4m3); store y, s."""
    assert x.size(-1) % block_size == 0
    xf = x.detach().float().reshape(-1, block_size)
    s = xf.abs().amax(dim=1) / 448.0
    y = (xf / s[:, None]).to(F8)
    return y.reshape(x.shape), s.reshape(*x.shape[:-1], x.size(-1) // block_size)


def fp8_gemm(a, a_s, b, b_s, out_dtype=torch.bfloat16, exact_dot=True):
    """kernel.py:186-211: accumulator(fp32) += dot(a_kb, b_kb) * a_s[:, kb] * b_s[n // 128, kb] for every 128-deep K
    block, then the cast to the output dtype. The block dot products are taken in float64 (`exact_dot`: products of
    two e4m3 values are exact in fp32 and a 128-term sum is exact in fp64), so the oracle is independent of any
    accumulation order; promotion and the running sum are fp32 like the kernel's."""
    K = a.size(-1)
    M = a.numel() // K
    N = b.size(0)
    af = a.reshape(M, K).float()
    bf = b.float()
    nkb = (K + 127) // 128
    a_s = a_s.reshape(M, nkb).float()
    b_s = b_s.reshape((N + 127) // 128, nkb).float()
    acc = torch.zeros(M, N, dtype=torch.float32)
    bs_cols = b_s.repeat_interleave(128, dim=0)[:N]            # [N, nkb]: offs_n // 128
    for kb in range(nkb):
        sl = slice(kb * 128, min(K, (kb + 1) * 128))
        if exact_dot:
            d = (af[:, sl].double() @ bf[:, sl].double().t()).float()
        else:
            d = af[:, sl] @ bf[:, sl].t()
        acc += d * a_s[:, kb:kb + 1] * bs_cols[:, kb][None, :]
    return acc.to(out_dtype).reshape(*a.shape[:-1], N)


def block_wise_fp8_forward_func(x, w, w_scale, block_size, bias):
    """module_utils.py:41-46."""
    xq, scale = act_quant(x, block_size)
    y = fp8_gemm(xq, scale, w, w_scale).to(torch.bfloat16)
    if bias is not None:
        y += bias
    return y
