"""CPU proofs of the identities the CUDA fast chain relies on (lightcompress_b200/csrc/fq_math.cuh).

1. Markstein division: with r = RN(1/s), q0 = RN(x*r), e = fma(-s, q0, x), q = fma(e, r, q0)  ==  RN(x/s).
2. (q + 1.5*2^23) - 1.5*2^23 == rint(q) (half-to-even) for |q| < 2^22.
3. bf16/fp16 integers up to 256/2048 are exact, so "+z, clamp, -z" == clamp(t, qmin-z, qmax-z).
"""
import ctypes

import numpy as np
import torch

from oracle import quant_oracle as qo


def _all_16bit(dtype):
    bits = torch.arange(0, 65536, dtype=torch.int32).to(torch.int16)
    v = bits.view(dtype).float()
    return v[torch.isfinite(v)]


def _check(xs, ss):
    xs = np.ascontiguousarray(xs, dtype=np.float32)
    ss = np.ascontiguousarray(ss, dtype=np.float32)
    return qo.lib().lco_check_markstein(xs.ctypes.data_as(ctypes.c_void_p), xs.size,
                                        ss.ctypes.data_as(ctypes.c_void_p), ss.size)


def test_markstein_bf16_and_fp16_operands():
    rng = np.random.default_rng(0)
    for dtype in (torch.bfloat16, torch.float16):
        xs = _all_16bit(dtype).numpy()
        cand = _all_16bit(dtype)
        cand = cand[(cand >= 1e-20) & (cand <= 1e20)].numpy()   # the kernel's fast-path window for s
        ss = rng.choice(cand, size=1500, replace=False)
        # |x| < 1e8 keeps x*r finite (the kernel guards the rest). The checker ignores quotients below 2^-40:
        # there the residual e = x - s*q0 may underflow, but every output format rounds such a quotient to 0.
        assert _check(xs[np.abs(xs) < 1e8], ss) == 0


def test_markstein_fp32_operands():
    rng = np.random.default_rng(1)
    xs = (rng.standard_normal(20000) * np.exp(rng.uniform(-20, 20, 20000))).astype(np.float32)
    ss = np.exp(rng.uniform(-40, 40, 4000)).astype(np.float32)

    # adversarial divisors: all-ones / near-power-of-two significands
    adv = np.array([0x3f7fffff, 0x3fffffff, 0x3f800001, 0x3f800000, 0x3effffff, 0x40490fdb, 0x3a83126f],
                   dtype=np.uint32).view(np.float32)
    assert _check(xs, np.concatenate([ss, adv])) == 0


def test_magic_round_is_rint():
    rng = np.random.default_rng(2)
    q = np.concatenate([rng.uniform(-300, 300, 200000), np.arange(-1000, 1000) + 0.5,
                        rng.uniform(-4e6, 4e6, 100000)]).astype(np.float32)
    assert qo.lib().lco_check_magic_round(q.ctypes.data_as(ctypes.c_void_p), q.size) == 0


def test_small_integers_exact_in_16bit():
    i = torch.arange(-256, 257, dtype=torch.float32)
    assert torch.equal(i.to(torch.bfloat16).float(), i)
    i = torch.arange(-2048, 2049, dtype=torch.float32)
    assert torch.equal(i.to(torch.float16).float(), i)
    # beyond 256 a bf16 rounding never re-enters [-256, 255]: integers 257..65535 stay >= 256
    j = torch.arange(257, 65536, dtype=torch.float32)
    assert (j.to(torch.bfloat16).float() >= 256).all()
