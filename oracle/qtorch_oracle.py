"""TEST INFRASTRUCTURE ONLY — never imported by lightcompress_b200/ or by a timed region.

Restatement of `qtorch.quant.float_quantize(x, exp, man, rounding='nearest')`, the call behind the reference's
FloatQuantizer.quant when `use_qtorch: True` (llmc/compression/quantization/quant.py:9, 1063-1070).

qtorch (QPyTorch, PyPI `qtorch`, un-pinned in the reference's requirements) is absent from /root/reference and from both
machines, and there is no network: this file restates the PUBLISHED algorithm of QPyTorch 0.3.0
(`qtorch/quant/quant_cpu/quant_cpu.cpp: float_quantize_nearest`, `bit_helper.cpp: round_bitwise_nearest, clip_exponent`)
as recalled — it could not be executed against the library here. PARITY UNPINNED: it is used only to QUANTIFY how far the
product's rounding (fp32 -> fp8 round-to-nearest-even with saturation at the format's maximum, i.e. `Tensor.to(float8_*)`
plus clamp, the rounding of the reference's own real-quant and Triton paths) is from what a qtorch-equipped reference
would compute in its fake-quant path (tests/test_qtorch_oracle.py).

Algorithm (per fp32 element, bit-level):
  * normal range: add half an ulp of the target mantissa to the fp32 bit pattern and truncate -> round to nearest, TIES
    AWAY FROM ZERO in magnitude; then clip_exponent: IEEE-style exponent range, bias 2^(e-1) - 1, NO code reserved for
    inf / NaN beyond that: the largest magnitude is (2 - 2^-m) * 2^(2^(e-1) - 1)  (e4m3: 240, e5m2: 57344) — values above
    saturate to it;
  * below 2^min_exp (min_exp = -(2^(e-1) - 2)): the value is shifted by +-2^min_exp, rounded the same way at the shifted
    exponent and shifted back (a fixed-point grid of step 2^(min_exp - m): gradual underflow).
"""
import numpy as np


def _round_bitwise_nearest(target, man_bits):
    mask = np.uint32((1 << (23 - man_bits)) - 1)
    half = np.uint32(1 << (23 - man_bits - 1))
    return (target + half) & ~mask


def _clip_exponent(exp_bits, man_bits, old, q):
    q = q.copy()
    qexp = ((q << np.uint32(1)) >> np.uint32(24)).astype(np.int64)
    max_store = (1 << (exp_bits - 1)) - 1 + 127
    min_store = -((1 << (exp_bits - 1)) - 2) + 127
    sign = old & np.uint32(0x80000000)
    max_man = np.uint32(((0xFFFFFFFF << 9) & 0xFFFFFFFF) >> 9 >> (23 - man_bits) << (23 - man_bits))
    max_num = np.uint32((max_store << 23)) | max_man
    over = (qexp > max_store) & (q != 0)
    q[over] = sign[over] | max_num
    under = (qexp < min_store) & (q != 0)
    min_num = np.uint32(min_store << 23)
    mid_num = np.uint32((min_store - 1) << 23)
    mag = q & np.uint32(0x7FFFFFFF)
    up = under & (mag > mid_num)
    q[up] = sign[up] | min_num
    dn = under & ~(mag > mid_num)
    q[dn] = sign[dn]
    return q


def float_quantize(x, exp_bits, man_bits):
    """x: float32 array (finite values) -> float32 array on the (exp_bits, man_bits) grid, rounding='nearest'."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    target = x.view(np.uint32)
    texp = (((target << np.uint32(1)) >> np.uint32(24)).astype(np.int64)) - 127
    min_exp = -((1 << (exp_bits - 1)) - 2)
    sub = texp < min_exp
    out = np.empty_like(x)
    # normal range
    qn = _clip_exponent(exp_bits, man_bits, target, _round_bitwise_nearest(target, man_bits))
    out[...] = qn.view(np.float32)
    # subnormal range of the target format
    if sub.any():
        shift_bits = (np.uint32((127 + min_exp) << 23) | (target[sub] & np.uint32(0x80000000)))
        shift = shift_bits.view(np.float32)
        val = (x[sub] + shift).astype(np.float32)
        qb = _round_bitwise_nearest(val.view(np.uint32), man_bits)
        out[sub] = (qb.view(np.float32) - shift).astype(np.float32)
    return out


def max_value(exp_bits, man_bits):
    return (2.0 - 2.0 ** (-man_bits)) * 2.0 ** ((1 << (exp_bits - 1)) - 1)
