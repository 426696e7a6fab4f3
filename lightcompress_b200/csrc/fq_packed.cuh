// fq_packed.cuh — the 16-bit packed element chains and the cheap bf16 qparams of the grouped quant-dequant kernels.
//
// The reference chain (quant.py:699-717) on a bf16 / fp16 tensor is "fp32 op, round to T" per torch op:
//     t1 = rd(x / s); t2 = rint(t1); q = clamp(rd(t2 + z), qmin, qmax); out = rd(rd(q - z) * s)
// For integer ranges that fit T exactly this equals
//     y = rd(x / s); t = rint(clamp(y, qmin - z, qmax - z)); out = rd(t * s)
// (rint is monotonic and the bounds are integers, so clamp and rint commute; "+z ... -z" is exact on small integers).
// Everything after the division therefore runs on PACKED 16-bit pairs: clamp = 2 HMNMX2, rint = magic add/sub in T
// (1.5*2^7 for bf16 when |t| < 64, 1.5*2^10 for fp16 when |t| < 512), the dequant product t*s of two T values is exact
// in fp32, so one packed multiply with a single rounding to T is the reference's rd(fp32 product).
//
// Division. bf16: x and s both carry 8 significant bits, so x/s can never equal a bf16 rounding midpoint (x = mid*s
// would need >= 9 bits) and is at least 2^-17 (relative) away from every midpoint, while x*RCP(s) is within 2^-22 of
// x/s for any reciprocal good to 2 ulp. Hence rd_bf16(x * r) == rd_bf16(RN(x / s)) with r = MUFU.RCP(s): one FMUL.
// The same argument covers s = rd(d / div) (div an integer whose odd part is <= 1024) and the zero-point quotient
// rd(mn / s). fp16 (11 bits) leaves only a factor-2 margin, so fp16 keeps the Markstein-corrected quotient.
// tests/test_fq_math_identity.py checks both claims exhaustively / by dense sampling on the CPU.
#pragma once
#include "fq_math.cuh"

namespace lcb {

// per-unit constants of the packed chains
struct UnitP {
  uint32_t lo2, hi2, s2;  // (qmin - z, qmin - z), (qmax - z, ...), (s_out, s_out) as packed T
};

__device__ __forceinline__ uint32_t pack_bf16x2(float hi, float lo) {
  uint32_t d;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  return d;
}
__device__ __forceinline__ uint32_t pack_f16x2(float hi, float lo) {
  uint32_t d;
  asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  return d;
}

#define LCB_P2(name, op)                                                        \
  __device__ __forceinline__ uint32_t name(uint32_t a, uint32_t b) {            \
    uint32_t d;                                                                 \
    asm(op " %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));                          \
    return d;                                                                   \
  }
LCB_P2(bf2_max, "max.bf16x2")
LCB_P2(bf2_min, "min.bf16x2")
LCB_P2(bf2_add, "add.rn.bf16x2")
LCB_P2(bf2_sub, "sub.rn.bf16x2")
LCB_P2(bf2_mul, "mul.rn.bf16x2")
LCB_P2(h2_max, "max.f16x2")
LCB_P2(h2_min, "min.f16x2")
LCB_P2(h2_add, "add.rn.f16x2")
LCB_P2(h2_sub, "sub.rn.f16x2")
LCB_P2(h2_mul, "mul.rn.f16x2")
#undef LCB_P2

// ---- bf16, integer codes, |qmin - z|, |qmax - z| < 64 : 5 instructions per element ---------------------------------
__device__ __forceinline__ uint32_t fq2_bf16_small(uint32_t w, float r, const UnitP& p) {
  const float q0 = __fmul_rn(__uint_as_float(w << 16), r);
  const float q1 = __fmul_rn(__uint_as_float(w & 0xffff0000u), r);
  uint32_t y = pack_bf16x2(q1, q0);                 // rd(x / s)
  y = bf2_min(bf2_max(y, p.lo2), p.hi2);            // clamp (commutes with rint)
  const uint32_t M = 0x43404340u;                   // (192, 192) = 1.5 * 2^7: RNE to integer for |y| < 64
  y = bf2_sub(bf2_add(y, M), M);
  return bf2_mul(y, p.s2);                          // rd(t * s)
}

// ---- bf16, integer codes up to 8 bit: clamp packed, rint + dequant in fp32 ----------------------------------------
__device__ __forceinline__ uint32_t fq2_bf16_wide(uint32_t w, float r, const UnitP& p, float s_out) {
  const float q0 = __fmul_rn(__uint_as_float(w << 16), r);
  const float q1 = __fmul_rn(__uint_as_float(w & 0xffff0000u), r);
  uint32_t y = pack_bf16x2(q1, q0);
  y = bf2_min(bf2_max(y, p.lo2), p.hi2);
  const float MAGIC = 12582912.0f;
  float t0 = __fsub_rn(__fadd_rn(__uint_as_float(y << 16), MAGIC), MAGIC);
  float t1 = __fsub_rn(__fadd_rn(__uint_as_float(y & 0xffff0000u), MAGIC), MAGIC);
  return pack_bf16x2(__fmul_rn(t1, s_out), __fmul_rn(t0, s_out));
}

// ---- fp16, integer codes, |qmin - z|, |qmax - z| < 512 -------------------------------------------------------------
__device__ __forceinline__ uint32_t fq2_f16(uint32_t w, float s, float r, const UnitP& p) {
  const float2 x = __half22float2(*reinterpret_cast<const __half2*>(&w));
  float q0 = __fmul_rn(x.x, r), q1 = __fmul_rn(x.y, r);
  q0 = __fmaf_rn(__fmaf_rn(-s, q0, x.x), r, q0);    // Markstein: RN(x / s)
  q1 = __fmaf_rn(__fmaf_rn(-s, q1, x.y), r, q1);
  uint32_t y = pack_f16x2(q1, q0);
  y = h2_min(h2_max(y, p.lo2), p.hi2);
  const uint32_t M = 0x66006600u;                   // (1536, 1536) = 1.5 * 2^10
  y = h2_sub(h2_add(y, M), M);
  return h2_mul(y, p.s2);
}

enum { CHAIN_F32 = 0, CHAIN_BF16_SMALL = 1, CHAIN_BF16_WIDE = 2, CHAIN_F16 = 3 };

// cheap bf16 qparams (quant.py:545-559 with T = math = qparam dtype = bf16); see the header for why FMUL suffices
__device__ __forceinline__ float rd_bf16(float x) { return __uint_as_float(pack_bf16x2(0.f, x) << 16); }

}  // namespace lcb
