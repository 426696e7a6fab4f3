// fq_group.cuh — the per_group / per_token / per_channel kernel for units of up to 32 lanes x 16 x 128-bit.
//
// A sub-warp of L lanes (4..32) owns one unit; every lane holds NV 128-bit vectors of it. A warp works on
// U * (32 / L) consecutive units per iteration (U * NV = 8..16 vectors per lane, all loads issued before any use, one
// contiguous span per warp). Per unit: packed (max, -min) reduce in registers, log2(L) xor-shuffles, the qparams
// (computed redundantly by every lane: cheaper than broadcasting them), then round + clamp + dequant on packed 16-bit
// pairs (fq_packed.cuh) and one 128-bit streaming store per vector. The loop body is a few hundred instructions:
// the previous revision unrolled 16 units x 3 code paths into 180 KB of straight-line SASS and ran at the speed of
// the instruction cache (profiles/r1_fakequant.md).
#pragma once
#include "fq_packed.cuh"

namespace lcb {

// ---- packed (max, -min) range type ---------------------------------------------------------------------
template <typename T>
struct Range;  // P: register type carrying (max, -min)

template <>
struct Range<float> {
  struct P { float mx, nmn; };
  __device__ __forceinline__ static P init() { return {-INFINITY, -INFINITY}; }
  __device__ __forceinline__ static P from_vec(const uint4& v) {
    float f[4];
    TI<float>::unpack(v, f);
    float a = max_nan(max_nan(f[0], f[1]), max_nan(f[2], f[3]));
    float b = min_nan(min_nan(f[0], f[1]), min_nan(f[2], f[3]));
    return {a, -b};
  }
  __device__ __forceinline__ static P join(P a, P b) { return {max_nan(a.mx, b.mx), max_nan(a.nmn, b.nmn)}; }
  __device__ __forceinline__ static P shfl_xor(P a, int o) {
    return {__shfl_xor_sync(0xffffffffu, a.mx, o), __shfl_xor_sync(0xffffffffu, a.nmn, o)};
  }
  __device__ __forceinline__ static float mx(P a) { return a.mx; }
  __device__ __forceinline__ static float mn(P a) { return -a.nmn; }
};

template <typename H2>
__device__ __forceinline__ H2 as_h2(uint32_t w) { return *reinterpret_cast<H2*>(&w); }
template <typename H2>
__device__ __forceinline__ uint32_t as_u32(H2 h) { return *reinterpret_cast<uint32_t*>(&h); }

template <typename T, typename H2>
struct Range16 {
  using P = uint32_t;  // low half: max, high half: -min
  __device__ __forceinline__ static P init() { return TI<T>::ID == LCB_BF16 ? 0xff80ff80u : 0xfc00fc00u; }
  __device__ __forceinline__ static P from_vec(const uint4& v) {
    H2 a = __hmax2_nan(__hmax2_nan(as_h2<H2>(v.x), as_h2<H2>(v.y)), __hmax2_nan(as_h2<H2>(v.z), as_h2<H2>(v.w)));
    H2 b = __hmin2_nan(__hmin2_nan(as_h2<H2>(v.x), as_h2<H2>(v.y)), __hmin2_nan(as_h2<H2>(v.z), as_h2<H2>(v.w)));
    const uint32_t ua = as_u32(a), ub = as_u32(b) ^ 0x80008000u;           // (a0,a1), (-b0,-b1)
    const uint32_t p = __byte_perm(ua, ub, 0x5410), q = __byte_perm(ua, ub, 0x7632);  // (a0,-b0), (a1,-b1)
    return as_u32(__hmax2_nan(as_h2<H2>(p), as_h2<H2>(q)));
  }
  __device__ __forceinline__ static P join(P a, P b) { return as_u32(__hmax2_nan(as_h2<H2>(a), as_h2<H2>(b))); }
  __device__ __forceinline__ static P shfl_xor(P a, int o) { return __shfl_xor_sync(0xffffffffu, a, o); }
  __device__ __forceinline__ static float mx(P a) {
    uint16_t h = (uint16_t)(a & 0xffffu);
    return TI<T>::load1(reinterpret_cast<const T*>(&h));
  }
  __device__ __forceinline__ static float mn(P a) {
    uint16_t h = (uint16_t)((a >> 16) ^ 0x8000u);
    return TI<T>::load1(reinterpret_cast<const T*>(&h));
  }
};
template <>
struct Range<__half> : Range16<__half, __half2> {};
template <>
struct Range<__nv_bfloat16> : Range16<__nv_bfloat16, __nv_bfloat162> {};

// Markstein step: RN(a / b) from r = RN(1 / b)
__device__ __forceinline__ float div_mk(float a, float b, float r) {
  float q0 = __fmul_rn(a, r);
  float e = __fmaf_rn(-b, q0, a);
  return __fmaf_rn(e, r, q0);
}

struct FastSpec {   // host-precomputed constants of the fast chains
  float qmin, qmax, div, rdiv;  // div = qmax (sym / fp8) or qmax - qmin (asym); rdiv = RN(1 / div)
  float eps;                    // rd_T(1e-5f)
  int chain;                    // CHAIN_*: element chain of the LCB_OUT_FAKE path
  int cheap_qp;                 // 1: bf16 qparams by FMUL + MUFU.RCP (fq_packed.cuh), 0: Markstein / IEEE reciprocal
};

// qparams of one unit from its range; math dtype == storage dtype T (quant.py:545-559)
template <typename T>
__device__ __forceinline__ void qparams_unit(float mn, float mx, const SpecD& sp, const FastSpec& fs, UnitQ& u) {
  float s, z = 0.f, r;   // sp.clip was consumed by fq_finish_unit (the vectors and the range arrive clamped)
  const bool symlike = sp.sym || sp.kind != LCB_KIND_INT;
  // rd(max(a, 1e-5f)) == max(a, rd(1e-5f)) for a already in T
  const float a = symlike ? max_nan(max_nan(fabsf(mx), fabsf(mn)), fs.eps) : max_nan(TI<T>::rd(__fsub_rn(mx, mn)), fs.eps);
  if (fs.cheap_qp) {
    s = TI<T>::rd(__fmul_rn(a, fs.rdiv));
    if (sp.kind != LCB_KIND_INT && s == 0.f && sp.out_mode != LCB_OUT_NONE) s = 1.f;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(s));  // s >= rd(1e-5)/div: never subnormal
    if (!symlike) {
      const float t = TI<T>::rd(__fmul_rn(mn, r));
      z = clamp_nan(TI<T>::rd(__fsub_rn(fs.qmin, rintf(t))), fs.qmin, fs.qmax);
    }
  } else {
    s = TI<T>::rd(div_mk(a, fs.div, fs.rdiv));
    if (sp.kind != LCB_KIND_INT && s == 0.f && sp.out_mode != LCB_OUT_NONE) s = 1.f;
    r = __frcp_rn(s);
    if (!symlike) {
      const float t = TI<T>::rd(div_mk(mn, s, r));
      z = clamp_nan(TI<T>::rd(__fsub_rn(fs.qmin, rintf(t))), fs.qmin, fs.qmax);
    }
  }
  u.s = s;
  u.r = r;
  u.z = z;
  u.s_out = (sp.osf != 1.0f) ? TI<T>::rd(__fmul_rn(s, sp.osf)) : s;
  u.lo = __fsub_rn(fs.qmin, z);
  u.hi = __fsub_rn(fs.qmax, z);
  // window in which every step above and in the element chains is exact (tests/test_fq_math_identity.py)
  const bool ok = sp.fast_ok && (s >= 1e-20f) && (s <= 1e20f) && finite_f(mx) && finite_f(mn) &&
                  (__fmul_rn(max_nan(fabsf(mx), fabsf(mn)), r) < 1e30f);
  u.exact = ok ? 0 : 1;
}

// ---- out-of-line rare paths (kept out of the loop body so that the hot code stays a few KB) -------------------------
// literal arithmetic for a whole vector (non-finite ranges, round_zp=False, exotic specs)
template <typename T>
__device__ __noinline__ void quant_store_vec_literal(uint4 raw, float mn, float mx, SpecD sp, void* out, int64_t e0) {
  constexpr int VEC = TI<T>::VEC;
  UnitQ u;
  qparams_from_range(mn, mx, sp, u);
  finish_unit(sp, u);
  float f[VEC], val[VEC], code[VEC];
  TI<T>::unpack(raw, f);
#pragma unroll 1
  for (int i = 0; i < VEC; ++i) {
    float2 vc = fq_elem_literal(f[i], u, sp);
    val[i] = vc.x;
    code[i] = vc.y;
  }
  store_vec<T>(out, e0, val, code, sp);
}
__device__ __noinline__ void write_qparams_literal(float mn, float mx, SpecD sp, void* scales, void* zeros, int64_t unit) {
  UnitQ u;
  qparams_from_range(mn, mx, sp, u);
  finish_unit(sp, u);
  write_qparams(scales, zeros, unit, u, sp);
}
// fp32 fast chain + runtime out_mode store (codes / fp8 / wide integer ranges)
template <typename T, int KINDC>
__device__ __noinline__ void quant_store_vec_generic(uint4 raw, UnitQ u, SpecD sp, void* out, int64_t e0) {
  float f[TI<T>::VEC];
  TI<T>::unpack(raw, f);
  quant_store_vec<T, T, KINDC>(f, u, sp, out, e0);
}

// FP8 fake-quant of a packed pair of 16-bit values (FloatQuantizer per_group / per_token activations, quant.py:1061-1081):
// q = rd_T(x / s) (+ 0: the reference's `+ zeros` turns -0 into +0), fp8 RNE with saturation, decode, * s, rd_T.
// Same operations as fq_elem_fast_fp8, two elements per conversion instruction: cvt.rn.{bf16x2,f16x2}.f32 packs the pair,
// one packed add applies the +0, cvt.rn.satfinite.e4m3x2 / e5m2x2 encodes both, cvt.rn.f16x2.e4m3x2 decodes both.
template <typename T>
__device__ __forceinline__ uint32_t fq2_fp8_fake(uint32_t w, const UnitQ& u, bool e5) {
  float x0, x1;
  if (TI<T>::ID == LCB_BF16) {
    x0 = __uint_as_float(w << 16);
    x1 = __uint_as_float(w & 0xffff0000u);
  } else {
    const float2 t = __half22float2(*reinterpret_cast<const __half2*>(&w));
    x0 = t.x;
    x1 = t.y;
  }
  float q0 = __fmul_rn(x0, u.r), q1 = __fmul_rn(x1, u.r);
  q0 = __fmaf_rn(__fmaf_rn(-u.s, q0, x0), u.r, q0);   // == RN(x / s) (Markstein step on the correctly rounded reciprocal)
  q1 = __fmaf_rn(__fmaf_rn(-u.s, q1, x1), u.r, q1);
  uint16_t c;
  if (TI<T>::ID == LCB_BF16) {
    __nv_bfloat162 p = __floats2bfloat162_rn(q0, q1);
    p = __hadd2(p, __floats2bfloat162_rn(0.f, 0.f));
    const uint32_t pw = *reinterpret_cast<const uint32_t*>(&p);
    const float r0 = __uint_as_float(pw << 16), r1 = __uint_as_float(pw & 0xffff0000u);
    if (e5) asm("cvt.rn.satfinite.e5m2x2.f32 %0, %1, %2;" : "=h"(c) : "f"(r1), "f"(r0));
    else asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(c) : "f"(r1), "f"(r0));
  } else {
    __half2 p = __floats2half2_rn(q0, q1);
    p = __hadd2(p, __floats2half2_rn(0.f, 0.f));
    const uint32_t pw = *reinterpret_cast<const uint32_t*>(&p);
    if (e5) asm("cvt.rn.satfinite.e5m2x2.f16x2 %0, %1;" : "=h"(c) : "r"(pw));
    else asm("cvt.rn.satfinite.e4m3x2.f16x2 %0, %1;" : "=h"(c) : "r"(pw));
  }
  uint32_t h;
  if (e5) asm("cvt.rn.f16x2.e5m2x2 %0, %1;" : "=r"(h) : "h"(c));
  else asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(h) : "h"(c));
  const float2 v = __half22float2(*reinterpret_cast<const __half2*>(&h));
  const float o0 = __fmul_rn(v.x, u.s_out), o1 = __fmul_rn(v.y, u.s_out);
  if (TI<T>::ID == LCB_BF16) {
    const __nv_bfloat162 o = __floats2bfloat162_rn(o0, o1);
    return *reinterpret_cast<const uint32_t*>(&o);
  }
  const __half2 o = __floats2half2_rn(o0, o1);
  return *reinterpret_cast<const uint32_t*>(&o);
}

// fp32 fast chain, LCB_OUT_FAKE only, inline (fp32 tensors; fp8 fake-quant)
template <typename T, int KINDC>
__device__ __forceinline__ uint4 quant_vec_fast_f32(const uint4& raw, const UnitQ& u, int kind) {
  constexpr int VEC = TI<T>::VEC;
  if (KINDC == 1 && sizeof(T) == 2) {
    const bool e5 = kind == LCB_KIND_FP8_E5M2;
    return make_uint4(fq2_fp8_fake<T>(raw.x, u, e5), fq2_fp8_fake<T>(raw.y, u, e5), fq2_fp8_fake<T>(raw.z, u, e5),
                      fq2_fp8_fake<T>(raw.w, u, e5));
  }
  float f[VEC], val[VEC], code;
  TI<T>::unpack(raw, f);
#pragma unroll
  for (int i = 0; i < VEC; ++i)
    val[i] = KINDC == 0 ? fq_elem_fast_int<T>(f[i], u, &code) : fq_elem_fast_fp8<T>(f[i], u, kind, &code);
  return TI<T>::pack(val);
}

template <typename T>
__device__ __noinline__ uint4 clamp_vec_noinline(uint4 raw, float lo, float hi) {
  float f[TI<T>::VEC];
  TI<T>::unpack(raw, f);
#pragma unroll
  for (int i = 0; i < TI<T>::VEC; ++i) f[i] = clamp_nan(f[i], lo, hi);   // exact: values and bounds are T-representable
  return TI<T>::pack(f);
}

enum { CHAIN_ANY_OUT = -1 };  // template value: runtime out_mode, out-of-line generic chain

// CHAIN: element chain compiled in. CHAIN_ANY_OUT serves every out_mode through quant_store_vec_generic; the others
// are LCB_OUT_FAKE only (the host guarantees it).
// qparams -> (rare) literal path or the compiled-in element chain -> stores, for the NV vectors one lane holds of a unit.
// Lane's vector j sits at vector index v0 + j * vstride of the unit and at element offset e0 + j * vstride * VEC.
template <typename T, int NV, int KINDC, int CHAIN>
__device__ __forceinline__ void fq_finish_unit(const uint4 (&raw_in)[NV], float mn, float mx, int64_t unit, int64_t e0, int v0,
                                               int vstride, int nvec, bool writer, void* __restrict__ out,
                                               void* __restrict__ scales, void* __restrict__ zeros, const SpecD& sp_in,
                                               const FastSpec& fs) {
  constexpr int VEC = TI<T>::VEC;
  uint4 raw[NV];
#pragma unroll
  for (int j = 0; j < NV; ++j) raw[j] = raw_in[j];
  const SpecD& sp = sp_in;
  float mnc = mn, mxc = mx;
  {  // AutoClip v1 (rare, uniform): clamp the unit to its shrunk range; the fast chains then see the clamped tensor.
     // The literal path below gets the original vectors and range (qparams_from_range / fq_elem_literal clip themselves).
    float lo, hi;
    if (clip_bounds(mnc, mxc, sp.clip, TI<T>::ID, lo, hi)) {
#pragma unroll
      for (int j = 0; j < NV; ++j) raw[j] = clamp_vec_noinline<T>(raw[j], lo, hi);
    }
  }
  UnitQ u;
  qparams_unit<T>(mnc, mxc, sp, fs, u);
  if (u.exact) {  // unit-uniform, rare: literal chain, qparams recomputed literally as well
    if (writer) write_qparams_literal(mn, mx, sp, scales, zeros, unit);
    if (sp.out_mode != LCB_OUT_NONE) {
#pragma unroll
      for (int j = 0; j < NV; ++j)
        if (v0 + j * vstride < nvec) quant_store_vec_literal<T>(raw_in[j], mn, mx, sp, out, e0 + (int64_t)j * vstride * VEC);
    }
    return;
  }
  if (writer) write_qparams(scales, zeros, unit, u, sp);
  if (CHAIN == CHAIN_ANY_OUT) {
    if (sp.out_mode != LCB_OUT_NONE) {
#pragma unroll
      for (int j = 0; j < NV; ++j)
        if (v0 + j * vstride < nvec)
          quant_store_vec_generic<T, KINDC>(raw[j], u, sp, out, e0 + (int64_t)j * vstride * VEC);
    }
    return;
  }
  T* o = reinterpret_cast<T*>(out) + e0;
  UnitP p;
  if (CHAIN == CHAIN_BF16_SMALL || CHAIN == CHAIN_BF16_WIDE) {
    p.lo2 = pack_bf16x2(u.lo, u.lo);
    p.hi2 = pack_bf16x2(u.hi, u.hi);
    p.s2 = pack_bf16x2(u.s_out, u.s_out);
  } else if (CHAIN == CHAIN_F16) {
    p.lo2 = pack_f16x2(u.lo, u.lo);
    p.hi2 = pack_f16x2(u.hi, u.hi);
    p.s2 = pack_f16x2(u.s_out, u.s_out);
  }
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const uint4 w = raw[j];
    uint4 r4;
    if (CHAIN == CHAIN_BF16_SMALL)
      r4 = make_uint4(fq2_bf16_small(w.x, u.r, p), fq2_bf16_small(w.y, u.r, p), fq2_bf16_small(w.z, u.r, p),
                      fq2_bf16_small(w.w, u.r, p));
    else if (CHAIN == CHAIN_BF16_WIDE)
      r4 = make_uint4(fq2_bf16_wide(w.x, u.r, p, u.s_out), fq2_bf16_wide(w.y, u.r, p, u.s_out),
                      fq2_bf16_wide(w.z, u.r, p, u.s_out), fq2_bf16_wide(w.w, u.r, p, u.s_out));
    else if (CHAIN == CHAIN_F16)
      r4 = make_uint4(fq2_f16(w.x, u.s, u.r, p), fq2_f16(w.y, u.s, u.r, p), fq2_f16(w.z, u.s, u.r, p),
                      fq2_f16(w.w, u.s, u.r, p));
    else
      r4 = quant_vec_fast_f32<T, KINDC>(w, u, sp.kind);
    if (v0 + j * vstride < nvec) st_stream(o + (int64_t)j * vstride * VEC, r4);
  }
}

// One warp-iteration: the warp's chunk of U * (32 / lanes) consecutive units of the tensor (in, out, n_units).
template <typename T, int NV, int KINDC, int CHAIN>
__device__ __forceinline__ void fq_group_chunk(const T* __restrict__ in, void* __restrict__ out, void* __restrict__ scales,
                                               void* __restrict__ zeros, int64_t n_units, int unit_len, int lanes,
                                               int64_t chunk, const SpecD& sp, const FastSpec& fs) {
  constexpr int VEC = TI<T>::VEC;
  constexpr int U = NV >= 8 ? 1 : 8 / NV;
  using R = Range<T>;
  const int nvec = unit_len / VEC;
  const int lane = threadIdx.x & 31;
  const int lane_in = lane & (lanes - 1);
  const int spw = 32 / lanes;          // sub-warps (units) per warp per k
  const int sub_in_warp = lane / lanes;
  const int64_t upw = (int64_t)U * spw;  // units per warp per iteration
  const int64_t unit0 = chunk * upw + sub_in_warp;  // unit of k: unit0 + k * spw
  uint4 raw[U][NV];
#pragma unroll
  for (int k = 0; k < U; ++k) {
    const int64_t unit = unit0 + (int64_t)k * spw;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int v = j * lanes + lane_in;
      raw[k][j] = (unit < n_units && v < nvec) ? ld_stream(in + unit * unit_len + (int64_t)v * VEC)
                                               : make_uint4(0u, 0u, 0u, 0u);
    }
  }
#pragma unroll
  for (int k = 0; k < U; ++k) {
    const int64_t unit = unit0 + (int64_t)k * spw;
    typename R::P rg = R::init();
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int v = j * lanes + lane_in;
      if (v < nvec) rg = R::join(rg, R::from_vec(raw[k][j]));
    }
    for (int o = lanes >> 1; o > 0; o >>= 1) rg = R::join(rg, R::shfl_xor(rg, o));
    const float mx = R::mx(rg), mn = R::mn(rg);
    if (unit >= n_units) continue;  // sub-warp uniform; no shuffles below
    fq_finish_unit<T, NV, KINDC, CHAIN>(raw[k], mn, mx, unit, unit * unit_len + (int64_t)lane_in * VEC, lane_in, lanes, nvec,
                                        lane_in == 0, out, scales, zeros, sp, fs);
  }
}

// Programmatic dependent launch: the kernel lets its successor in the stream be scheduled right away
// (launch_dependents) and waits for its predecessor to COMPLETE and flush (wait) before touching memory — ordinary stream
// semantics, but the successor's launch latency and CTA rasterisation overlap this kernel's tail. Back-to-back quantizer
// calls (one per linear at deploy time, one per linear per forward for activations) are a few microseconds each.
__device__ __forceinline__ void pdl_enter() {
  asm volatile("griddepcontrol.launch_dependents;");
  asm volatile("griddepcontrol.wait;" ::: "memory");
}

template <typename T, int NV, int KINDC, int CHAIN>
__global__ void __launch_bounds__(256, 4)
    fq_group_kernel(const T* __restrict__ in, void* __restrict__ out, void* __restrict__ scales,
                    void* __restrict__ zeros, int64_t n_units, int unit_len, int lanes, SpecD sp, FastSpec fs) {
  constexpr int U = NV >= 8 ? 1 : 8 / NV;
  pdl_enter();
  const int64_t warp_id = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const int64_t upw = (int64_t)U * (32 / lanes);
  const int64_t chunks = (n_units + upw - 1) / upw;
  for (int64_t c = warp_id; c < chunks; c += n_warps)
    fq_group_chunk<T, NV, KINDC, CHAIN>(in, out, scales, zeros, n_units, unit_len, lanes, c, sp, fs);
}

// Several tensors in ONE launch (same dtype / spec / unit length): e.g. the 7 linear weights of a transformer block.
// Chunks of all tensors are laid end to end; a warp finds its tensor with a warp-uniform scan of the prefix table.
constexpr int kMaxMulti = 16;
struct MultiDesc {
  const void* in[kMaxMulti];
  void* out[kMaxMulti];
  int64_t n_units[kMaxMulti];
  int64_t chunk_start[kMaxMulti + 1];
  int n;
};

template <typename T, int NV, int KINDC, int CHAIN>
__global__ void __launch_bounds__(256, 4)
    fq_group_multi_kernel(const __grid_constant__ MultiDesc md, int unit_len, int lanes, SpecD sp, FastSpec fs) {
  pdl_enter();
  const int64_t warp_id = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const int64_t chunks = md.chunk_start[md.n];
  int t = 0;
  for (int64_t c = warp_id; c < chunks; c += n_warps) {
    while (c >= md.chunk_start[t + 1]) ++t;  // warp-uniform, monotone in c
    fq_group_chunk<T, NV, KINDC, CHAIN>(reinterpret_cast<const T*>(md.in[t]), md.out[t], nullptr, nullptr, md.n_units[t],
                                        unit_len, lanes, c - md.chunk_start[t], sp, fs);
  }
}

// Rows longer than one warp's registers (2048 < elements <= 16384 for 16-bit types): W warps share a row, NV = 8 vectors
// per lane, cross-warp range through shared memory and ONE named barrier per row (slots double-buffered by iteration).
// CTA = 8 warps = 8 / W rows at a time. Same element chains as the sub-warp kernel.
template <typename T, int KINDC, int CHAIN, int W, int NV = 8>
__global__ void __launch_bounds__(256, W == 8 ? 3 : 2)   // 8 warps per row = one row per CTA: a third CTA per SM is a third
                                                         // row in flight (+18 %); 2 / 4 warps per row lose 3 % to spills at 3
    fq_wrow_kernel(const T* __restrict__ in, void* __restrict__ out, void* __restrict__ scales, void* __restrict__ zeros,
                   int64_t n_units, int unit_len, SpecD sp, FastSpec fs) {
  constexpr int VEC = TI<T>::VEC;
  constexpr int RPC = 8 / W;  // rows per CTA per iteration
  using R = Range<T>;
  __shared__ typename R::P slot[2][RPC][W];
  pdl_enter();
  const int nvec = unit_len / VEC;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int rsub = warp / W, wsub = warp % W;
  const int v0 = wsub * 32 + lane;
  const int64_t n_iter = (n_units + (int64_t)gridDim.x * RPC - 1) / ((int64_t)gridDim.x * RPC);
  for (int64_t it = 0; it < n_iter; ++it) {
    const int64_t unit = (it * gridDim.x + blockIdx.x) * RPC + rsub;
    const bool valid = unit < n_units;
    uint4 raw[NV];
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int v = v0 + j * W * 32;
      raw[j] = (valid && v < nvec) ? ld_stream(in + unit * unit_len + (int64_t)v * VEC) : make_uint4(0u, 0u, 0u, 0u);
    }
    typename R::P rg = R::init();
#pragma unroll
    for (int j = 0; j < NV; ++j)
      if (v0 + j * W * 32 < nvec) rg = R::join(rg, R::from_vec(raw[j]));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) rg = R::join(rg, R::shfl_xor(rg, o));
    const int par = (int)(it & 1);
    if (lane == 0) slot[par][rsub][wsub] = rg;
    asm volatile("bar.sync %0, %1;" ::"r"(rsub + 1), "r"(W * 32) : "memory");
    rg = slot[par][rsub][0];
#pragma unroll
    for (int w = 1; w < W; ++w) rg = R::join(rg, slot[par][rsub][w]);
    if (!valid) continue;  // row-group uniform
    fq_finish_unit<T, NV, KINDC, CHAIN>(raw, R::mn(rg), R::mx(rg), unit, unit * unit_len + (int64_t)v0 * VEC, v0, W * 32, nvec,
                                        wsub == 0 && lane == 0, out, scales, zeros, sp, fs);
  }
}

}  // namespace lcb
