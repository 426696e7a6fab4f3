// fq_group.cuh — the per_group / short-row kernel (units of <= 2048 16-bit or 1024 fp32 elements).
//
// A sub-warp of L lanes (4..32) owns U units per iteration (U*NV <= 16 x 128-bit vectors per lane, all loads issued
// before any use). Per unit every lane reduces its 8*NV elements to a packed (max, -min) pair; a TRANSPOSING xor
// butterfly then leaves lane i with the range of unit i (U-1 shuffles for U units instead of U*log2 L), lane i
// computes that unit's qparams ONCE (division-free: Markstein steps on host-rounded reciprocals), and the four
// per-unit constants are broadcast back with sub-warp shuffles. Element work is the fast chain of fq_math.cuh.
#pragma once
#include "fq_math.cuh"

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
  __device__ __forceinline__ static P select(bool c, P a, P b) { return {c ? a.mx : b.mx, c ? a.nmn : b.nmn}; }
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
  __device__ __forceinline__ static P init() { return T16_NEG_INF2(); }
  __device__ __forceinline__ static uint32_t T16_NEG_INF2() {
    return sizeof(T) == 2 && TI<T>::ID == LCB_BF16 ? 0xff80ff80u : 0xfc00fc00u;
  }
  __device__ __forceinline__ static P from_vec(const uint4& v) {
    H2 a = __hmax2_nan(__hmax2_nan(as_h2<H2>(v.x), as_h2<H2>(v.y)), __hmax2_nan(as_h2<H2>(v.z), as_h2<H2>(v.w)));
    H2 b = __hmin2_nan(__hmin2_nan(as_h2<H2>(v.x), as_h2<H2>(v.y)), __hmin2_nan(as_h2<H2>(v.z), as_h2<H2>(v.w)));
    const uint32_t ua = as_u32(a), ub = as_u32(b) ^ 0x80008000u;           // (a0,a1), (-b0,-b1)
    const uint32_t p = __byte_perm(ua, ub, 0x5410), q = __byte_perm(ua, ub, 0x7632);  // (a0,-b0), (a1,-b1)
    return as_u32(__hmax2_nan(as_h2<H2>(p), as_h2<H2>(q)));
  }
  __device__ __forceinline__ static P join(P a, P b) { return as_u32(__hmax2_nan(as_h2<H2>(a), as_h2<H2>(b))); }
  __device__ __forceinline__ static P shfl_xor(P a, int o) { return __shfl_xor_sync(0xffffffffu, a, o); }
  __device__ __forceinline__ static P select(bool c, P a, P b) { return c ? a : b; }
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

// ---- division-free qparams (math dtype == storage dtype T) -------------------------------------------------
// Markstein step: RN(a / b) from r = RN(1 / b)
__device__ __forceinline__ float div_mk(float a, float b, float r) {
  float q0 = __fmul_rn(a, r);
  float e = __fmaf_rn(-b, q0, a);
  return __fmaf_rn(e, r, q0);
}

struct FastSpec {   // host-precomputed constants of the fast chain
  float qmin, qmax, div, rdiv;  // div = qmax (sym / fp8) or qmax - qmin (asym); rdiv = RN(1 / div)
  float eps;                    // rd_T(1e-5f)
};

template <typename T>
__device__ __forceinline__ void qparams_fast(float mn, float mx, const SpecD& sp, const FastSpec& fs, UnitQ& u) {
  if (sp.clip != 1.0f) {
    mx = TI<T>::rd(__fmul_rn(mx, sp.clip));
    mn = TI<T>::rd(__fmul_rn(mn, sp.clip));
  }
  float s, z = 0.f;
  if (sp.sym || sp.kind != LCB_KIND_INT) {
    float a = max_nan(max_nan(fabsf(mx), fabsf(mn)), fs.eps);  // rd(max(a, 1e-5f)) == max(a, rd(1e-5f)) for a in T
    s = TI<T>::rd(div_mk(a, fs.div, fs.rdiv));
    if (sp.kind != LCB_KIND_INT && s == 0.f && sp.out_mode != LCB_OUT_NONE) s = 1.f;
    u.r = __frcp_rn(s);
  } else {
    float d = max_nan(TI<T>::rd(__fsub_rn(mx, mn)), fs.eps);
    s = TI<T>::rd(div_mk(d, fs.div, fs.rdiv));
    u.r = __frcp_rn(s);
    float t = TI<T>::rd(div_mk(mn, s, u.r));
    z = clamp_nan(TI<T>::rd(__fsub_rn(fs.qmin, rintf(t))), fs.qmin, fs.qmax);
  }
  u.s = s;
  u.z = z;
  u.s_out = (sp.osf != 1.0f) ? TI<T>::rd(__fmul_rn(s, sp.osf)) : s;
  u.lo = __fsub_rn(fs.qmin, z);
  u.hi = __fsub_rn(fs.qmax, z);
  // window in which every Markstein step above and in the element chain is exact (tests/test_fq_math_identity.py)
  const bool ok = sp.fast_ok && (s >= 1e-20f) && (s <= 1e20f) && finite_f(mx) && finite_f(mn);
  u.exact = ok ? 0 : 1;
}

template <typename T, int NV, int U, int KINDC>
__global__ void __launch_bounds__(256)
    fq_group_kernel(const T* __restrict__ in, void* __restrict__ out, void* __restrict__ scales,
                    void* __restrict__ zeros, int64_t n_units, int unit_len, int lanes, SpecD sp, FastSpec fs) {
  constexpr int VEC = TI<T>::VEC;
  using R = Range<T>;
  const int nvec = unit_len / VEC;
  const int lane = threadIdx.x & 31;
  const int lane_in = lane & (lanes - 1);
  const unsigned sub_mask = (lanes == 32 ? 0xffffffffu : ((1u << lanes) - 1u)) << (lane & ~(lanes - 1));
  const int64_t sub_id = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / lanes;
  const int64_t n_sub = ((int64_t)gridDim.x * blockDim.x) / lanes;
  const int64_t iters = (n_units + n_sub * U - 1) / (n_sub * U);

  for (int64_t it = 0; it < iters; ++it) {
    // consecutive sub-warps take consecutive blocks of U units: a warp reads one contiguous span per k
    const int64_t unit0 = (it * n_sub + sub_id) * U;
    uint4 raw[U][NV];
#pragma unroll
    for (int k = 0; k < U; ++k) {
#pragma unroll
      for (int j = 0; j < NV; ++j) {
        const int v = j * lanes + lane_in;
        if (unit0 + k < n_units && v < nvec) raw[k][j] = ld_stream(in + (unit0 + k) * unit_len + (int64_t)v * VEC);
      }
    }
    typename R::P rg[U];
#pragma unroll
    for (int k = 0; k < U; ++k) {
      rg[k] = R::init();
#pragma unroll
      for (int j = 0; j < NV; ++j) {
        const int v = j * lanes + lane_in;
        if (unit0 + k < n_units && v < nvec) rg[k] = R::join(rg[k], R::from_vec(raw[k][j]));
      }
    }
    // plain butterfly while more lanes than units, then the transposing butterfly
    for (int o = lanes >> 1; o >= U; o >>= 1) {
#pragma unroll
      for (int k = 0; k < U; ++k) rg[k] = R::join(rg[k], R::shfl_xor(rg[k], o));
    }
#pragma unroll
    for (int o = U / 2; o >= 1; o >>= 1) {
      const bool upper = (lane_in & o) != 0;
#pragma unroll
      for (int i = 0; i < o; ++i) {
        typename R::P send = R::select(upper, rg[i], rg[i + o]);
        typename R::P keep = R::select(upper, rg[i + o], rg[i]);
        rg[i] = R::join(keep, R::shfl_xor(send, o));
      }
    }
    // lane_in (mod U) now owns unit (lane_in mod U)
    const int own = lane_in & (U - 1);
    UnitQ mine;
    qparams_fast<T>(R::mn(rg[0]), R::mx(rg[0]), sp, fs, mine);
    if (lane_in < U && unit0 + own < n_units) write_qparams(scales, zeros, unit0 + own, mine, sp);
    if (sp.out_mode == LCB_OUT_NONE) continue;
    const float r_flag = mine.exact ? -1.0f : mine.r;
#pragma unroll
    for (int k = 0; k < U; ++k) {
      UnitQ u;
      u.s = __shfl_sync(0xffffffffu, mine.s, k, lanes);
      u.r = __shfl_sync(0xffffffffu, r_flag, k, lanes);
      u.lo = __shfl_sync(0xffffffffu, mine.lo, k, lanes);
      u.hi = __shfl_sync(0xffffffffu, mine.hi, k, lanes);
      u.s_out = (sp.osf != 1.0f) ? __shfl_sync(0xffffffffu, mine.s_out, k, lanes) : u.s;
      u.z = __fsub_rn(fs.qmin, u.lo);
      u.exact = 0;
      if (unit0 + k >= n_units) continue;  // sub-warp uniform
      if (u.r < 0.f) {                     // sub-warp uniform: literal chain for this unit
        const float mx = __shfl_sync(sub_mask, R::mx(rg[0]), k, lanes);
        const float mn = __shfl_sync(sub_mask, R::mn(rg[0]), k, lanes);
        qparams_from_range(mn, mx, sp, u);
        finish_unit(sp, u);
        u.exact = 1;
      }
#pragma unroll
      for (int j = 0; j < NV; ++j) {
        const int v = j * lanes + lane_in;
        if (v < nvec) {
          float f[VEC];
          TI<T>::unpack(raw[k][j], f);
          quant_store_vec<T, T, KINDC>(f, u, sp, out, (unit0 + k) * unit_len + (int64_t)v * VEC);
        }
      }
    }
  }
}

}  // namespace lcb
