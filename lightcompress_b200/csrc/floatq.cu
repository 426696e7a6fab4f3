// floatq.cu — FloatQuantizer WITHOUT qtorch (`use_qtorch` unset; configs/quantization/methods/FP_Quant/rtn_we4m3ae4m3.yml,
// rtn_we5m2ae5m2.yml): every ELEMENT gets its own scale 2^(floor(log2|x| + bias) - m - bias), the unit (row / group /
// token) only contributes its absolute maximum through `bias`. Reference: FloatQuantizer.get_float_qparams
// (quant.py:1005-1027), quant / dequant (quant.py:1061-1076) — a chain of torch ops on CUDA tensors, one rounding to the
// tensor dtype after every op (e_bits < 5) or an fp32 chain with one final cast (e_bits >= 5: `maxval.to(float32)`
// promotes everything downstream, quant.py:1016-1017).
//
//   maxval = max(max_val, -min_val)                                   per unit
//   bias   = 2^e - log2(maxval) + log2(2 - 2^-m) - 1                   per unit, three roundings
//   xc     = min(max(x, -maxval), maxval)
//   ls     = clamp(floor(log2|xc| + bias), min = 1)
//   s      = 2.0 ** (ls - m - bias);  s[s == 0] = 1
//   y      = (round(xc / s + 0) - 0) * s
//
// The 0-dim CUDA tensors `e_bits` / `m_bits` are operands like any other: TensorIterator casts them to the common dtype
// of the op (the tensor dtype) when they are loaded, so log2(2 - 2^-m) enters the bf16 chain as a bf16 number. Python
// scalars (1, 2.0) enter as fp32. log2 / pow are libdevice's log2f / powf on the fp32 image of the operand, as in ATen's
// CUDA kernels (compiled here without fast-math, so the same routines). An all-zero unit gives log2(0) + inf = NaN and the
// reference returns NaN for the whole unit; so does this kernel.
//
// One pass over HBM: a unit is read once into registers / re-read from L1 for the element chain. The chain holds a
// division, a log2f and a powf per element (~150 instructions): the kernel is bound by the FP32 / SFU pipes, not by HBM
// (measured number in DESIGN.md), which is what the arithmetic of the reference dictates.
#include <algorithm>
#include <cstdlib>
#include <type_traits>
#include "common.cuh"

namespace lcb {
namespace {

template <typename T, bool F32CHAIN>
struct Chain {
  // rounding after every op: identity on the fp32 chain
  __device__ __forceinline__ static float rd(float v) { return F32CHAIN ? v : TI<T>::rd(v); }

  // bias of a unit from its maxval (exact in T / fp32)
  __device__ __forceinline__ static float bias_of(float maxval, float two_e, float c_log) {
    float b = rd(__fsub_rn(two_e, rd(log2f(maxval))));
    b = rd(__fadd_rn(b, c_log));
    return rd(__fsub_rn(b, 1.0f));
  }

  __device__ __forceinline__ static float clamp_unit(float x, float maxval) {
    return min_nan(max_nan(x, -maxval), maxval);
  }

  // element scale from the clamped element
  __device__ __forceinline__ static float scale_of(float xc, float bias, float m_bits) {
    float t = rd(__fadd_rn(rd(log2f(fabsf(xc))), bias));
    t = floorf(t);
    t = (t != t) ? t : fmaxf(t, 1.0f);          // torch.clamp(min=1.0) keeps NaN
    float e = rd(__fsub_rn(rd(__fsub_rn(t, m_bits)), bias));
    return rd(powf(2.0f, e));
  }

  // (round(x / s + 0) - 0) * s with s == 0 replaced by 1 (quant.py:1061-1076)
  __device__ __forceinline__ static float qdq(float x, float s) {
    if (s == 0.0f) s = 1.0f;
    float q = rd(__fadd_rn(rd(__fdiv_rn(x, s)), 0.0f));
    q = nearbyintf(q);
    return rd(__fmul_rn(rd(__fsub_rn(q, 0.0f)), s));
  }
};

struct FParams {
  const void* in;
  void* out;        // T
  void* xc_out;     // T (e < 5) or fp32 (e >= 5), may be null
  void* s_out;      // same dtype rule, may be null
  int64_t n_units, unit;
  float e_bits, m_bits;
};

template <bool F32CHAIN, typename T>
__device__ __forceinline__ void store_aux(void* base, int64_t i, float v) {
  if (F32CHAIN) reinterpret_cast<float*>(base)[i] = v;
  else TI<T>::store1(reinterpret_cast<T*>(base) + i, v);
}

// LANES threads per unit (a power of two <= 32, or a whole CTA when LANES == 0)
template <typename T, bool F32CHAIN, int LANES>
__global__ void __launch_bounds__(256) floatq_dynamic_kernel(FParams p) {
  using C = Chain<T, F32CHAIN>;
  const T* in = reinterpret_cast<const T*>(p.in);
  T* out = reinterpret_cast<T*>(p.out);
  constexpr int TPU = LANES == 0 ? 256 : LANES;   // threads per unit
  const int upb = 256 / TPU;                      // units per CTA pass
  const int sub = threadIdx.x / TPU, lane = threadIdx.x % TPU;
  __shared__ float red_mx[8], red_mn[8];
  // the two 0-dim operands, computed with the device's own routines like the reference's 0-dim CUDA tensors:
  // 2 ** e_bits and log2(2 - 2 ** (-m_bits)) (fp32), the latter rounded to the chain's dtype when the consuming op loads it
  const float two_e = powf(2.0f, p.e_bits);
  const float c_log = C::rd(log2f(__fsub_rn(2.0f, powf(2.0f, -p.m_bits))));
  for (int64_t u0 = (int64_t)blockIdx.x * upb; u0 < p.n_units; u0 += (int64_t)gridDim.x * upb) {
    const int64_t u = u0 + sub;
    const bool live = u < p.n_units;
    const T* row = in + u * p.unit;
    float mx = -INFINITY, mn = INFINITY;
    if (live) {
      for (int64_t i = lane; i < p.unit; i += TPU) {
        const float x = TI<T>::load1(row + i);
        mx = max_nan(mx, x);
        mn = min_nan(mn, x);
      }
    }
    if (LANES != 0) {
#pragma unroll
      for (int o = TPU / 2; o > 0; o >>= 1) {
        mx = max_nan(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        mn = min_nan(mn, __shfl_xor_sync(0xffffffffu, mn, o));
      }
    } else {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        mx = max_nan(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        mn = min_nan(mn, __shfl_xor_sync(0xffffffffu, mn, o));
      }
      __syncthreads();   // the previous unit's readers are done
      if ((threadIdx.x & 31) == 0) {
        red_mx[threadIdx.x >> 5] = mx;
        red_mn[threadIdx.x >> 5] = mn;
      }
      __syncthreads();
      mx = red_mx[0];
      mn = red_mn[0];
#pragma unroll
      for (int w = 1; w < 8; ++w) {
        mx = max_nan(mx, red_mx[w]);
        mn = min_nan(mn, red_mn[w]);
      }
    }
    if (!live) continue;
    const float maxval = max_nan(mx, -mn);
    const float bias = C::bias_of(maxval, two_e, c_log);
    for (int64_t i = lane; i < p.unit; i += TPU) {
      const float xc = C::clamp_unit(TI<T>::load1(row + i), maxval);
      const float s = C::scale_of(xc, bias, p.m_bits);
      const int64_t e = u * p.unit + i;
      if (p.xc_out) store_aux<F32CHAIN, T>(p.xc_out, e, xc);
      if (p.s_out) store_aux<F32CHAIN, T>(p.s_out, e, s);
      if (out) TI<T>::store1(out + e, C::qdq(xc, s));    // fp32 chain: the final .to(org_dtype)
    }
  }
}

// static: y = (round(x / s + 0) - 0) * s with per-element scales; math dtype = promote(T, scales dtype)
template <typename T, typename TS, bool F32CHAIN>
__global__ void __launch_bounds__(256) floatq_static_kernel(const T* __restrict__ in, const TS* __restrict__ scales,
                                                           T* __restrict__ out, int64_t n) {
  using C = Chain<T, F32CHAIN>;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    TI<T>::store1(out + i, C::qdq(TI<T>::load1(in + i), TI<TS>::load1(scales + i)));
}

// ---- fast kernel: 16-bit tensors, 16-byte aligned units of <= 16384 elements, fake-quant output only ----------------------
// Same arithmetic, fewer instructions. The element scale is a function of the unit's bias and of the INTEGER
// ls = clamp(floor(log2|xc| + bias), 1) only, and ls <= K := ls(maxval) <= 2^e + 1: the threads of a unit build a table
// (s_k, RN(1 / s_k)) for k = 1 .. K once per unit (K powf calls per unit instead of one per element) and the element
// divides by Markstein's correction on the tabulated reciprocal (q0 = x r; e = fma(-s, q0, x); q = fma(e, r, q0) == RN(x / s)
// for s in [1e-20, 1e20] and |x| in [2^-40, 1e8): tests/test_fq_math_identity.py). Whatever falls outside — a table index
// that is not in 1 .. K (NaN, or log2f not monotone in its last bit), a scale or an element outside the window — takes the
// literal chain for that element. The unit lives in registers between the range pass and the element pass.
constexpr int kTabMax = 48;

// floor(t) for |t| < 2^22 without the conversion pipe: rint by the 1.5 * 2^23 magic add, minus one where it rounded up
__device__ __forceinline__ float floor_small(float t) {
  const float r = __fsub_rn(__fadd_rn(t, 12582912.0f), 12582912.0f);
  return r > t ? __fsub_rn(r, 1.0f) : r;
}

// One element on the fast path. Returns the fake-quantized value and sets `ok` when it is certified to be the value of the
// literal chain; otherwise the caller recomputes the element with elem_slow. Branch-free: eight of these interleave.
//
// ls = clamp(floor(rd(rd(log2f(ax)) + bias)), 1) comes from MUFU.LG2 when the answer cannot depend on the last bits of the
// logarithm: lg2.approx is within 2^-22 (absolute for results in (-1, 1), relative otherwise) of log2 and libdevice's log2f
// within one ulp; both ends of an interval four times that wide must round to the same number (16-bit chain: the first
// rounding is the only place the logarithm's low bits can matter; fp32 chain: the floor). Not certified: zero, subnormal,
// inf, NaN, |x| outside [2^-40, 1e8) (the window of the division identity), an index outside the unit's table.
template <typename T, bool F32CHAIN>
__device__ __forceinline__ float elem_fast(float x, float maxval, float bias, float Kf, const float2* tab, bool& ok) {
  using C = Chain<T, F32CHAIN>;
  const float xc = C::clamp_unit(x, maxval);
  const float ax = fabsf(xc);
  float lp;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(lp) : "f"(ax));
  const float mg = __fmaf_rn(fabsf(lp), 9.5367431640625e-07f, 9.5367431640625e-07f);   // 2^-20 (|lp| + 1)
  float t;
  bool sure;
  if (F32CHAIN) {
    const float f_lo = floor_small(__fadd_rn(__fsub_rn(lp, mg), bias));
    t = floor_small(__fadd_rn(__fadd_rn(lp, mg), bias));
    sure = f_lo == t;
  } else {
    const float r_lo = C::rd(__fsub_rn(lp, mg)), r_hi = C::rd(__fadd_rn(lp, mg));
    t = floor_small(C::rd(__fadd_rn(r_lo, bias)));      // |bias| <= 512 and ax normal: |t| < 2^22
    sure = r_lo == r_hi;
  }
  t = fmaxf(t, 1.0f);                                   // NaN -> 1
  const float ti = fminf(t, Kf);                        // Kf = 0 (no table for this unit): entry 0, never certified
  // 2^-40 <= ax < 1e8 in one unsigned compare on the bit patterns (also rejects inf / NaN / zero / subnormals)
  const bool in_window = (__float_as_uint(ax) - 0x2b800000u) < (0x4cbebc20u - 0x2b800000u);
  ok = in_window && sure && ti == t;
  const float2 sr = tab[(int)ti];
  const float q0 = __fmul_rn(xc, sr.y);
  const float er = __fmaf_rn(-sr.x, q0, xc);
  float q = __fadd_rn(C::rd(__fmaf_rn(er, sr.y, q0)), 0.0f);          // rd(x / s) + 0  (a T-valued sum: no second rounding)
  // |q| < 2^(m + 9): ls is the element's own exponent (or larger, when clamped to 1) and the roundings of the chain move
  // log2(s) by a few units at most for |bias| <= 512. torch.round = magic add (half to even), the sign survives
  // (-0.3 -> -0.0); "- 0" is the identity
  q = copysignf(__fsub_rn(__fadd_rn(q, 12582912.0f), 12582912.0f), q);
  return __fmul_rn(C::rd(q), sr.x);                     // the caller's pack rounds to T (the chain's last rd)
}

// what elem_fast does not certify
template <typename T, bool F32CHAIN>
__device__ __noinline__ float elem_slow(float x, float maxval, float bias, float m_bits, float Kf, const float2* tab) {
  using C = Chain<T, F32CHAIN>;
  const float xc = C::clamp_unit(x, maxval);
  // a zero element: ls = 1, 0 / s_1 = 0, "+ 0" makes it +0 whatever its sign, 0 * s_1 = +0 for a finite scale
  if (xc == 0.0f && Kf >= 1.0f) return 0.0f;
  return C::qdq(xc, C::scale_of(xc, bias, m_bits));
}

template <typename T, bool F32CHAIN, int TPU, int NV>   // TPU threads per unit (8 | 32 | 256), NV 16-byte vectors per thread
__global__ void __launch_bounds__(256) floatq_fast_kernel(FParams p) {
  using C = Chain<T, F32CHAIN>;
  constexpr int VEC = 8;
  constexpr int UPB = 256 / TPU;
  __shared__ float2 tab[UPB][kTabMax];
  __shared__ float red_mx[8], red_mn[8];
  const int sub = threadIdx.x / TPU, lane = threadIdx.x % TPU;
  const float two_e = powf(2.0f, p.e_bits);
  const float c_log = C::rd(log2f(__fsub_rn(2.0f, powf(2.0f, -p.m_bits))));
  const int nvec = (int)(p.unit / VEC);
  const T* in = reinterpret_cast<const T*>(p.in);
  T* out = reinterpret_cast<T*>(p.out);
  for (int64_t u0 = (int64_t)blockIdx.x * UPB; u0 < p.n_units; u0 += (int64_t)gridDim.x * UPB) {
    const int64_t u = u0 + sub;
    const bool live = u < p.n_units;
    const uint4* row = reinterpret_cast<const uint4*>(in + u * p.unit);
    uint4 v[NV];
    float mx = -INFINITY, mn = INFINITY;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int j = lane + i * TPU;
      if (live && j < nvec) {
        v[i] = ld_stream(row + j);
        float f[VEC];
        TI<T>::unpack(v[i], f);
#pragma unroll
        for (int k = 0; k < VEC; ++k) {
          mx = max_nan(mx, f[k]);
          mn = min_nan(mn, f[k]);
        }
      }
    }
#pragma unroll
    for (int o = (TPU < 32 ? TPU : 32) / 2; o > 0; o >>= 1) {
      mx = max_nan(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      mn = min_nan(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    }
    if (TPU == 256) {
      __syncthreads();   // the previous unit's table readers are done
      if ((threadIdx.x & 31) == 0) {
        red_mx[threadIdx.x >> 5] = mx;
        red_mn[threadIdx.x >> 5] = mn;
      }
      __syncthreads();
      mx = red_mx[0];
      mn = red_mn[0];
#pragma unroll
      for (int w = 1; w < 8; ++w) {
        mx = max_nan(mx, red_mx[w]);
        mn = min_nan(mn, red_mn[w]);
      }
    } else {
      __syncwarp();      // the previous unit's table readers are done
    }
    const float maxval = max_nan(mx, -mn);
    const float bias = C::bias_of(maxval, two_e, c_log);
    // K = ls(maxval): the largest table index an element of this unit can ask for
    float tk = floorf(C::rd(__fadd_rn(C::rd(log2f(maxval)), bias)));
    tk = (tk != tk) ? tk : fmaxf(tk, 1.0f);
    // 0: no table (NaN / inf / implausible bias): literal chain
    int K = (live && tk >= 1.0f && tk < (float)kTabMax && fabsf(bias) <= 512.0f) ? (int)tk : 0;
    bool all_ok = true;
    for (int k = lane; k <= K; k += TPU) {                 // entry 0: a harmless stand-in for units without a table
      const float e2 = C::rd(__fsub_rn(C::rd(__fsub_rn((float)k, p.m_bits)), bias));
      float sk = k == 0 ? 1.0f : C::rd(powf(2.0f, e2));
      if (sk == 0.0f) sk = 1.0f;
      all_ok = all_ok && sk >= 1e-20f && sk <= 1e20f;      // the window of the division identity
      tab[sub][k] = make_float2(sk, __frcp_rn(sk));
    }
    // one scale outside the window: the whole unit takes the literal chain (K = 0)
    if (TPU == 256) {
      if (!__syncthreads_and(all_ok)) K = 0;
    } else {
      unsigned bad = __ballot_sync(0xffffffffu, !all_ok);
      if (TPU < 32) bad &= ((1u << TPU) - 1u) << ((threadIdx.x & 31) / TPU * TPU);
      if (bad) K = 0;
      __syncwarp();
    }
    if (!live) continue;
    const float Kf = (float)K;
    uint4* orow = reinterpret_cast<uint4*>(out + u * p.unit);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int j = lane + i * TPU;
      if (j < nvec) {
        float f[VEC];
        TI<T>::unpack(v[i], f);
        float y[VEC];
        unsigned redo = 0;
#pragma unroll
        for (int k = 0; k < VEC; ++k) {
          bool ok;
          y[k] = elem_fast<T, F32CHAIN>(f[k], maxval, bias, Kf, tab[sub], ok);
          redo |= ok ? 0u : (1u << k);
        }
        if (redo) {
#pragma unroll
          for (int k = 0; k < VEC; ++k)
            if (redo & (1u << k)) y[k] = elem_slow<T, F32CHAIN>(f[k], maxval, bias, p.m_bits, Kf, tab[sub]);
        }
#pragma unroll
        for (int k = 0; k < VEC; ++k) f[k] = y[k];
        st_stream(orow + j, TI<T>::pack(f));
      }
    }
  }
}

template <typename T, bool F32CHAIN>
bool launch_fast(const FParams& p, cudaStream_t st) {
  if (p.xc_out || p.s_out || !p.out || p.unit % 8 || p.unit > 16384) return false;
  if ((reinterpret_cast<uintptr_t>(p.in) | reinterpret_cast<uintptr_t>(p.out)) & 15) return false;
  const int nvec = (int)(p.unit / 8);
  auto grid = [&](int upb) { return (int)std::min<int64_t>((p.n_units + upb - 1) / upb, (int64_t)kNumSMs * 16); };
  if (nvec <= 8) floatq_fast_kernel<T, F32CHAIN, 8, 1><<<grid(32), 256, 0, st>>>(p);
  else if (nvec <= 32) floatq_fast_kernel<T, F32CHAIN, 32, 1><<<grid(8), 256, 0, st>>>(p);
  else if (nvec <= 64) floatq_fast_kernel<T, F32CHAIN, 32, 2><<<grid(8), 256, 0, st>>>(p);
  else if (nvec <= 128) floatq_fast_kernel<T, F32CHAIN, 32, 4><<<grid(8), 256, 0, st>>>(p);
  else if (nvec <= 512) floatq_fast_kernel<T, F32CHAIN, 256, 2><<<grid(1), 256, 0, st>>>(p);
  else if (nvec <= 1024) floatq_fast_kernel<T, F32CHAIN, 256, 4><<<grid(1), 256, 0, st>>>(p);
  else floatq_fast_kernel<T, F32CHAIN, 256, 8><<<grid(1), 256, 0, st>>>(p);
  return true;
}

template <typename T>
int launch_dynamic(const FParams& p, bool f32chain, cudaStream_t st) {
  const int64_t unit = p.unit;
  if (sizeof(T) == 2 && !getenv("LCB_FLOATQ_LITERAL")) {
    using T16 = typename std::conditional<sizeof(T) == 2, T, __nv_bfloat16>::type;
    if (f32chain ? launch_fast<T16, true>(p, st) : launch_fast<T16, false>(p, st)) {
      LCB_CUDA_CHECK_LAUNCH("floatq_fast");
      return LCB_OK;
    }
  }
  auto grid = [&](int upb) { return (int)std::min<int64_t>((p.n_units + upb - 1) / upb, (int64_t)kNumSMs * 8); };
#define LCB_FQ_LAUNCH(F32, L, UPB) floatq_dynamic_kernel<T, F32, L><<<grid(UPB), 256, 0, st>>>(p)
  if (f32chain) {
    if (unit <= 64) LCB_FQ_LAUNCH(true, 8, 32);
    else if (unit <= 1024) LCB_FQ_LAUNCH(true, 32, 8);
    else LCB_FQ_LAUNCH(true, 0, 1);
  } else {
    if (unit <= 64) LCB_FQ_LAUNCH(false, 8, 32);
    else if (unit <= 1024) LCB_FQ_LAUNCH(false, 32, 8);
    else LCB_FQ_LAUNCH(false, 0, 1);
  }
#undef LCB_FQ_LAUNCH
  LCB_CUDA_CHECK_LAUNCH("floatq_dynamic");
  return LCB_OK;
}

}  // namespace
}  // namespace lcb

extern "C" int lcb_floatquant_elem(const void* in, void* out, void* xc_out, void* scales_out, int64_t n_units,
                                   int64_t unit_size, int32_t e_bits, int32_t m_bits, int32_t dtype, void* stream) {
  using namespace lcb;
  if (!in || n_units < 0 || unit_size < 1 || e_bits < 1 || e_bits > 8 || m_bits < 0 || m_bits > 10) return LCB_E_INVALID;
  if (!out && !xc_out && !scales_out) return LCB_E_INVALID;
  if (dtype != LCB_F32 && dtype != LCB_F16 && dtype != LCB_BF16) return LCB_E_UNSUPPORTED;
  if (n_units == 0) return LCB_OK;
  const bool f32chain = e_bits >= 5 || dtype == LCB_F32;
  FParams p;
  p.in = in;
  p.out = out;
  p.xc_out = xc_out;
  p.s_out = scales_out;
  p.n_units = n_units;
  p.unit = unit_size;
  p.e_bits = (float)e_bits;
  p.m_bits = (float)m_bits;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  switch (dtype) {
    case LCB_F32: return launch_dynamic<float>(p, true, st);
    case LCB_F16: return launch_dynamic<__half>(p, f32chain, st);
    default: return launch_dynamic<__nv_bfloat16>(p, f32chain, st);
  }
}

extern "C" int lcb_floatquant_elem_static(const void* in, const void* scales, void* out, int64_t n, int32_t dtype,
                                          int32_t scales_dtype, void* stream) {
  using namespace lcb;
  if (!in || !scales || !out || n < 0) return LCB_E_INVALID;
  if (n == 0) return LCB_OK;
  if (scales_dtype != dtype && scales_dtype != LCB_F32) return LCB_E_UNSUPPORTED;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int grid = (int)std::min<int64_t>((n + 255) / 256, (int64_t)kNumSMs * 16);
  const bool f32 = scales_dtype == LCB_F32;
  switch (dtype) {
    case LCB_F32:
      floatq_static_kernel<float, float, true><<<grid, 256, 0, st>>>((const float*)in, (const float*)scales, (float*)out, n);
      break;
    case LCB_F16:
      if (f32) floatq_static_kernel<__half, float, true><<<grid, 256, 0, st>>>((const __half*)in, (const float*)scales, (__half*)out, n);
      else floatq_static_kernel<__half, __half, false><<<grid, 256, 0, st>>>((const __half*)in, (const __half*)scales, (__half*)out, n);
      break;
    case LCB_BF16:
      if (f32) floatq_static_kernel<__nv_bfloat16, float, true><<<grid, 256, 0, st>>>((const __nv_bfloat16*)in, (const float*)scales, (__nv_bfloat16*)out, n);
      else floatq_static_kernel<__nv_bfloat16, __nv_bfloat16, false><<<grid, 256, 0, st>>>((const __nv_bfloat16*)in, (const __nv_bfloat16*)scales, (__nv_bfloat16*)out, n);
      break;
    default: return LCB_E_UNSUPPORTED;
  }
  LCB_CUDA_CHECK_LAUNCH("floatq_static");
  return LCB_OK;
}
