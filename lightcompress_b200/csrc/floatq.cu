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

template <typename T>
int launch_dynamic(const FParams& p, bool f32chain, cudaStream_t st) {
  const int64_t unit = p.unit;
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
