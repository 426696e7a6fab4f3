// mse_range.cu — calib_algo = 'mse': the shrink-grid range search of BaseQuantizer.get_mse_range (quant.py:145-203).
//
// Reference, per quantization unit (a row of the reshaped [units, L] tensor), everything in fp32:
//   (mn, mx) = (amin, amax);  best = inf
//   for i in 0 .. int(maxshrink * mse_grid) - 1:   p = 1 - i / mse_grid          (Python double, used as an fp32 scalar)
//       (xmin, xmax) = (p * mn, p * mx);  qparams = get_qparams(xmin, xmax)
//       err = sum_j |quant_dequant(x_j) - x_j| ** norm
//       if err < best: best = err; (mn, mx) = (xmin, xmax)     <- the reference writes through views of min_val / max_val
//                                                                 (quant.py:164, 193-195), so later candidates shrink the
//                                                                 KEPT range; reproduced here on purpose.
// The torch version makes ~12 passes over an fp32 copy of the tensor per grid step (80 steps by default). Here the unit is
// read from HBM once, parked in shared memory as fp32, and all steps run out of shared memory: the kernel is bound by the
// fp32 pipes (one IEEE-exact quotient, one rint, one powf per element and step), not by memory.
//   L <= 2048 : one warp per unit, 8 units per CTA        L > 2048 : one CTA per unit (row of a per_channel weight)
#include "../../include/lcb200.h"
#include "fq_math.cuh"

namespace lcb {
namespace {

constexpr int MSE_WARP_MAX_L = 2048;

__device__ __forceinline__ float warp_sum_all(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);   // butterfly: every lane ends with the same bits
  return v;
}
__device__ __forceinline__ float warp_max_nan_all(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = max_nan(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_min_nan_all(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = min_nan(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// sum over this thread's elements of |qdq(x) - x| ** norm for the candidate range (xmin, xmax)
template <int STRIDE>
__device__ __forceinline__ float candidate_err(const float* __restrict__ xs, int L, int first, float xmin, float xmax,
                                               const SpecD& sp, float norm) {
  UnitQ u;
  qparams_from_range(xmin, xmax, sp, u);
  finish_unit(sp, u);
  float err = 0.f;
  if (!u.exact) {
#pragma unroll 4
    for (int j = first; j < L; j += STRIDE) {
      const float x = xs[j];
      float code;
      const float dq = fq_elem_fast_int<float>(x, u, &code);
      err += powf(fabsf(__fsub_rn(dq, x)), norm);
    }
  } else {
    for (int j = first; j < L; j += STRIDE) {
      const float x = xs[j];
      const float dq = fq_elem_literal(x, u, sp).x;
      err += powf(fabsf(__fsub_rn(dq, x)), norm);
    }
  }
  return err;
}

__device__ __forceinline__ float shrink_factor(int i, double grid) { return (float)(1.0 - (double)i / grid); }

template <typename T>
__global__ void __launch_bounds__(256)
    mse_range_warp_kernel(const T* __restrict__ x, int64_t n_units, int L, SpecD sp, int n_steps, double grid, float norm,
                          float* __restrict__ mins, float* __restrict__ maxs) {
  extern __shared__ float smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* xs = smem + (size_t)warp * L;
  for (int64_t unit = (int64_t)blockIdx.x * 8 + warp; unit < n_units; unit += (int64_t)gridDim.x * 8) {
    const T* src = x + unit * L;
    float mn = INFINITY, mx = -INFINITY;
    __syncwarp();
    for (int j = lane; j < L; j += 32) {
      const float v = TI<T>::load1(src + j);
      xs[j] = v;
      mn = min_nan(mn, v);
      mx = max_nan(mx, v);
    }
    mn = warp_min_nan_all(mn);
    mx = warp_max_nan_all(mx);
    __syncwarp();
    float best = INFINITY;
    for (int i = 0; i < n_steps; ++i) {
      const float p = shrink_factor(i, grid);
      const float xmin = __fmul_rn(p, mn), xmax = __fmul_rn(p, mx);
      const float err = warp_sum_all(candidate_err<32>(xs, L, lane, xmin, xmax, sp, norm));
      if (err < best) {
        best = err;
        mn = xmin;
        mx = xmax;
      }
    }
    if (lane == 0) {
      mins[unit] = mn;
      maxs[unit] = mx;
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
    mse_range_cta_kernel(const T* __restrict__ x, int64_t n_units, int L, SpecD sp, int n_steps, double grid, float norm,
                         float* __restrict__ mins, float* __restrict__ maxs) {
  extern __shared__ float smem[];
  __shared__ float part[2][8];
  __shared__ float part_mn[8], part_mx[8];
  float* xs = smem;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int64_t unit = blockIdx.x; unit < n_units; unit += gridDim.x) {
    const T* src = x + unit * L;
    float mn = INFINITY, mx = -INFINITY;
    __syncthreads();
    for (int j = threadIdx.x; j < L; j += 256) {
      const float v = TI<T>::load1(src + j);
      xs[j] = v;
      mn = min_nan(mn, v);
      mx = max_nan(mx, v);
    }
    mn = warp_min_nan_all(mn);
    mx = warp_max_nan_all(mx);
    if (lane == 0) {
      part_mn[warp] = mn;
      part_mx[warp] = mx;
    }
    __syncthreads();
    mn = part_mn[0];
    mx = part_mx[0];
#pragma unroll
    for (int w = 1; w < 8; ++w) {
      mn = min_nan(mn, part_mn[w]);
      mx = max_nan(mx, part_mx[w]);
    }
    float best = INFINITY;
    for (int i = 0; i < n_steps; ++i) {
      const float p = shrink_factor(i, grid);
      const float xmin = __fmul_rn(p, mn), xmax = __fmul_rn(p, mx);
      const float e = warp_sum_all(candidate_err<256>(xs, L, threadIdx.x, xmin, xmax, sp, norm));
      float* pp = part[i & 1];     // double-buffered partials: one barrier per step
      if (lane == 0) pp[warp] = e;
      __syncthreads();
      float err = pp[0];
#pragma unroll
      for (int w = 1; w < 8; ++w) err += pp[w];      // same order in every thread: the branch below is CTA-uniform
      if (err < best) {
        best = err;
        mn = xmin;
        mx = xmax;
      }
    }
    if (threadIdx.x == 0) {
      mins[unit] = mn;
      maxs[unit] = mx;
    }
  }
}

template <typename T>
int launch(const void* x, int64_t n_units, int64_t L, const SpecD& sp, int n_steps, double grid, float norm, float* mins,
           float* maxs, cudaStream_t st) {
  if (L <= MSE_WARP_MAX_L) {
    const size_t smem = (size_t)8 * L * sizeof(float);
    if (smem > 48 * 1024)
      cudaFuncSetAttribute(mse_range_warp_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const int64_t ctas = (n_units + 7) / 8;
    const int64_t cap = (int64_t)kNumSMs * 8;
    mse_range_warp_kernel<T><<<(int)(ctas < cap ? ctas : cap), 256, smem, st>>>((const T*)x, n_units, (int)L, sp, n_steps, grid,
                                                                              norm, mins, maxs);
  } else {
    const size_t smem = (size_t)L * sizeof(float);
    if (smem > 200 * 1024) return LCB_E_UNSUPPORTED;
    if (smem > 48 * 1024)
      cudaFuncSetAttribute(mse_range_cta_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const int64_t cap = (int64_t)kNumSMs * 4;
    mse_range_cta_kernel<T><<<(int)(n_units < cap ? n_units : cap), 256, smem, st>>>((const T*)x, n_units, (int)L, sp, n_steps,
                                                                                    grid, norm, mins, maxs);
  }
  LCB_CUDA_CHECK_LAUNCH("mse_range");
  return LCB_OK;
}

}  // namespace
}  // namespace lcb

extern "C" int lcb_mse_range(const void* in, void* mins, void* maxs, int64_t n_units, int64_t unit_len,
                             const lcb_quant_spec* spec, int32_t n_steps, double mse_grid, float norm, void* stream) {
  using namespace lcb;
  if (!in || !mins || !maxs || !spec || n_units < 0 || unit_len < 1 || n_steps < 0 || !(mse_grid > 0.0)) return LCB_E_INVALID;
  if (spec->dtype < LCB_F32 || spec->dtype > LCB_BF16) return LCB_E_UNSUPPORTED;
  // FloatQuantizer: a shrunk range overflows the fp8 grid and what happens then is qtorch.float_quantize's business
  // (not installable offline, so it cannot be pinned): integer grids only
  if (spec->kind != LCB_KIND_INT) return LCB_E_UNSUPPORTED;
  if (n_units == 0) return LCB_OK;
  SpecD sp;   // the search runs on tensor.float() (quant.py:151): fp32 math, fp32 ranges
  sp.dtype = spec->dtype;
  sp.math_dtype = sp.qparam_dtype = sp.qp_math = LCB_F32;
  sp.kind = spec->kind;
  sp.sym = spec->sym;
  sp.round_zp = spec->round_zp;
  sp.out_mode = LCB_OUT_FAKE;
  sp.qmin = spec->qmin;
  sp.qmax = spec->qmax;
  sp.osf = 1.0f;
  sp.clip = 1.0f;
  const bool ints = (floorf(sp.qmin) == sp.qmin) && (floorf(sp.qmax) == sp.qmax) && fabsf(sp.qmin) <= 256.f &&
                    fabsf(sp.qmax) <= 256.f;
  sp.fast_ok = sp.round_zp && ints;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  switch (spec->dtype) {
    case LCB_F32: return launch<float>(in, n_units, unit_len, sp, n_steps, mse_grid, norm, (float*)mins, (float*)maxs, st);
    case LCB_F16: return launch<__half>(in, n_units, unit_len, sp, n_steps, mse_grid, norm, (float*)mins, (float*)maxs, st);
    default: return launch<__nv_bfloat16>(in, n_units, unit_len, sp, n_steps, mse_grid, norm, (float*)mins, (float*)maxs, st);
  }
}
