// fakequant.cu — single-pass grouped quant-dequant (subsystem 1 of the north star).
//
// One HBM read + one HBM write per element. Kernels:
//   fq_group_kernel   : (fq_group.cuh) units of <= 32 lanes x 8 x 16 B; a sub-warp of 4..32 lanes owns a unit, 8 vectors
//                       in flight per lane, 128-bit coalesced loads, packed xor-shuffle min/max, round + clamp + dequant
//                       on packed 16-bit pairs in registers. fq_group_multi_kernel: several tensors in one launch.
//   fq_wrow_kernel    : rows of up to 2048 vectors (per_channel / per_token): 2 / 4 / 8 warps per row, one named barrier.
//   fq_row_stream_kernel : longer rows, CTA per row, second pass re-reads L2-resident data.
//   fq_block2d_kernel : per_block 128x128 tiles (fp32 math), CTA per tile, ragged edges zero-padded.
//   fq_static_kernel  : scales / zeros given (fake_quant_*_static, per_tensor phase 2), pure elementwise.
//   fq_generic_kernel : any unit shape / alignment / exotic spec; literal arithmetic, warp per unit.
//   minmax kernels    : per_tensor phase 1.
#include <type_traits>
#include "tc.cuh"
#include "fq_math.cuh"

namespace lcb {

// ------------------------------------------------------------------------------------------------
// store VEC results
// ------------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ void store_vec(void* out, int64_t e0, const float* val, const float* code,
                                          const SpecD& sp) {
  constexpr int VEC = TI<T>::VEC;
  switch (sp.out_mode) {
    case LCB_OUT_FAKE: {
      uint4 p = TI<T>::pack(val);
      st_stream(reinterpret_cast<T*>(out) + e0, p);
    } break;
    case LCB_OUT_CODE_T: {
      uint4 p = TI<T>::pack(code);
      st_stream(reinterpret_cast<T*>(out) + e0, p);
    } break;
    case LCB_OUT_I8:
    case LCB_OUT_U8: {
      uint32_t w[VEC / 4];
#pragma unroll
      for (int i = 0; i < VEC / 4; ++i) {
        uint32_t b = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) b |= ((uint32_t)((int)code[4 * i + j]) & 0xffu) << (8 * j);
        w[i] = b;
      }
      if (VEC == 8)
        *reinterpret_cast<uint2*>(reinterpret_cast<uint8_t*>(out) + e0) = make_uint2(w[0], w[VEC / 4 - 1]);
      else
        *reinterpret_cast<uint32_t*>(reinterpret_cast<uint8_t*>(out) + e0) = w[0];
    } break;
    case LCB_OUT_FP8: {
      uint32_t w[VEC / 4];
#pragma unroll
      for (int i = 0; i < VEC / 4; ++i) {
        uint32_t b = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) b |= (uint32_t)f32_to_fp8_byte(code[4 * i + j], sp.kind) << (8 * j);
        w[i] = b;
      }
      if (VEC == 8)
        *reinterpret_cast<uint2*>(reinterpret_cast<uint8_t*>(out) + e0) = make_uint2(w[0], w[VEC / 4 - 1]);
      else
        *reinterpret_cast<uint32_t*>(reinterpret_cast<uint8_t*>(out) + e0) = w[0];
    } break;
    case LCB_OUT_I32: {
      int32_t* o = reinterpret_cast<int32_t*>(out) + e0;
#pragma unroll
      for (int i = 0; i < VEC / 4; ++i)
        st_stream(o + 4 * i, make_uint4((uint32_t)(int)code[4 * i], (uint32_t)(int)code[4 * i + 1],
                                        (uint32_t)(int)code[4 * i + 2], (uint32_t)(int)code[4 * i + 3]));
    } break;
    default:
      break;
  }
}

// quantize VEC fp32 values of one unit and store them
template <typename T, typename M, int KINDC>  // KINDC: 0 int, 1 fp8
__device__ __forceinline__ void quant_store_vec(const float* f, const UnitQ& u, const SpecD& sp, void* out,
                                                int64_t e0) {
  constexpr int VEC = TI<T>::VEC;
  float val[VEC], code[VEC];
  if (!u.exact) {
#pragma unroll
    for (int i = 0; i < VEC; ++i) {
      if (KINDC == 0)
        val[i] = fq_elem_fast_int<M>(f[i], u, &code[i]);
      else
        val[i] = fq_elem_fast_fp8<M>(f[i], u, sp.kind, &code[i]);
    }
    if (KINDC == 0 && sp.out_mode != LCB_OUT_FAKE) {
#pragma unroll
      for (int i = 0; i < VEC; ++i) code[i] = __fadd_rn(code[i], u.z);
    }
  } else {
#pragma unroll
    for (int i = 0; i < VEC; ++i) {
      float2 vc = fq_elem_literal(f[i], u, sp);
      val[i] = vc.x;
      code[i] = vc.y;
    }
  }
  store_vec<T>(out, e0, val, code, sp);
}

__device__ __forceinline__ void write_qparams(void* scales, void* zeros, int64_t unit, const UnitQ& u,
                                              const SpecD& sp) {
  if (scales) store_rt(scales, unit, u.s_out, sp.qparam_dtype);
  if (zeros && !sp.sym && sp.kind == LCB_KIND_INT) store_rt(zeros, unit, u.z, sp.qparam_dtype);
}

}  // namespace lcb
#include "fq_group.cuh"
namespace lcb {

// ------------------------------------------------------------------------------------------------
// row kernel: one CTA per unit, row cached in registers (NV vectors per thread)
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void block_minmax(float& mx, float& mn, float* smem /*[2*33]*/) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    mx = max_nan(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    mn = min_nan(mn, __shfl_xor_sync(0xffffffffu, mn, o));
  }
  const int w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  if ((threadIdx.x & 31) == 0) {
    smem[w] = mx;
    smem[32 + w] = mn;
  }
  __syncthreads();
  float a = (int)(threadIdx.x & 31) < nw ? smem[threadIdx.x & 31] : -INFINITY;
  float b = (int)(threadIdx.x & 31) < nw ? smem[32 + (threadIdx.x & 31)] : INFINITY;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    a = max_nan(a, __shfl_xor_sync(0xffffffffu, a, o));
    b = min_nan(b, __shfl_xor_sync(0xffffffffu, b, o));
  }
  mx = a;
  mn = b;
  __syncthreads();
}

// streaming variant for rows that do not fit the register cache: second pass re-reads (L2-resident) data
template <typename T, int KINDC>
__global__ void __launch_bounds__(512)
    fq_row_stream_kernel(const T* __restrict__ in, void* __restrict__ out, void* __restrict__ scales,
                         void* __restrict__ zeros, int64_t n_units, int64_t unit_len, SpecD sp) {
  constexpr int VEC = TI<T>::VEC;
  __shared__ float red[64];
  const int64_t nvec = unit_len / VEC;
  for (int64_t unit = blockIdx.x; unit < n_units; unit += gridDim.x) {
    const T* src = in + unit * unit_len;
    float mx = -INFINITY, mn = INFINITY;
    for (int64_t v = threadIdx.x; v < nvec; v += blockDim.x) {
      float f[VEC];
      uint4 raw = *reinterpret_cast<const uint4*>(src + v * VEC);
      TI<T>::unpack(raw, f);
#pragma unroll
      for (int i = 0; i < VEC; ++i) {
        mx = max_nan(mx, f[i]);
        mn = min_nan(mn, f[i]);
      }
    }
    block_minmax(mx, mn, red);
    UnitQ u;
    qparams_from_range(mn, mx, sp, u);
    finish_unit(sp, u);
    if (!(finite_f(mx) && finite_f(mn))) u.exact = 1;
    if (threadIdx.x == 0) write_qparams(scales, zeros, unit, u, sp);
    if (sp.out_mode == LCB_OUT_NONE) continue;
    for (int64_t v = threadIdx.x; v < nvec; v += blockDim.x) {
      float f[VEC];
      uint4 raw = *reinterpret_cast<const uint4*>(src + v * VEC);
      TI<T>::unpack(raw, f);
      quant_store_vec<T, T, KINDC>(f, u, sp, out, unit * unit_len + v * VEC);
    }
  }
}

}  // namespace lcb
#include "fakequant_part2.cuh"
