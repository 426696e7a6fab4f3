// misc.cu — small bandwidth-bound helpers: block-wise FP8 dequant, column |x| statistics.
#include "common.cuh"

namespace lcb {

// y[r, c] = fp8(codes[r, c]) * scales[r / bs, c / bs]   (kernel.py:86-138 / quant.py:18-31)
// 16 codes per thread: one 16-byte load, packed fp8x2 -> f16x2 conversions (exact), fp32 product, 16-byte stores
template <typename TO, int E5M2>
__global__ void __launch_bounds__(256)
    fp8_block_dequant_vec16_kernel(const uint8_t* __restrict__ codes, const float* __restrict__ scales, TO* __restrict__ out,
                                   int64_t rows, int64_t cols, int64_t bs) {
  const int64_t ncb = (cols + bs - 1) / bs;
  const int64_t nv = rows * cols / 16;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nv; v += stride) {
    const int64_t e = v * 16, r = e / cols, c = e % cols;
    const float s = __ldg(scales + (r / bs) * ncb + c / bs);
    const uint4 raw = ld_stream(codes + e);
    const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
    float f[16];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const uint16_t b2 = (uint16_t)(w[i >> 1] >> (16 * (i & 1)));
      uint32_t h2;
      if (E5M2) asm("cvt.rn.f16x2.e5m2x2 %0, %1;" : "=r"(h2) : "h"(b2));
      else asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(h2) : "h"(b2));
      const float2 t = __half22float2(*reinterpret_cast<const __half2*>(&h2));
      f[2 * i] = __fmul_rn(t.x, s);
      f[2 * i + 1] = __fmul_rn(t.y, s);
    }
    constexpr int V = TI<TO>::VEC;
#pragma unroll
    for (int k = 0; k < 16 / V; ++k) st_stream(out + e + k * V, TI<TO>::pack(f + k * V));
  }
}

template <typename TO>
__global__ void __launch_bounds__(256)
    fp8_block_dequant_kernel(const uint8_t* __restrict__ codes, const float* __restrict__ scales, TO* __restrict__ out,
                             int64_t rows, int64_t cols, int64_t bs, int kind) {
  const int64_t ncb = (cols + bs - 1) / bs;
  const int64_t total = rows * cols;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const bool vec = (cols % 8 == 0) && (bs % 8 == 0);
  if (vec) {
    const int64_t nv = total / 8;
    for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nv; v += stride) {
      const int64_t e = v * 8, r = e / cols, c = e % cols;
      const float s = scales[(r / bs) * ncb + c / bs];
      const uint2 raw = *reinterpret_cast<const uint2*>(codes + e);
      const uint32_t w[2] = {raw.x, raw.y};
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float f = fp8_byte_to_f32((uint8_t)(w[i >> 2] >> (8 * (i & 3))), kind);
        TI<TO>::store1(out + e + i, __fmul_rn(f, s));
      }
    }
  } else {
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += stride) {
      const int64_t r = e / cols, c = e % cols;
      TI<TO>::store1(out + e, __fmul_rn(fp8_byte_to_f32(codes[e], kind), scales[(r / bs) * ncb + c / bs]));
    }
  }
}

// out[c] = max(out[c], max_r |x[r, c]|)   or   out[c] += sum_r |x[r, c]|
// grid: x = column tiles of 256 columns (threads = 128, 2 columns per thread via 32-bit loads), y = row slabs.
template <typename T>
__global__ void __launch_bounds__(128)
    col_absstat_kernel(const T* __restrict__ x, float* __restrict__ out, int64_t rows, int64_t cols, int mean) {
  const int64_t c = (int64_t)blockIdx.x * 128 + threadIdx.x;
  if (c >= cols) return;
  const int64_t r0 = (int64_t)blockIdx.y * rows / gridDim.y, r1 = (int64_t)(blockIdx.y + 1) * rows / gridDim.y;
  float acc = 0.f;
  int64_t r = r0;
  for (; r + 4 <= r1; r += 4) {
    float v0 = fabsf(TI<T>::load1(x + r * cols + c)), v1 = fabsf(TI<T>::load1(x + (r + 1) * cols + c));
    float v2 = fabsf(TI<T>::load1(x + (r + 2) * cols + c)), v3 = fabsf(TI<T>::load1(x + (r + 3) * cols + c));
    if (mean) acc += (v0 + v1) + (v2 + v3);
    else acc = max_nan(max_nan(acc, max_nan(v0, v1)), max_nan(v2, v3));
  }
  for (; r < r1; ++r) {
    float v = fabsf(TI<T>::load1(x + r * cols + c));
    if (mean) acc += v; else acc = max_nan(acc, v);
  }
  if (mean) atomicAdd(out + c, acc);
  else atomicMax(reinterpret_cast<int*>(out + c), __float_as_int(acc));  // non-negative floats order like ints
}

}  // namespace lcb

extern "C" int lcb_fp8_block_dequant(const void* codes, const float* scales, void* out, int64_t rows, int64_t cols,
                                     int64_t block, int32_t kind, int32_t out_dtype, void* stream) {
  using namespace lcb;
  if (!codes || !scales || !out || block < 1) return LCB_E_INVALID;
  if (kind != LCB_KIND_FP8_E4M3 && kind != LCB_KIND_FP8_E5M2) return LCB_E_UNSUPPORTED;
  if (rows == 0 || cols == 0) return LCB_OK;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  int64_t work = (rows * cols / 8 + 255) / 256;
  int grid = (int)(work < 148 * 8 ? (work < 1 ? 1 : work) : 148 * 8);
  const uint8_t* c = reinterpret_cast<const uint8_t*>(codes);
  if (cols % 16 == 0 && block % 16 == 0 && !(reinterpret_cast<uintptr_t>(codes) & 15) && !(reinterpret_cast<uintptr_t>(out) & 15)) {
    int64_t w16 = (rows * cols / 16 + 255) / 256;
    const int g16 = (int)(w16 < 148 * 8 ? (w16 < 1 ? 1 : w16) : 148 * 8);
    const bool e5 = kind == LCB_KIND_FP8_E5M2;
#define LCB_DQ16(TO_)                                                                                                  \
  {                                                                                                                    \
    if (e5) fp8_block_dequant_vec16_kernel<TO_, 1><<<g16, 256, 0, st>>>(c, scales, (TO_*)out, rows, cols, block);       \
    else fp8_block_dequant_vec16_kernel<TO_, 0><<<g16, 256, 0, st>>>(c, scales, (TO_*)out, rows, cols, block);          \
    LCB_CUDA_CHECK_LAUNCH("fp8_block_dequant");                                                                        \
    return LCB_OK;                                                                                                     \
  }
    if (out_dtype == LCB_BF16) LCB_DQ16(__nv_bfloat16)
    if (out_dtype == LCB_F16) LCB_DQ16(__half)
    if (out_dtype == LCB_F32) LCB_DQ16(float)
#undef LCB_DQ16
    return LCB_E_UNSUPPORTED;
  }
  if (out_dtype == LCB_BF16) fp8_block_dequant_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(c, scales, (__nv_bfloat16*)out, rows, cols, block, kind);
  else if (out_dtype == LCB_F16) fp8_block_dequant_kernel<__half><<<grid, 256, 0, st>>>(c, scales, (__half*)out, rows, cols, block, kind);
  else if (out_dtype == LCB_F32) fp8_block_dequant_kernel<float><<<grid, 256, 0, st>>>(c, scales, (float*)out, rows, cols, block, kind);
  else return LCB_E_UNSUPPORTED;
  LCB_CUDA_CHECK_LAUNCH("fp8_block_dequant");
  return LCB_OK;
}

extern "C" int lcb_col_absstat(const void* x, float* out, int64_t rows, int64_t cols, int32_t dtype, int32_t mean,
                               void* stream) {
  using namespace lcb;
  if (!x || !out) return LCB_E_INVALID;
  if (rows == 0 || cols == 0) return LCB_OK;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  int gx = (int)((cols + 127) / 128);
  int64_t want = (148 * 8 + gx - 1) / gx;
  int gy = (int)(want < 1 ? 1 : (want > rows ? rows : want));
  dim3 grid(gx, gy);
  if (dtype == LCB_BF16) col_absstat_kernel<__nv_bfloat16><<<grid, 128, 0, st>>>((const __nv_bfloat16*)x, out, rows, cols, mean);
  else if (dtype == LCB_F16) col_absstat_kernel<__half><<<grid, 128, 0, st>>>((const __half*)x, out, rows, cols, mean);
  else if (dtype == LCB_F32) col_absstat_kernel<float><<<grid, 128, 0, st>>>((const float*)x, out, rows, cols, mean);
  else return LCB_E_UNSUPPORTED;
  LCB_CUDA_CHECK_LAUNCH("col_absstat");
  return LCB_OK;
}

// ---- sum((a - b)^2) with the reference's roundings: calculate_loss of the AWQ search (awq.py:134-145) ---------------------
// `(org_out - out).float().pow(2).mean()` is four passes over [tokens, hidden] in torch (bf16 difference, fp32 copy, fp32
// square, reduction: 6.4 GB for a 65536 x 4096 output); here one pass reads the two tensors once. Per element:
// d = rd(a - b) in the tensor dtype, d * d in fp32 — the same two roundings; the fp32 partial sums of a thread (a few hundred
// terms) are combined in fp64, so the result is within fp32 rounding of any summation order.
namespace lcb {
namespace {
template <typename T>
__global__ void __launch_bounds__(256) sqdiff_sum_kernel(const T* __restrict__ a, const T* __restrict__ b, double* __restrict__ out,
                                                        int64_t n) {
  constexpr int VEC = TI<T>::VEC;
  const int64_t nvec = n / VEC;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  float acc = 0.f;
  const bool vec_ok = ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b)) & 15) == 0;
  int64_t done = 0;
  if (vec_ok) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
      float fa[VEC], fb[VEC];
      TI<T>::unpack(ld_stream(reinterpret_cast<const uint4*>(a) + i), fa);
      TI<T>::unpack(ld_stream(reinterpret_cast<const uint4*>(b) + i), fb);
#pragma unroll
      for (int k = 0; k < VEC; ++k) {
        const float d = TI<T>::rd(__fsub_rn(fa[k], fb[k]));
        acc = __fadd_rn(acc, __fmul_rn(d, d));
      }
    }
    done = nvec * VEC;
  }
  for (int64_t i = done + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float d = TI<T>::rd(__fsub_rn(TI<T>::load1(a + i), TI<T>::load1(b + i)));
    acc = __fadd_rn(acc, __fmul_rn(d, d));
  }
  double s = (double)acc;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  __shared__ double part[8];
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += part[w];
    atomicAdd(out, t);
  }
}
}  // namespace
}  // namespace lcb

extern "C" int lcb_sqdiff_sum(const void* a, const void* b, double* out, int64_t n, int32_t dtype, void* stream) {
  using namespace lcb;
  if (!a || !b || !out || n < 0) return LCB_E_INVALID;
  if (n == 0) return LCB_OK;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int64_t want = (n / 8 + 255) / 256;
  const int grid = (int)(want < 1 ? 1 : (want > kNumSMs * 8 ? kNumSMs * 8 : want));
  if (dtype == LCB_BF16) sqdiff_sum_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)a, (const __nv_bfloat16*)b, out, n);
  else if (dtype == LCB_F16) sqdiff_sum_kernel<__half><<<grid, 256, 0, st>>>((const __half*)a, (const __half*)b, out, n);
  else if (dtype == LCB_F32) sqdiff_sum_kernel<float><<<grid, 256, 0, st>>>((const float*)a, (const float*)b, out, n);
  else return LCB_E_UNSUPPORTED;
  LCB_CUDA_CHECK_LAUNCH("sqdiff_sum");
  return LCB_OK;
}
