// hessian.cu — GPTQ Hessian accumulation S += X^T X on tcgen05 tensor cores (subsystem 2 of the north star).
//
// Reference: GPTQ.add_batch (gptq.py:254-295): H <- H*n/(n+b) + (2/(n+b)) X^T X per calibration batch, fp32 SGEMM.
// Here the running state is the UN-NORMALISED sum S = sum_tokens x x^T (fp32) and the scale 2/n is applied once by
// lcb_hessian_finalize; the two are the same number up to fp32 rounding (H_k = (2/k) * sum_{j<=k} X_j^T X_j).
//
// Kernel (one CTA per SM, persistent, warp-specialised):
//   warp 0  : TMA producer. X is [T, C] row-major, so both MMA operands are MN-major slices of the same matrix:
//             A = X[k0:k0+64, i*128 : +128], B = X[k0:k0+64, j*256 : +256], loaded as 64-column boxes with the
//             128-byte swizzle into a 4-stage ring (48 KB / stage). Out-of-range rows / columns are zero-filled by TMA.
//   warp 1  : MMA issuer. One thread issues tcgen05.mma.kind::f16 (M=128, N=256, K=16, fp32 accumulate in TMEM);
//             bf16*bf16 and fp16*fp16 products are exact in fp32, so the result is an fp32 sum of exact products like
//             the reference's SGEMM. Two 256-column accumulators (all 512 TMEM columns) double-buffer the epilogue.
//   warps 2-5: epilogue. tcgen05.ld 32 lanes x 32 columns -> swizzled smem -> TMA reduce-add (f32) into S. The L2 does
//             the read-modify-write, which also makes split-K (several CTAs per tile) free.
// Only tiles that touch the upper triangle (128-row block i, 128-column block jb >= i) are computed and written
// (SYRK): half the FLOPs and half the S traffic of the reference's full GEMM. fp32 inputs are split into bf16 hi + lo
// (x = hi + lo to 2^-17) and accumulated as hi*hi + hi*lo + lo*hi in the same kernel (three K segments).
#include <algorithm>
#include "tc_gemm.cuh"

namespace lcb {

namespace {
// ---- fp32 -> bf16 hi / lo split ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) split_f32_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ hi,
                                                        __nv_bfloat16* __restrict__ lo, int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float v = x[i];
    const __nv_bfloat16 h = __float2bfloat16_rn(v);
    hi[i] = h;
    lo[i] = __float2bfloat16_rn(v - __bfloat162float(h));
  }
}

// ---- SIMT fallback for shapes TMA cannot describe (C % 8 != 0): upper 128-blocks, same output contract ------------------
template <typename T>
__global__ void __launch_bounds__(256) hessian_simt_kernel(const T* __restrict__ x, float* __restrict__ S, int64_t T_, int64_t C) {
  __shared__ float xa[32][33], xb[32][33];
  const int bi = blockIdx.y, bj = blockIdx.x;
  if ((bj * 32) / 128 < (bi * 32) / 128) return;  // lower 128-block: not part of the contract
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  for (int64_t k0 = 0; k0 < T_; k0 += 32) {
    for (int r = ty; r < 32; r += 8) {
      const int64_t k = k0 + r;
      const int64_t ca = (int64_t)bi * 32 + tx, cb = (int64_t)bj * 32 + tx;
      xa[r][tx] = (k < T_ && ca < C) ? TI<T>::load1(x + k * C + ca) : 0.f;
      xb[r][tx] = (k < T_ && cb < C) ? TI<T>::load1(x + k * C + cb) : 0.f;
    }
    __syncthreads();
#pragma unroll 8
    for (int k = 0; k < 32; ++k) {
      const float b = xb[k][tx];
#pragma unroll
      for (int i = 0; i < 4; ++i) acc[i] = fmaf(xa[k][ty * 4 + i], b, acc[i]);
    }
    __syncthreads();
  }
  for (int i = 0; i < 4; ++i) {
    const int64_t r = (int64_t)bi * 32 + ty * 4 + i, c = (int64_t)bj * 32 + tx;
    if (r < C && c < C) S[r * C + c] += acc[i];
  }
}

// ---- finalize: H = scale * sym(S) in place: every element below the diagonal is mirrored from above it ----------------
__global__ void __launch_bounds__(256) hessian_finalize_kernel(float* __restrict__ S, int64_t C, float scale) {
  __shared__ float t[32][33];
  const int bi = blockIdx.y, bj = blockIdx.x;  // 32 x 32 tiles; tiles with bj >= bi do the work
  if (bj < bi) return;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int r = ty; r < 32; r += 8) {
    const int64_t gr = (int64_t)bi * 32 + r, gc = (int64_t)bj * 32 + tx;
    float v = 0.f;
    if (gr < C && gc < C) {
      v = S[gr * C + gc] * scale;
      if (bi != bj || tx >= r) S[gr * C + gc] = v;
    }
    t[r][tx] = v;
  }
  __syncthreads();
  for (int r = ty; r < 32; r += 8) {
    const int64_t gr = (int64_t)bj * 32 + r, gc = (int64_t)bi * 32 + tx;
    if (gr < C && gc < C && (bi != bj || tx < r)) S[gr * C + gc] = t[tx][r];
  }
}

// ---- pack / unpack of the maintained part of S (128-row block i keeps columns i*128 ..): the payload of the one all-reduce
// of a token-sharded Hessian. Packed layout: block rows end to end, row-major inside a block row; one extra float at the
// end carries the batch count, so that sum and count travel in ONE collective.
__device__ __forceinline__ int64_t packed_row_offset(int64_t r, int64_t C) {
  const int64_t bi = r >> 7, rin = r & 127;
  // elements of block rows 0 .. bi-1: sum_{b<bi} rows_b * (C - 128 b) with rows_b = 128
  const int64_t before = 128 * (bi * C - 128 * (bi * (bi - 1) / 2));
  return before + rin * (C - 128 * bi);
}

template <bool PACK>
__global__ void __launch_bounds__(256) hessian_pack_kernel(float* __restrict__ S, float* __restrict__ P, int64_t C, float count) {
  const int64_t r = blockIdx.x;
  const int64_t c0 = (r >> 7) << 7;
  float* srow = S + r * C + c0;
  float* prow = P + packed_row_offset(r, C);
  const int64_t n = C - c0;
  if ((n & 3) == 0 && ((reinterpret_cast<uintptr_t>(srow) | reinterpret_cast<uintptr_t>(prow)) & 15) == 0) {
    for (int64_t i = threadIdx.x * 4; i < n; i += 1024) {
      if (PACK) *reinterpret_cast<float4*>(prow + i) = *reinterpret_cast<const float4*>(srow + i);
      else *reinterpret_cast<float4*>(srow + i) = *reinterpret_cast<const float4*>(prow + i);
    }
  } else {
    for (int64_t i = threadIdx.x; i < n; i += 256) {
      if (PACK) prow[i] = srow[i];
      else srow[i] = prow[i];
    }
  }
  if (PACK && r == 0 && threadIdx.x == 0) P[packed_row_offset(C, C)] = count;
}

__global__ void __launch_bounds__(256) hessian_finalize_dev_kernel(float* __restrict__ S, int64_t C, const float* __restrict__ count) {
  __shared__ float t[32][33];
  const float scale = 2.0f / *count;
  const int bi = blockIdx.y, bj = blockIdx.x;
  if (bj < bi) return;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int r = ty; r < 32; r += 8) {
    const int64_t gr = (int64_t)bi * 32 + r, gc = (int64_t)bj * 32 + tx;
    float v = 0.f;
    if (gr < C && gc < C) {
      v = S[gr * C + gc] * scale;
      if (bi != bj || tx >= r) S[gr * C + gc] = v;
    }
    t[r][tx] = v;
  }
  __syncthreads();
  for (int r = ty; r < 32; r += 8) {
    const int64_t gr = (int64_t)bj * 32 + r, gc = (int64_t)bi * 32 + tx;
    if (gr < C && gc < C && (bi != bj || tx < r)) S[gr * C + gc] = t[tx][r];
  }
}

inline int64_t packed_elems_host(int64_t C) {
  int64_t n = 0;
  for (int64_t r0 = 0; r0 < C; r0 += 128) n += std::min<int64_t>(128, C - r0) * (C - r0);
  return n;
}

}  // namespace

}  // namespace lcb

extern "C" int64_t lcb_hessian_packed_elems(int64_t C) { return C < 1 ? 0 : lcb::packed_elems_host(C) + 1; }

extern "C" int lcb_hessian_pack(const float* S, float* packed, int64_t C, double count, void* stream) {
  using namespace lcb;
  if (!S || !packed || C < 1) return LCB_E_INVALID;
  hessian_pack_kernel<true><<<(unsigned)C, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(const_cast<float*>(S), packed, C, (float)count);
  LCB_CUDA_CHECK_LAUNCH("hessian_pack");
  return LCB_OK;
}

extern "C" int lcb_hessian_unpack_finalize(float* S, const float* packed, int64_t C, void* stream) {
  using namespace lcb;
  if (!S || !packed || C < 1) return LCB_E_INVALID;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  hessian_pack_kernel<false><<<(unsigned)C, 256, 0, st>>>(S, const_cast<float*>(packed), C, 0.f);
  LCB_CUDA_CHECK_LAUNCH("hessian_unpack");
  dim3 grid((unsigned)((C + 31) / 32), (unsigned)((C + 31) / 32));
  hessian_finalize_dev_kernel<<<grid, 256, 0, st>>>(S, C, packed + (lcb_hessian_packed_elems(C) - 1));
  LCB_CUDA_CHECK_LAUNCH("hessian_finalize_dev");
  return LCB_OK;
}

extern "C" int64_t lcb_hessian_workspace_bytes(int64_t tokens, int64_t C, int32_t x_dtype) {
  if (x_dtype == LCB_F32) return tokens * C * 4 + 512;  // bf16 hi + lo copies
  return 0;
}

extern "C" int lcb_hessian_accum(float* S, const void* X, int64_t tokens, int64_t C, int32_t x_dtype, void* workspace,
                                 void* stream) {
  using namespace lcb;
  if (!S || !X || tokens < 0 || C < 1) return LCB_E_INVALID;
  if (x_dtype != LCB_F32 && x_dtype != LCB_F16 && x_dtype != LCB_BF16) return LCB_E_UNSUPPORTED;
  if (tokens == 0) return LCB_OK;
  if (C > (1 << 24) || tokens > (1ll << 30)) return LCB_E_SHAPE;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const bool tma_ok = (C % 8 == 0) && ((reinterpret_cast<uintptr_t>(X) & 15) == 0) && ((reinterpret_cast<uintptr_t>(S) & 15) == 0);
  if (!tma_ok) {
    dim3 grid((unsigned)((C + 31) / 32), (unsigned)((C + 31) / 32));
    if (x_dtype == LCB_F32) hessian_simt_kernel<float><<<grid, 256, 0, st>>>((const float*)X, S, tokens, C);
    else if (x_dtype == LCB_F16) hessian_simt_kernel<__half><<<grid, 256, 0, st>>>((const __half*)X, S, tokens, C);
    else hessian_simt_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)X, S, tokens, C);
    LCB_CUDA_CHECK_LAUNCH("hessian_simt");
    return LCB_OK;
  }
  const int fmt = x_dtype == LCB_F16 ? 0 : 1;
  if (x_dtype == LCB_F32) {
    if (!workspace) return LCB_E_INVALID;
    __nv_bfloat16* hi = reinterpret_cast<__nv_bfloat16*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~uintptr_t(255));
    __nv_bfloat16* lo = hi + tokens * C;
    const int64_t n = tokens * C;
    split_f32_kernel<<<(int)std::min<int64_t>((n + 255) / 256, kNumSMs * 16), 256, 0, st>>>((const float*)X, hi, lo, n);
    LCB_CUDA_CHECK_LAUNCH("hessian_split");
    const void* aa[3] = {hi, hi, lo};  // hi*hi + hi*lo + lo*hi
    const void* bb[3] = {hi, lo, hi};
    return tc_gemm_launch(false, true, false, 1, aa, bb, 3, S, C, C, tokens, C, C, C, st);
  }
  const void* aa[1] = {X};
  return tc_gemm_launch(false, true, false, fmt, aa, aa, 1, S, C, C, tokens, C, C, C, st);
}

extern "C" int lcb_hessian_finalize(float* S, int64_t C, double scale, void* stream) {
  using namespace lcb;
  if (!S || C < 1) return LCB_E_INVALID;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  dim3 grid((unsigned)((C + 31) / 32), (unsigned)((C + 31) / 32));
  hessian_finalize_kernel<<<grid, 256, 0, st>>>(S, C, (float)scale);
  LCB_CUDA_CHECK_LAUNCH("hessian_finalize");
  return LCB_OK;
}

// ---- allocation-free variant of the fp32 path (no counterpart in the reference)
// Callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it) get the bf16
// hi / lo split buffers from a file-scope device array instead of a caller workspace. Every GPU that loads the library
// holds the array until exit; every call shares it, so calls that overlap on one GPU (two streams, two threads) race.
// Nothing in it has to start at zero: split_f32_kernel writes each element the GEMM reads before the GEMM reads it.
namespace lcb {
namespace {
constexpr int64_t kNoAllocMaxElems = LCB_HESSIAN_NOALLOC_MAX_ELEMS;                       // tokens * C of one call
static __device__ __align__(256) unsigned char g_hessian_scratch[kNoAllocMaxElems * 4];  // bf16 hi + bf16 lo = 16 MiB

// In host code g_hessian_scratch is only a stand-in for the symbol, not a device address: the address a kernel can use
// comes from cudaGetSymbolAddress, asked once per device with that device current.
constexpr int kMaxDevices = 64;
std::atomic<void*> g_hessian_scratch_addr[kMaxDevices];

int hessian_scratch_for_current_device(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) {
    set_last_error("lcb_hessian_accum_f32_noalloc: cudaGetDevice", e);
    return LCB_E_CUDA;
  }
  const bool cached = dev >= 0 && dev < kMaxDevices;
  void* p = cached ? g_hessian_scratch_addr[dev].load(std::memory_order_acquire) : nullptr;
  if (!p) {
    e = cudaGetSymbolAddress(&p, g_hessian_scratch);
    if (e != cudaSuccess || !p) {
      set_last_error("lcb_hessian_accum_f32_noalloc: cudaGetSymbolAddress(g_hessian_scratch)", e);
      return LCB_E_CUDA;
    }
    if (cached) g_hessian_scratch_addr[dev].store(p, std::memory_order_release);  // racing threads store the same address
  }
  *out = p;
  return LCB_OK;
}
}  // namespace
}  // namespace lcb

extern "C" int lcb_hessian_accum_f32_noalloc(float* S, const float* X, int64_t tokens, int64_t C, void* stream) {
  using namespace lcb;
  if (!S || !X || tokens < 0 || C < 1) return LCB_E_INVALID;
  if (tokens == 0) return LCB_OK;
  if (tokens > kNoAllocMaxElems || C > kNoAllocMaxElems || tokens * C > kNoAllocMaxElems) {
    set_last_error("lcb_hessian_accum_f32_noalloc: tokens * C exceeds LCB_HESSIAN_NOALLOC_MAX_ELEMS (4194304); "
                   "split the tokens over several calls or use lcb_hessian_accum with a workspace",
                   cudaErrorInvalidValue);
    return LCB_E_SHAPE;
  }
  void* scratch = nullptr;
  const int rc = hessian_scratch_for_current_device(&scratch);
  if (rc != LCB_OK) return rc;
  return lcb_hessian_accum(S, X, tokens, C, LCB_F32, scratch, stream);
}
