// fakequant_part2.cuh — per_block, static, generic and per_tensor kernels (included by fakequant.cu)
#pragma once

namespace lcb {

// ------------------------------------------------------------------------------------------------
// per_block 128x128 (quant.py:137-138: |x| upcast to fp32 -> fp32 scales, fp32 math)
// ------------------------------------------------------------------------------------------------
template <typename T, int KINDC>
__global__ void __launch_bounds__(256)
    fq_block2d_kernel(const T* __restrict__ in, void* __restrict__ out, void* __restrict__ scales, int64_t rows,
                      int64_t cols, SpecD sp) {
  constexpr int VEC = TI<T>::VEC;
  constexpr int BS = 128;
  constexpr int VPR = BS / VEC;        // vectors per tile row
  constexpr int RPP = 256 / VPR;       // rows per pass
  constexpr int NP = BS / RPP;         // passes
  __shared__ float red[64];
  const int64_t ncb = (cols + BS - 1) / BS;
  const int64_t bi = blockIdx.x / ncb, bj = blockIdx.x % ncb;
  const int vr = threadIdx.x / VPR, vc = threadIdx.x % VPR;
  const int64_t col = bj * BS + vc * VEC;
  uint4 raw[NP];
  bool ok[NP];
#pragma unroll
  for (int p = 0; p < NP; ++p) {
    const int64_t row = bi * BS + p * RPP + vr;
    ok[p] = (row < rows) && (col < cols);
    if (ok[p]) raw[p] = ld_stream(in + row * cols + col);
  }
  float mx = 0.f;
#pragma unroll
  for (int p = 0; p < NP; ++p) {
    if (ok[p]) {
      float f[VEC];
      TI<T>::unpack(raw[p], f);
#pragma unroll
      for (int i = 0; i < VEC; ++i) mx = max_nan(mx, fabsf(f[i]));
    }
  }
  float mn = mx;
  block_minmax(mx, mn, red);
  UnitQ u;
  qparams_from_range(mx, mx, sp, u);
  finish_unit(sp, u);
  if (!finite_f(mx)) u.exact = 1;
  if (threadIdx.x == 0 && scales) store_rt(scales, blockIdx.x, u.s_out, sp.qparam_dtype);
  if (sp.out_mode == LCB_OUT_NONE) return;
#pragma unroll
  for (int p = 0; p < NP; ++p) {
    if (ok[p]) {
      const int64_t row = bi * BS + p * RPP + vr;
      float f[VEC];
      TI<T>::unpack(raw[p], f);
      quant_store_vec<T, float, KINDC>(f, u, sp, out, row * cols + col);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// static qparams: elementwise
// ------------------------------------------------------------------------------------------------
template <typename T, typename M, int KINDC>
__global__ void __launch_bounds__(256)
    fq_static_kernel(const T* __restrict__ in, void* __restrict__ out, const void* __restrict__ scales,
                     const void* __restrict__ zeros, int64_t rows, int64_t cols, int64_t ur, int64_t uc,
                     SpecD sp) {
  constexpr int VEC = TI<T>::VEC;
  const int64_t vpr = cols / VEC;
  const int64_t total = rows * vpr;
  const int64_t ncu = (cols + uc - 1) / uc;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t v0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v0 < total; v0 += 4 * stride) {
    uint4 raw[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int64_t v = v0 + k * stride;
      if (v < total) raw[k] = ld_stream(in + v * VEC);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int64_t v = v0 + k * stride;
      if (v >= total) continue;
      const int64_t row = v / vpr, col = (v % vpr) * VEC;
      const int64_t unit = (row / ur) * ncu + col / uc;
      UnitQ u;
      u.s = load_rt(scales, unit, sp.qparam_dtype);
      u.z = (zeros && !sp.sym && sp.kind == LCB_KIND_INT) ? load_rt(zeros, unit, sp.qparam_dtype) : 0.f;
      if (sp.kind != LCB_KIND_INT && u.s == 0.f) u.s = 1.f;
      finish_unit(sp, u);
      float f[VEC];
      TI<T>::unpack(raw[k], f);
      float am = 0.f;
#pragma unroll
      for (int i = 0; i < VEC; ++i) am = max_nan(am, fabsf(f[i]));
      if (!(__fmul_rn(am, u.r) < 1e30f)) u.exact = 1;
      quant_store_vec<T, M, KINDC>(f, u, sp, out, v * VEC);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// per_block 128x128, FP8 (DeepSeek-style block-wise weight quant, quant.py:1061-1081 with fp32 math): compile-time format
// and output, packed conversions. Per pair of elements: 2 x Markstein division (3 FP32 ops each), "+ 0" (-0 -> +0),
// ONE cvt.rn.satfinite.e4m3x2.f32 (or e5m2x2), then either the two bytes are the real-quant code (LCB_OUT_FP8) or ONE
// cvt.rn.f16x2.e4m3x2 brings them back for the dequant product (LCB_OUT_FAKE).
// ------------------------------------------------------------------------------------------------
template <int E5M2>
__device__ __forceinline__ uint32_t f32x2_to_fp8x2(float lo, float hi) {   // byte 0 = fp8(lo), byte 1 = fp8(hi)
  uint16_t r;
  if (E5M2) asm("cvt.rn.satfinite.e5m2x2.f32 %0, %1, %2;" : "=h"(r) : "f"(hi), "f"(lo));
  else asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(r) : "f"(hi), "f"(lo));
  return r;
}
template <int E5M2>
__device__ __forceinline__ float2 fp8x2_to_f32x2(uint32_t b) {
  uint32_t h2;
  const uint16_t s = (uint16_t)b;
  if (E5M2) asm("cvt.rn.f16x2.e5m2x2 %0, %1;" : "=r"(h2) : "h"(s));
  else asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(h2) : "h"(s));
  return __half22float2(*reinterpret_cast<const __half2*>(&h2));
}
template <typename T>
__device__ __noinline__ void block_vec_slow(uint4 raw, UnitQ u, SpecD sp, void* out, int64_t e0) {
  float f[TI<T>::VEC];
  TI<T>::unpack(raw, f);
  quant_store_vec<T, float, 1>(f, u, sp, out, e0);
}

template <typename T, int E5M2, int OUT_CODES>
__global__ void __launch_bounds__(256)
    fq_block2d_fp8_kernel(const T* __restrict__ in, void* __restrict__ out, void* __restrict__ scales, int64_t rows,
                          int64_t cols, SpecD sp) {
  constexpr int VEC = TI<T>::VEC;
  constexpr int BS = 128;
  constexpr int VPR = BS / VEC;        // vectors per tile row
  constexpr int RPP = 256 / VPR;       // rows per pass
  constexpr int NP = BS / RPP;         // passes
  __shared__ float red[64];
  const int64_t ncb = (cols + BS - 1) / BS;
  const int64_t bi = blockIdx.x / ncb, bj = blockIdx.x % ncb;
  const int vr = threadIdx.x / VPR, vc = threadIdx.x % VPR;
  const int64_t col = bj * BS + vc * VEC;
  uint4 raw[NP];
  bool ok[NP];
#pragma unroll
  for (int p = 0; p < NP; ++p) {
    const int64_t row = bi * BS + p * RPP + vr;
    ok[p] = (row < rows) && (col < cols);
    raw[p] = ok[p] ? ld_stream(in + row * cols + col) : make_uint4(0u, 0u, 0u, 0u);
  }
  float mx = 0.f;
#pragma unroll
  for (int p = 0; p < NP; ++p) {
    float f[VEC];
    TI<T>::unpack(raw[p], f);
#pragma unroll
    for (int i = 0; i < VEC; ++i) mx = max_nan(mx, fabsf(f[i]));
  }
  float mn = mx;
  block_minmax(mx, mn, red);
  UnitQ u;
  qparams_from_range(mx, mx, sp, u);
  finish_unit(sp, u);
  if (!finite_f(mx)) u.exact = 1;
  if (threadIdx.x == 0 && scales) store_rt(scales, blockIdx.x, u.s_out, sp.qparam_dtype);
  if (u.exact) {   // CTA-uniform, rare (non-finite range): the literal chain
#pragma unroll 1
    for (int p = 0; p < NP; ++p)
      if (ok[p]) block_vec_slow<T>(raw[p], u, sp, out, (bi * BS + p * RPP + vr) * cols + col);
    return;
  }
  const float s = u.s, r = u.r, so = u.s_out;
#pragma unroll
  for (int p = 0; p < NP; ++p) {
    if (!ok[p]) continue;
    const int64_t e0 = (bi * BS + p * RPP + vr) * cols + col;
    float f[VEC];
    TI<T>::unpack(raw[p], f);
    uint32_t b2[VEC / 2];
#pragma unroll
    for (int i = 0; i < VEC; i += 2) {
      float q[2];
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const float q0 = __fmul_rn(f[i + k], r);
        q[k] = __fadd_rn(__fmaf_rn(__fmaf_rn(-s, q0, f[i + k]), r, q0), 0.0f);   // RN(x / s), -0 -> +0
      }
      b2[i / 2] = f32x2_to_fp8x2<E5M2>(q[0], q[1]);
    }
    if (OUT_CODES) {
      uint32_t w[VEC / 4];
#pragma unroll
      for (int i = 0; i < VEC / 4; ++i) w[i] = b2[2 * i] | (b2[2 * i + 1] << 16);
      if (VEC == 8) *reinterpret_cast<uint2*>(reinterpret_cast<uint8_t*>(out) + e0) = make_uint2(w[0], w[VEC / 4 - 1]);
      else *reinterpret_cast<uint32_t*>(reinterpret_cast<uint8_t*>(out) + e0) = w[0];
    } else {
      float val[VEC];
#pragma unroll
      for (int i = 0; i < VEC; i += 2) {
        const float2 v = fp8x2_to_f32x2<E5M2>(b2[i / 2]);
        val[i] = __fmul_rn(v.x, so);
        val[i + 1] = __fmul_rn(v.y, so);
      }
      st_stream(reinterpret_cast<T*>(out) + e0, TI<T>::pack(val));
    }
  }
}

// Same arithmetic, 16-bit inputs, tile STAGED IN SHARED MEMORY BY TMA (one cp.async.bulk.tensor of the 128 x 128 box per
// CTA: ragged edges arrive zero-filled = the reference's zero padding, quant.py:632-642). The register version above
// keeps the tile in 8 vectors per thread across the CTA reduction, which costs it 128 registers and leaves 2 CTAs
// (64 KB of loads) per SM; here nothing is live across the reduction (both passes read shared memory), so 5-6 CTAs of
// 32 KB are resident per SM.
template <typename T, int E5M2, int OUT_CODES>
__device__ __forceinline__ void block2d_fp8_tma_tile(const CUtensorMap* tmp, void* __restrict__ out, void* __restrict__ scales,
                                                     int64_t rows, int64_t cols, int64_t tile_idx, const SpecD& sp) {
  constexpr int VEC = 8, BS = 128, VPR = BS / VEC, RPP = 256 / VPR, NP = BS / RPP;
  static_assert(sizeof(T) == 2, "16-bit inputs");
  __shared__ __align__(128) uint8_t tile[BS * BS * 2];
  __shared__ __align__(8) uint64_t bar;
  __shared__ float red[64];
  const int64_t ncb = (cols + BS - 1) / BS;
  const int64_t bi = tile_idx / ncb, bj = tile_idx % ncb;
  const int vr = threadIdx.x / VPR, vc = threadIdx.x % VPR;
  if (threadIdx.x == 0) {
    tc::mbar_init(&bar, 1);
    tc::fence_barrier_init();
    tc::mbar_expect_tx(&bar, BS * BS * 2);
    tc::tma_load_2d(tile, tmp, &bar, (int32_t)(bj * BS), (int32_t)(bi * BS));
  }
  __syncthreads();
  tc::mbar_wait(&bar, 0);
  const uint8_t* mine = tile + vr * (BS * 2) + vc * 16;      // pass p: + p * RPP rows
  // pass 1: |x| max on the packed words (sign bits cleared, NaN-propagating packed max)
  uint32_t m2 = 0u;
#pragma unroll
  for (int p = 0; p < NP; ++p) {
    const uint4 v = *reinterpret_cast<const uint4*>(mine + p * RPP * (BS * 2));
    const uint32_t msk = 0x7fff7fffu;
    uint32_t a = v.x & msk, b = v.y & msk, c = v.z & msk, d = v.w & msk;
    if (TI<T>::ID == LCB_BF16) {
      asm("max.NaN.bf16x2 %0, %1, %2;" : "=r"(a) : "r"(a), "r"(b));
      asm("max.NaN.bf16x2 %0, %1, %2;" : "=r"(c) : "r"(c), "r"(d));
      asm("max.NaN.bf16x2 %0, %1, %2;" : "=r"(a) : "r"(a), "r"(c));
      asm("max.NaN.bf16x2 %0, %1, %2;" : "=r"(m2) : "r"(m2), "r"(a));
    } else {
      asm("max.NaN.f16x2 %0, %1, %2;" : "=r"(a) : "r"(a), "r"(b));
      asm("max.NaN.f16x2 %0, %1, %2;" : "=r"(c) : "r"(c), "r"(d));
      asm("max.NaN.f16x2 %0, %1, %2;" : "=r"(a) : "r"(a), "r"(c));
      asm("max.NaN.f16x2 %0, %1, %2;" : "=r"(m2) : "r"(m2), "r"(a));
    }
  }
  float mx;
  if (TI<T>::ID == LCB_BF16) {
    mx = max_nan(__uint_as_float(m2 << 16), __uint_as_float(m2 & 0xffff0000u));
  } else {
    const float2 t = __half22float2(*reinterpret_cast<const __half2*>(&m2));
    mx = max_nan(t.x, t.y);
  }
  float mn = mx;
  block_minmax(mx, mn, red);
  UnitQ u;
  qparams_from_range(mx, mx, sp, u);
  finish_unit(sp, u);
  if (!finite_f(mx)) u.exact = 1;
  if (threadIdx.x == 0 && scales) store_rt(scales, tile_idx, u.s_out, sp.qparam_dtype);
  const int64_t col = bj * BS + vc * VEC;
  if (u.exact) {   // CTA-uniform, rare (non-finite range): the literal chain
#pragma unroll 1
    for (int p = 0; p < NP; ++p) {
      const int64_t row = bi * BS + p * RPP + vr;
      if (row < rows && col < cols)
        block_vec_slow<T>(*reinterpret_cast<const uint4*>(mine + p * RPP * (BS * 2)), u, sp, out, row * cols + col);
    }
    return;
  }
  const float s = u.s, r = u.r, so = u.s_out;
  if (col >= cols) return;
#pragma unroll 2
  for (int p = 0; p < NP; ++p) {
    const int64_t row = bi * BS + p * RPP + vr;
    if (row >= rows) break;
    const int64_t e0 = row * cols + col;
    float f[VEC];
    TI<T>::unpack(*reinterpret_cast<const uint4*>(mine + p * RPP * (BS * 2)), f);
    uint32_t b2[VEC / 2];
#pragma unroll
    for (int i = 0; i < VEC; i += 2) {
      float q[2];
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const float q0 = __fmul_rn(f[i + k], r);
        q[k] = __fadd_rn(__fmaf_rn(__fmaf_rn(-s, q0, f[i + k]), r, q0), 0.0f);   // RN(x / s), -0 -> +0
      }
      b2[i / 2] = f32x2_to_fp8x2<E5M2>(q[0], q[1]);
    }
    if (OUT_CODES) {
      *reinterpret_cast<uint2*>(reinterpret_cast<uint8_t*>(out) + e0) = make_uint2(b2[0] | (b2[1] << 16), b2[2] | (b2[3] << 16));
    } else {
      float val[VEC];
#pragma unroll
      for (int i = 0; i < VEC; i += 2) {
        const float2 v = fp8x2_to_f32x2<E5M2>(b2[i / 2]);
        val[i] = __fmul_rn(v.x, so);
        val[i + 1] = __fmul_rn(v.y, so);
      }
      st_stream(reinterpret_cast<T*>(out) + e0, TI<T>::pack(val));
    }
  }
}

template <typename T, int E5M2, int OUT_CODES>
__global__ void __launch_bounds__(256, 5)
    fq_block2d_fp8_tma_kernel(const __grid_constant__ CUtensorMap tm, void* __restrict__ out, void* __restrict__ scales,
                              int64_t rows, int64_t cols, SpecD sp) {
  block2d_fp8_tma_tile<T, E5M2, OUT_CODES>(&tm, out, scales, rows, cols, blockIdx.x, sp);
}

// The same staging for every other per_block spec of a 16-bit tensor (integer grids, output_scale_factor != 1, code
// outputs): pass 2 runs the generic per-vector chain (quant_store_vec) from shared memory.
template <typename T, int KINDC>
__global__ void __launch_bounds__(256)
    fq_block2d_tma_kernel(const __grid_constant__ CUtensorMap tm, void* __restrict__ out, void* __restrict__ scales,
                          int64_t rows, int64_t cols, SpecD sp) {
  constexpr int VEC = 8, BS = 128, VPR = BS / VEC, RPP = 256 / VPR, NP = BS / RPP;
  static_assert(sizeof(T) == 2, "16-bit inputs");
  __shared__ __align__(128) uint8_t tile[BS * BS * 2];
  __shared__ __align__(8) uint64_t bar;
  __shared__ float red[64];
  const int64_t ncb = (cols + BS - 1) / BS;
  const int64_t bi = blockIdx.x / ncb, bj = blockIdx.x % ncb;
  const int vr = threadIdx.x / VPR, vc = threadIdx.x % VPR;
  if (threadIdx.x == 0) {
    tc::mbar_init(&bar, 1);
    tc::fence_barrier_init();
    tc::mbar_expect_tx(&bar, BS * BS * 2);
    tc::tma_load_2d(tile, &tm, &bar, (int32_t)(bj * BS), (int32_t)(bi * BS));
  }
  __syncthreads();
  tc::mbar_wait(&bar, 0);
  const uint8_t* mine = tile + vr * (BS * 2) + vc * 16;
  float mx = 0.f;
#pragma unroll
  for (int p = 0; p < NP; ++p) {
    float f[VEC];
    TI<T>::unpack(*reinterpret_cast<const uint4*>(mine + p * RPP * (BS * 2)), f);
#pragma unroll
    for (int i = 0; i < VEC; ++i) mx = max_nan(mx, fabsf(f[i]));
  }
  float mn = mx;
  block_minmax(mx, mn, red);
  UnitQ u;
  qparams_from_range(mx, mx, sp, u);
  finish_unit(sp, u);
  if (!finite_f(mx)) u.exact = 1;
  if (threadIdx.x == 0 && scales) store_rt(scales, blockIdx.x, u.s_out, sp.qparam_dtype);
  if (sp.out_mode == LCB_OUT_NONE) return;
  const int64_t col = bj * BS + vc * VEC;
  if (col >= cols) return;
#pragma unroll 2
  for (int p = 0; p < NP; ++p) {
    const int64_t row = bi * BS + p * RPP + vr;
    if (row >= rows) break;
    float f[VEC];
    TI<T>::unpack(*reinterpret_cast<const uint4*>(mine + p * RPP * (BS * 2)), f);
    quant_store_vec<T, float, KINDC>(f, u, sp, out, row * cols + col);
  }
}

// Several weights in ONE launch (the expert matrices of a MoE block: 2.9 M elements each for DeepSeek-V2-Lite, ~1 us of
// kernel against ~10 us of host work per call). Tiles of all tensors are laid end to end.
constexpr int kMaxBlockMulti = 64;   // 64 x (128 B tensor map + 48 B) = 11 KB of kernel parameters (limit 32 KB)
struct BlockMultiDesc {
  CUtensorMap tm[kMaxBlockMulti];
  void* out[kMaxBlockMulti];
  void* scales[kMaxBlockMulti];
  int64_t rows[kMaxBlockMulti];
  int64_t cols[kMaxBlockMulti];
  int64_t tile_start[kMaxBlockMulti + 1];
  int n;
};
template <typename T, int E5M2, int OUT_CODES>
__global__ void __launch_bounds__(256) fq_block2d_fp8_tma_multi_kernel(const __grid_constant__ BlockMultiDesc md, SpecD sp) {
  int t = 0;   // last tensor whose first tile is <= blockIdx.x (CTA-uniform binary search; empty tensors own no tile)
#pragma unroll
  for (int step = kMaxBlockMulti / 2; step > 0; step >>= 1)
    if (t + step < md.n && (int64_t)blockIdx.x >= md.tile_start[t + step]) t += step;
  block2d_fp8_tma_tile<T, E5M2, OUT_CODES>(&md.tm[t], md.out[t], md.scales[t], md.rows[t], md.cols[t],
                                           (int64_t)blockIdx.x - md.tile_start[t], sp);
}

// ------------------------------------------------------------------------------------------------
// generic: warp per 2-D unit, scalar accesses, literal arithmetic. Any shape, any alignment, any spec.
// ------------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ void store_scalar(void* out, int64_t e, float val, float code, const SpecD& sp) {
  switch (sp.out_mode) {
    case LCB_OUT_DEQUANT:
    case LCB_OUT_FAKE: TI<T>::store1(reinterpret_cast<T*>(out) + e, val); break;
    case LCB_OUT_CODE_T: TI<T>::store1(reinterpret_cast<T*>(out) + e, code); break;
    case LCB_OUT_I8: reinterpret_cast<int8_t*>(out)[e] = (int8_t)(int)code; break;
    case LCB_OUT_U8: reinterpret_cast<uint8_t*>(out)[e] = (uint8_t)(int)code; break;
    case LCB_OUT_I32: reinterpret_cast<int32_t*>(out)[e] = (int)code; break;
    case LCB_OUT_FP8: reinterpret_cast<uint8_t*>(out)[e] = f32_to_fp8_byte(code, sp.kind); break;
    default: break;
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
    fq_generic_kernel(const T* __restrict__ in, void* __restrict__ out, void* __restrict__ scales,
                      void* __restrict__ zeros, int64_t rows, int64_t cols, int64_t ur, int64_t uc, int dynamic,
                      int abs_range, SpecD sp) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const int64_t nru = (rows + ur - 1) / ur, ncu = (cols + uc - 1) / uc;
  sp.fast_ok = 0;
  for (int64_t unit = warp; unit < nru * ncu; unit += nwarps) {
    const int64_t r0 = (unit / ncu) * ur, c0 = (unit % ncu) * uc;
    const int64_t nr = min(ur, rows - r0), nc = min(uc, cols - c0);
    const int64_t n = nr * nc;
    UnitQ u;
    if (dynamic) {
      float mx = -INFINITY, mn = INFINITY;
      for (int64_t i = lane; i < n; i += 32) {
        float x = TI<T>::load1(in + (r0 + i / nc) * cols + c0 + i % nc);
        if (abs_range) x = fabsf(x);
        mx = max_nan(mx, x);
        mn = min_nan(mn, x);
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        mx = max_nan(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        mn = min_nan(mn, __shfl_xor_sync(0xffffffffu, mn, o));
      }
      if (abs_range && (nr < ur || nc < uc)) mn = min_nan(mn, 0.f);  // zero padding (quant.py:637-638)
      qparams_from_range(mn, mx, sp, u);
      finish_unit(sp, u);
      if (lane == 0) write_qparams(scales, zeros, unit, u, sp);
    } else {
      u.s = load_rt(scales, unit, sp.qparam_dtype);
      u.z = (zeros && !sp.sym && sp.kind == LCB_KIND_INT) ? load_rt(zeros, unit, sp.qparam_dtype) : 0.f;
      if (sp.kind != LCB_KIND_INT && u.s == 0.f) u.s = 1.f;
      finish_unit(sp, u);
    }
    if (sp.out_mode == LCB_OUT_NONE) continue;
    for (int64_t i = lane; i < n; i += 32) {
      const int64_t e = (r0 + i / nc) * cols + c0 + i % nc;
      float2 vc = fq_elem_literal(TI<T>::load1(in + e), u, sp);
      store_scalar<T>(out, e, vc.x, vc.y, sp);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// per_tensor: phase 1 = grid min/max, finalize = qparams; phase 2 = fq_static_kernel
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) minmax_partial_kernel(const T* __restrict__ in, int64_t numel, float* ws) {
  constexpr int VEC = TI<T>::VEC;
  __shared__ float red[64];
  const int64_t nvec = numel / VEC;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  float mx = -INFINITY, mn = INFINITY;
  for (int64_t v0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v0 < nvec; v0 += 4 * stride) {
    uint4 raw[4];
#pragma unroll
    for (int k = 0; k < 4; ++k)
      if (v0 + k * stride < nvec) raw[k] = *reinterpret_cast<const uint4*>(in + (v0 + k * stride) * VEC);
#pragma unroll
    for (int k = 0; k < 4; ++k)
      if (v0 + k * stride < nvec) {
        float f[VEC];
        TI<T>::unpack(raw[k], f);
#pragma unroll
        for (int i = 0; i < VEC; ++i) {
          mx = max_nan(mx, f[i]);
          mn = min_nan(mn, f[i]);
        }
      }
  }
  block_minmax(mx, mn, red);
  if (threadIdx.x == 0) {
    ws[2 * blockIdx.x] = mx;
    ws[2 * blockIdx.x + 1] = mn;
  }
}

__global__ void __launch_bounds__(256)
    minmax_finalize_kernel(float* ws, int nparts, void* scales, void* zeros, SpecD sp) {
  __shared__ float red[64];
  float mx = -INFINITY, mn = INFINITY;
  for (int i = threadIdx.x; i < nparts; i += blockDim.x) {
    mx = max_nan(mx, ws[2 * i]);
    mn = min_nan(mn, ws[2 * i + 1]);
  }
  block_minmax(mx, mn, red);
  if (threadIdx.x == 0) {
    UnitQ u;
    qparams_from_range(mn, mx, sp, u);
    finish_unit(sp, u);
    float* q = ws + 2 * nparts;  // phase-2 qparams, fp32 (values already rounded to the math dtype)
    q[0] = u.s;
    q[1] = u.z;
    sp.osf = 1.0f;  // user-visible scale is the un-factored one; phase 2 applies osf itself
    write_qparams(scales, zeros, 0, u, sp);
  }
}

}  // namespace lcb
#include "fakequant_host.cuh"

namespace lcb {
// ---- get_minmax_range (quant.py:132-143): warp per unit -------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
    minmax_units_kernel(const T* __restrict__ in, void* __restrict__ mins, void* __restrict__ maxs, int64_t rows, int64_t cols,
                        int64_t ur, int64_t uc, int range_dtype, int abs_range) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const int64_t nru = (rows + ur - 1) / ur, ncu = (cols + uc - 1) / uc;
  for (int64_t unit = warp; unit < nru * ncu; unit += nwarps) {
    const int64_t r0 = (unit / ncu) * ur, c0 = (unit % ncu) * uc;
    const int64_t nr = min(ur, rows - r0), nc = min(uc, cols - c0);
    float mx = -INFINITY, mn = INFINITY;
    for (int64_t i = lane; i < nr * nc; i += 32) {
      float x = TI<T>::load1(in + (r0 + i / nc) * cols + c0 + i % nc);
      if (abs_range) x = fabsf(x);
      mx = max_nan(mx, x);
      mn = min_nan(mn, x);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      mx = max_nan(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      mn = min_nan(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    }
    if (abs_range && (nr < ur || nc < uc)) mn = min_nan(mn, 0.f);  // zero padding (quant.py:637-638)
    if (lane == 0) {
      store_rt(mins, unit, mn, range_dtype);
      store_rt(maxs, unit, mx, range_dtype);
    }
  }
}

// whole-tensor range, second stage: reduces the per-CTA (max, min) pairs of minmax_partial_kernel
__global__ void __launch_bounds__(256) minmax_finalize_range_kernel(const float* ws, int nparts, void* mins, void* maxs, int range_dtype) {
  __shared__ float red[64];
  float mx = -INFINITY, mn = INFINITY;
  for (int i = threadIdx.x; i < nparts; i += blockDim.x) {
    mx = max_nan(mx, ws[2 * i]);
    mn = min_nan(mn, ws[2 * i + 1]);
  }
  block_minmax(mx, mn, red);
  if (threadIdx.x == 0) {
    store_rt(mins, 0, mn, range_dtype);
    store_rt(maxs, 0, mx, range_dtype);
  }
}

// ---- get_qparams (quant.py:545-559) on given ranges: thread per unit, literal arithmetic --------------------------------
__global__ void __launch_bounds__(256)
    qparams_from_range_kernel(const void* __restrict__ mins, const void* __restrict__ maxs, void* __restrict__ scales,
                              void* __restrict__ zeros, int64_t n, SpecD sp) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  sp.fast_ok = 0;
  UnitQ u;
  qparams_from_range(load_rt(mins, i, sp.math_dtype), load_rt(maxs, i, sp.math_dtype), sp, u);
  finish_unit(sp, u);
  write_qparams(scales, zeros, i, u, sp);
}
}  // namespace lcb

template <typename T>
static int minmax_whole(const void* in, void* mins, void* maxs, int64_t numel, int range_dtype, float* ws, cudaStream_t st) {
  using namespace lcb;
  const int parts = grid_for((numel / TI<T>::VEC + 1023) / 1024, 8);
  minmax_partial_kernel<T><<<parts, 256, 0, st>>>((const T*)in, numel, ws);
  minmax_finalize_range_kernel<<<1, 256, 0, st>>>(ws, parts, mins, maxs, range_dtype);
  LCB_CUDA_CHECK_LAUNCH("minmax_whole");
  return LCB_OK;
}

extern "C" int lcb_minmax(const void* in, void* mins, void* maxs, int64_t rows, int64_t cols, int64_t unit_rows,
                          int64_t unit_cols, int32_t dtype, int32_t range_dtype, int32_t abs_range, void* workspace,
                          void* stream) {
  using namespace lcb;
  if (!in || !mins || !maxs || rows < 0 || cols < 0 || unit_rows < 1 || unit_cols < 1) return LCB_E_INVALID;
  if (dtype < LCB_F32 || dtype > LCB_BF16 || range_dtype < LCB_F32 || range_dtype > LCB_BF16) return LCB_E_UNSUPPORTED;
  if (rows == 0 || cols == 0) return LCB_OK;
  if (unit_rows == 1 && unit_cols < cols && cols % unit_cols != 0) return LCB_E_SHAPE;
  const int64_t ur = unit_rows > rows ? rows : unit_rows, uc = unit_cols > cols ? cols : unit_cols;
  const int64_t units = ((rows + ur - 1) / ur) * ((cols + uc - 1) / uc);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int64_t numel = rows * cols;
  const int vec = dtype == LCB_F32 ? 4 : 8;
  if (units == 1 && !abs_range && workspace && numel >= 4096 && numel % vec == 0 && aligned16(in)) {   // per_tensor
    float* ws = reinterpret_cast<float*>(workspace);
    if (dtype == LCB_F32) return minmax_whole<float>(in, mins, maxs, numel, range_dtype, ws, st);
    if (dtype == LCB_F16) return minmax_whole<__half>(in, mins, maxs, numel, range_dtype, ws, st);
    return minmax_whole<__nv_bfloat16>(in, mins, maxs, numel, range_dtype, ws, st);
  }
  const int grid = grid_for((units + 7) / 8, 8);
  if (dtype == LCB_F32)
    minmax_units_kernel<float><<<grid, 256, 0, st>>>((const float*)in, mins, maxs, rows, cols, ur, uc, range_dtype, abs_range);
  else if (dtype == LCB_F16)
    minmax_units_kernel<__half><<<grid, 256, 0, st>>>((const __half*)in, mins, maxs, rows, cols, ur, uc, range_dtype, abs_range);
  else
    minmax_units_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)in, mins, maxs, rows, cols, ur, uc, range_dtype,
                                                             abs_range);
  LCB_CUDA_CHECK_LAUNCH("minmax_units");
  return LCB_OK;
}

extern "C" int lcb_qparams_from_range(const void* mins, const void* maxs, void* scales, void* zeros, int64_t n,
                                      const lcb_quant_spec* spec, void* stream) {
  using namespace lcb;
  if (!mins || !maxs || !scales || !spec || n < 0) return LCB_E_INVALID;
  if (spec->kind < LCB_KIND_INT || spec->kind > LCB_KIND_FP8_E5M2) return LCB_E_UNSUPPORTED;
  if (!(spec->qmax > spec->qmin)) return LCB_E_INVALID;
  if (n == 0) return LCB_OK;
  SpecD sp = make_specd(spec);
  sp.out_mode = LCB_OUT_NONE;
  sp.clip = 1.0f;
  qparams_from_range_kernel<<<(int)((n + 255) / 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(mins, maxs, scales, zeros, n, sp);
  LCB_CUDA_CHECK_LAUNCH("qparams_from_range");
  return LCB_OK;
}
