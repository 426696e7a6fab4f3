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
#include "tc.cuh"

namespace lcb {

namespace {
constexpr int BM = 128, BN = 256, BK = 64, STAGES = 4;
constexpr int A_BYTES = BM * BK * 2, B_BYTES = BN * BK * 2, STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int BOX_BYTES = 64 * BK * 2;              // one TMA box: 64 k-rows x 64 elements
constexpr int EPI_BUF = 32 * 32 * 4;                // 32 rows x 32 fp32 columns
constexpr int SMEM_MAIN = STAGES * STAGE_BYTES;     // 196608
constexpr int SMEM_EPI = 4 * 2 * EPI_BUF;           // 32768
constexpr int SMEM_BYTES = SMEM_MAIN + SMEM_EPI + 256 + 1024;  // + barriers + alignment slack
constexpr int NTHREADS = 192;
constexpr int MAX_SEG = 3;

struct HessParams {
  int C, T;
  int nblk_i;        // 128-row blocks
  int nblk_j;        // 256-column blocks
  int n_tiles;       // upper-triangle tiles
  int nsplit;        // K splits per tile
  int kb_per_split;  // 64-token blocks per split (per segment)
  int n_seg;         // 1 (16-bit input) or 3 (fp32 split)
  uint32_t idesc;
};

struct HessMaps {
  CUtensorMap a[MAX_SEG];
  CUtensorMap b[MAX_SEG];
  CUtensorMap s;
};

// item -> (row block i, column block j): row block i owns column blocks j = i/2 .. nblk_j-1
__device__ __forceinline__ void tile_of(int t, int nblk_i, int nblk_j, int& ti, int& tj) {
  int i = 0;
  while (true) {
    const int cnt = nblk_j - (i >> 1);
    if (t < cnt) break;
    t -= cnt;
    ++i;
  }
  ti = i;
  tj = (i >> 1) + t;
}

__global__ void __launch_bounds__(NTHREADS, 1)
    hessian_syrk_kernel(const __grid_constant__ HessMaps maps, const HessParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* epi = smem + SMEM_MAIN;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + SMEM_MAIN + SMEM_EPI);
  uint64_t* empty = full + STAGES;
  uint64_t* tfull = empty + STAGES;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    for (int s = 0; s < p.n_seg; ++s) {
      tc::tma_prefetch_desc(&maps.a[s]);
      tc::tma_prefetch_desc(&maps.b[s]);
    }
    tc::tma_prefetch_desc(&maps.s);
    for (int s = 0; s < STAGES; ++s) {
      tc::mbar_init(&full[s], 1);
      tc::mbar_init(&empty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      tc::mbar_init(&tfull[a], 1);
      tc::mbar_init(&tempty[a], 4);
    }
    tc::fence_barrier_init();
  }
  if (warp == 1) tc::tmem_alloc(tmem_slot, 512);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int n_items = p.n_tiles * p.nsplit;
  const int kb_total = (p.T + BK - 1) / BK;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        int ti, tj;
        tile_of(item / p.nsplit, p.nblk_i, p.nblk_j, ti, tj);
        const int split = item % p.nsplit;
        const int kb0 = split * p.kb_per_split;
        const int kb1 = min(kb_total, kb0 + p.kb_per_split);
        for (int seg = 0; seg < p.n_seg; ++seg) {
          for (int kb = kb0; kb < kb1; ++kb) {
            tc::mbar_wait(&empty[stage], phase ^ 1);
            uint8_t* sa = smem + stage * STAGE_BYTES;
            uint8_t* sb = sa + A_BYTES;
            tc::mbar_expect_tx(&full[stage], STAGE_BYTES);
#pragma unroll
            for (int b = 0; b < BM / 64; ++b)
              tc::tma_load_2d(sa + b * BOX_BYTES, &maps.a[seg], &full[stage], ti * BM + b * 64, kb * BK);
#pragma unroll
            for (int b = 0; b < BN / 64; ++b)
              tc::tma_load_2d(sb + b * BOX_BYTES, &maps.b[seg], &full[stage], tj * BN + b * 64, kb * BK);
            if (++stage == STAGES) {
              stage = 0;
              phase ^= 1;
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    // MN-major SWIZZLE_128B: LBO = bytes between 64-element chunks (one TMA box), SBO = bytes between 8-k groups
    const uint64_t desc_base = tc::smem_desc_sw128(BOX_BYTES, 1024);
    for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
      const int split = item % p.nsplit;
      const int kb0 = split * p.kb_per_split;
      const int kb1 = min(kb_total, kb0 + p.kb_per_split);
      const int n_kb = (kb1 - kb0) * p.n_seg;
      tc::mbar_wait(&tempty[acc], acc_phase ^ 1);
      tc::tc_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)acc * BN;
      for (int it = 0; it < n_kb; ++it) {
        tc::mbar_wait(&full[stage], phase);
        tc::tc_fence_after();
        if (lane == 0) {
          const uint32_t sa = tc::smem_u32(smem + stage * STAGE_BYTES);
          const uint32_t sb = sa + A_BYTES;
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            // 16 k-rows = 2 swizzle atoms of 1024 B
            const uint64_t ad = tc::smem_desc_at(desc_base, sa + k * 2048);
            const uint64_t bd = tc::smem_desc_at(desc_base, sb + k * 2048);
            tc::mma_f16_ss(d_tmem, ad, bd, p.idesc, (it | k) != 0 ? 1u : 0u);
          }
          tc::mma_commit(&empty[stage]);                     // frees the smem slot once these MMAs have read it
          if (it == n_kb - 1) tc::mma_commit(&tfull[acc]);   // accumulator complete
        }
        __syncwarp();
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1;
        }
      }
      if (n_kb == 0 && lane == 0) tc::mbar_arrive(&tfull[acc]);  // cannot happen (host sizes splits), keeps the protocol total
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
  } else {
    // ===== epilogue warps: TMEM -> registers -> swizzled smem -> TMA reduce-add =====
    const int q = warp & 3;  // TMEM lane quarter this warp may access
    uint8_t* my_epi = epi + (warp - 2) * 2 * EPI_BUF;
    int acc = 0;
    uint32_t acc_phase = 0;
    int buf = 0;
    for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
      int ti, tj;
      tile_of(item / p.nsplit, p.nblk_i, p.nblk_j, ti, tj);
      tc::mbar_wait(&tfull[acc], acc_phase);
      tc::tc_fence_after();
      const int row0 = ti * BM + q * 32;
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        const int col0 = tj * BN + c * 32;
        const int jb = col0 >> 7;  // 128-column block of this chunk
        if (jb < ti || col0 >= p.C || row0 >= p.C) continue;  // warp-uniform: lower triangle / out of range
        uint32_t v[32];
        tc::tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BN + c * 32), v);
        tc::tmem_ld_wait();
        // staging buffer `buf` was handed to TMA two chunks ago: wait until that read is done
        if (lane == 0) tc::bulk_wait_read<1>();
        __syncwarp();
        uint8_t* dst = my_epi + buf * EPI_BUF + lane * 128;
#pragma unroll
        for (int ch = 0; ch < 8; ++ch) {
          const int pch = ch ^ (lane & 7);  // 128B swizzle: 16-byte chunk index xor (row mod 8)
          *reinterpret_cast<uint4*>(dst + pch * 16) = make_uint4(v[4 * ch], v[4 * ch + 1], v[4 * ch + 2], v[4 * ch + 3]);
        }
        tc::fence_async_smem();
        __syncwarp();
        if (lane == 0) {
          tc::tma_reduce_add_2d(&maps.s, my_epi + buf * EPI_BUF, col0, row0);
          tc::bulk_commit();
        }
        buf ^= 1;
      }
      tc::tc_fence_before();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&tempty[acc]);
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
    if (lane == 0) tc::bulk_wait_all();
  }

  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc::tc_fence_after();
    tc::tmem_dealloc(tmem_base, 512);
  }
}

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

static bool g_attr_set = false;

}  // namespace

// ---- tensor-map helpers (declared in tc.cuh) ----------------------------------------------------------------------------
lcb_encode_tiled_fn get_encode_tiled() {
  static lcb_encode_tiled_fn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<lcb_encode_tiled_fn>(p);
  }
  return fn;
}

bool make_tmap_2d(CUtensorMap* m, const void* base, CUtensorMapDataType dt, int elem_bytes, uint64_t rows, uint64_t cols,
                  uint64_t row_stride_bytes, uint32_t box_rows, uint32_t box_cols) {
  lcb_encode_tiled_fn enc = get_encode_tiled();
  if (!enc) return false;
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {row_stride_bytes};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  (void)elem_bytes;
  return enc(m, dt, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
             CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

}  // namespace lcb

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
  HessMaps maps;
  HessParams p;
  p.C = (int)C;
  p.T = (int)tokens;
  p.nblk_i = (int)((C + BM - 1) / BM);
  p.nblk_j = (int)((C + BN - 1) / BN);
  p.n_tiles = 0;
  for (int i = 0; i < p.nblk_i; ++i) p.n_tiles += p.nblk_j - (i >> 1);
  const int kb_total = (int)((tokens + BK - 1) / BK);
  // split K until the grid is filled about twice, keeping >= 8 k-blocks (512 tokens) per split
  int nsplit = 1;
  while (p.n_tiles * nsplit < 2 * kNumSMs && kb_total / (nsplit * 2) >= 8) nsplit *= 2;
  p.kb_per_split = (kb_total + nsplit - 1) / nsplit;
  p.nsplit = (kb_total + p.kb_per_split - 1) / p.kb_per_split;
  p.n_seg = 1;
  const int fmt = x_dtype == LCB_F16 ? 0 : 1;
  p.idesc = tc::idesc_f16(fmt, BM, BN, 1, 1);

  const CUtensorMapDataType dt16 = x_dtype == LCB_F16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  bool ok = true;
  if (x_dtype == LCB_F32) {
    if (!workspace) return LCB_E_INVALID;
    __nv_bfloat16* hi = reinterpret_cast<__nv_bfloat16*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~uintptr_t(255));
    __nv_bfloat16* lo = hi + tokens * C;
    const int64_t n = tokens * C;
    split_f32_kernel<<<(int)std::min<int64_t>((n + 255) / 256, kNumSMs * 16), 256, 0, st>>>((const float*)X, hi, lo, n);
    LCB_CUDA_CHECK_LAUNCH("hessian_split");
    p.n_seg = 3;  // hi*hi + hi*lo + lo*hi
    const __nv_bfloat16* aa[3] = {hi, hi, lo};
    const __nv_bfloat16* bb[3] = {hi, lo, hi};
    for (int s = 0; s < 3; ++s) {
      ok = ok && make_tmap_2d(&maps.a[s], aa[s], dt16, 2, tokens, C, C * 2, BK, 64);
      ok = ok && make_tmap_2d(&maps.b[s], bb[s], dt16, 2, tokens, C, C * 2, BK, 64);
    }
  } else {
    ok = ok && make_tmap_2d(&maps.a[0], X, dt16, 2, tokens, C, C * 2, BK, 64);
    maps.b[0] = maps.a[0];
    for (int s = 1; s < MAX_SEG; ++s) {
      maps.a[s] = maps.a[0];
      maps.b[s] = maps.a[0];
    }
  }
  ok = ok && make_tmap_2d(&maps.s, S, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, C, C, C * 4, 32, 32);
  if (!ok) {
    set_last_error("cuTensorMapEncodeTiled", cudaErrorInvalidValue);
    return LCB_E_CUDA;
  }
  if (!g_attr_set) {
    if (cudaFuncSetAttribute(hessian_syrk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES) != cudaSuccess) {
      set_last_error("cudaFuncSetAttribute(hessian)", cudaGetLastError());
      return LCB_E_CUDA;
    }
    g_attr_set = true;
  }
  const int n_items = p.n_tiles * p.nsplit;
  const int grid = n_items < kNumSMs ? n_items : kNumSMs;
  hessian_syrk_kernel<<<grid, NTHREADS, SMEM_BYTES, st>>>(maps, p);
  LCB_CUDA_CHECK_LAUNCH("hessian_syrk");
  return LCB_OK;
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
