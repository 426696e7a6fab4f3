// tc_gemm.cuh — the tcgen05 contraction shared by the Hessian (S += X^T X) and the GPTQ lazy-batch updates
// (W -= Err * Hinv):      D[M, N] (+/-)= sum_seg A_seg[M, K] * B_seg[K, N]      fp32 accumulate in TMEM, fp32 D.
//
// One CTA per SM, persistent, warp-specialised (192 threads):
//   warp 0   : TMA producer, 4-stage ring of (A 128 x 64, B 64 x 256) 16-bit tiles, 128-byte swizzle (48 KB / stage).
//              B is always [K, N] row-major (MN-major operand). A is either [K, M] row-major (A_KMAJOR = false: the
//              Hessian's X^T, MN-major) or [M, K] row-major (A_KMAJOR = true: the GPTQ Err block, K-major).
//   warp 1   : MMA issuer: one thread, tcgen05.mma.cta_group::1.kind::f16, M = 128, N = 256, K = 16 per instruction,
//              two 256-column TMEM accumulators so the epilogue of tile t overlaps the main loop of tile t + 1.
//   warps 2-5: epilogue: tcgen05.ld (32 lanes x 32 columns) -> 128B-swizzled smem -> TMA reduce-add into D. The L2
//              performs the read-modify-write, so split-K needs no workspace and no second pass.
// Segments let an fp32 operand enter as bf16 hi + lo: (Ah,Bh) + (Ah,Bl) + (Al,Bh) run as three K ranges into the
// same accumulator. Out-of-range rows / columns are zero-filled on load and clipped on store by TMA.
// SYRK = true: only tiles touching the upper triangle (128-row block i, 128-column block jb >= i) are computed/stored.
#pragma once
#include <algorithm>
#include "tc.cuh"

namespace lcb {
namespace tcg {

constexpr int BM = 128, BN = 256, BK = 64, STAGES = 4;
constexpr int A_BYTES = BM * BK * 2, B_BYTES = BN * BK * 2, STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int BOX_BYTES = 64 * BK * 2;              // one MN-major TMA box: 64 k-rows x 64 elements
constexpr int EPI_BUF = 32 * 32 * 4;                // 32 rows x 32 fp32 columns
constexpr int SMEM_MAIN = STAGES * STAGE_BYTES;     // 196608
constexpr int SMEM_EPI = 4 * 2 * EPI_BUF;           // 32768
constexpr int SMEM_BYTES = SMEM_MAIN + SMEM_EPI + 256 + 1024;  // + barriers + alignment slack
constexpr int NTHREADS = 192;
constexpr int MAX_SEG = 6;

// EPI_STORE16: org[g][row][col] = rd16(acc) (no D). EPI_ROWSSE: rowsse[row * n_grp + g] += sum_col rd16(rd16(rd16(acc) -
// org[g][row][col])^2) — the per-(output channel, group) squared error of AutoClip (auto_clip.py:150-156).
enum { EPI_REDUCE_ADD = 0, EPI_MSE = 1, EPI_STORE16 = 2, EPI_ROWSSE = 3 };

struct Params {
  int M, N, K;
  int nblk_i;        // 128-row blocks of D
  int nblk_j;        // 256-column blocks of D
  int n_tiles;
  int nsplit;        // K splits per tile
  int kb_per_split;  // 64-deep K blocks per split (per segment)
  int n_seg;
  int n_grp;         // grouped mode (n_grp > 1): group g contracts K blocks [g * kb_grp, (g + 1) * kb_grp) only — one
  int kb_grp;        // independent [M, N] product per group of input channels (AutoClip's (x * w).sum(-1))
  int64_t org_grp_stride;  // elements between consecutive groups of org
  double* rowsse;
  int tri_b;         // B[k, n] == 0 for k > n (upper-triangular B): column block tj only needs k < (tj + 1) * BN
  uint32_t idesc;
  // EPI_MSE: loss[0] += sum over the tile of rd16(rd16(org) - rd16(acc))^2   (awq.py:134-145 on 16-bit tensors)
  const void* org;   // [M, N] row-major 16-bit, leading dimension ld_org
  int64_t ld_org;
  int org_fmt;       // 0 = fp16, 1 = bf16
  double* loss;
  const void* bias;  // EPI_STORE16: optional [N] bias of the output dtype, added in fp32 before the one rounding
};

struct Maps {
  CUtensorMap a[MAX_SEG];
  CUtensorMap b[MAX_SEG];
  CUtensorMap d;
};

// item -> (row block i, column block j). SYRK: row block i owns column blocks j = i/2 .. nblk_j-1
template <bool SYRK>
__device__ __forceinline__ void tile_of(int t, int nblk_j, int& ti, int& tj) {
  if (!SYRK) {
    ti = t / nblk_j;
    tj = t % nblk_j;
    return;
  }
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

__device__ __forceinline__ int k_hi(const Params& p, int tj, int kb_total) {
  return p.tri_b ? min(kb_total, ((tj + 1) * BN + BK - 1) / BK) : kb_total;
}

struct Item {
  int ti, tj, g, kb0, kb1;   // tile, group, K-block range [kb0, kb1) (empty: the item is skipped by every role)
};
template <bool SYRK>
__device__ __forceinline__ Item decode_item(const Params& p, int item, int kb_total) {
  Item it;
  const int per_grp = p.n_tiles * p.nsplit;
  it.g = item / per_grp;
  const int r = item - it.g * per_grp;
  tile_of<SYRK>(r / p.nsplit, p.nblk_j, it.ti, it.tj);
  const int split = r % p.nsplit;
  it.kb0 = it.g * p.kb_grp + split * p.kb_per_split;
  it.kb1 = min(k_hi(p, it.tj, kb_total), it.kb0 + p.kb_per_split);
  return it;
}

template <int FMT>
__device__ __forceinline__ float rd16(float x) {
  if (FMT == 1) {
    uint32_t d;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(0.0f), "f"(x));
    return __uint_as_float(d << 16);
  }
  return __half2float(__float2half_rn(x));
}
template <int FMT>
__device__ __forceinline__ float2 unpack16x2(uint32_t w) {
  if (FMT == 1) return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u));
  return __half22float2(*reinterpret_cast<const __half2*>(&w));
}
// sum over 32 columns of rd16(org - rd16(acc))^2 for one row
template <int FMT>
__device__ __forceinline__ float mse_row32(const uint32_t (&v)[32], const uint8_t* org_row, int ncols) {
  float sum = 0.f;
#pragma unroll
  for (int c8 = 0; c8 < 4; ++c8) {
    if (c8 * 8 < ncols) {
      const uint4 o = *reinterpret_cast<const uint4*>(org_row + c8 * 16);
      const uint32_t ow[4] = {o.x, o.y, o.z, o.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 og = unpack16x2<FMT>(ow[j]);
        const float d0 = rd16<FMT>(__fsub_rn(og.x, rd16<FMT>(__uint_as_float(v[c8 * 8 + 2 * j]))));
        const float d1 = rd16<FMT>(__fsub_rn(og.y, rd16<FMT>(__uint_as_float(v[c8 * 8 + 2 * j + 1]))));
        sum = __fmaf_rn(d0, d0, sum);
        sum = __fmaf_rn(d1, d1, sum);
      }
    }
  }
  return sum;
}

// AutoClip: sum over 32 columns of rd16(d * d), d = rd16(rd16(acc) - org) (auto_clip.py:150: (cur_out - org_out).pow(2),
// every op rounded to the tensor dtype; the mean over tokens accumulates these in fp32)
template <int FMT>
__device__ __forceinline__ float sse2_row32(const uint32_t (&v)[32], const uint8_t* org_row, int ncols) {
  float sum = 0.f;
#pragma unroll
  for (int c8 = 0; c8 < 4; ++c8) {
    if (c8 * 8 < ncols) {
      const uint4 o = *reinterpret_cast<const uint4*>(org_row + c8 * 16);
      const uint32_t ow[4] = {o.x, o.y, o.z, o.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 og = unpack16x2<FMT>(ow[j]);
        const float d0 = rd16<FMT>(__fsub_rn(rd16<FMT>(__uint_as_float(v[c8 * 8 + 2 * j])), og.x));
        const float d1 = rd16<FMT>(__fsub_rn(rd16<FMT>(__uint_as_float(v[c8 * 8 + 2 * j + 1])), og.y));
        sum = __fadd_rn(sum, rd16<FMT>(__fmul_rn(d0, d0)));
        sum = __fadd_rn(sum, rd16<FMT>(__fmul_rn(d1, d1)));
      }
    }
  }
  return sum;
}
template <int FMT>
__device__ __forceinline__ void store16_row32(const uint32_t (&v)[32], uint8_t* out_row, int ncols) {
#pragma unroll
  for (int c8 = 0; c8 < 4; ++c8) {
    if (c8 * 8 < ncols) {
      uint32_t w[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float a = __uint_as_float(v[c8 * 8 + 2 * j]), b = __uint_as_float(v[c8 * 8 + 2 * j + 1]);
        if (FMT == 1) {
          asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(w[j]) : "f"(b), "f"(a));
        } else {
          const __half2 h = __floats2half2_rn(a, b);
          w[j] = *reinterpret_cast<const uint32_t*>(&h);
        }
      }
      *reinterpret_cast<uint4*>(out_row + c8 * 16) = make_uint4(w[0], w[1], w[2], w[3]);
    }
  }
}

template <bool A_KMAJOR, bool B_KMAJOR, bool SYRK, bool NEG, int EPI>
__global__ void __launch_bounds__(NTHREADS, 1) gemm_kernel(const __grid_constant__ Maps maps, const Params p) {
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
    tc::tma_prefetch_desc(&maps.d);
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

  const int n_items = p.n_tiles * p.nsplit * p.n_grp;
  const int kb_total = (p.K + BK - 1) / BK;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        const Item it = decode_item<SYRK>(p, item, kb_total);
        const int ti = it.ti, tj = it.tj, kb0 = it.kb0, kb1 = it.kb1;
        for (int seg = 0; seg < p.n_seg; ++seg) {
          for (int kb = kb0; kb < kb1; ++kb) {
            tc::mbar_wait(&empty[stage], phase ^ 1);
            uint8_t* sa = smem + stage * STAGE_BYTES;
            uint8_t* sb = sa + A_BYTES;
            tc::mbar_expect_tx(&full[stage], STAGE_BYTES);
            if (A_KMAJOR) {
              tc::tma_load_2d(sa, &maps.a[seg], &full[stage], kb * BK, ti * BM);  // box: 128 m-rows x 64 k
            } else {
#pragma unroll
              for (int b = 0; b < BM / 64; ++b)
                tc::tma_load_2d(sa + b * BOX_BYTES, &maps.a[seg], &full[stage], ti * BM + b * 64, kb * BK);
            }
            if (B_KMAJOR) {
              tc::tma_load_2d(sb, &maps.b[seg], &full[stage], kb * BK, tj * BN);  // box: 256 n-rows x 64 k
            } else {
#pragma unroll
              for (int b = 0; b < BN / 64; ++b)
                tc::tma_load_2d(sb + b * BOX_BYTES, &maps.b[seg], &full[stage], tj * BN + b * 64, kb * BK);
            }
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
    // MN-major SWIZZLE_128B: LBO = bytes between 64-element chunks (one TMA box), SBO = bytes between 8-k groups.
    // K-major SWIZZLE_128B : SBO = bytes between 8-row groups (1024), LBO unused; a K = 16 step is +32 B in the row.
    const uint64_t desc_mn = tc::smem_desc_sw128(BOX_BYTES, 1024);
    const uint64_t desc_k = tc::smem_desc_sw128(0, 1024);
    for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
      const Item it = decode_item<SYRK>(p, item, kb_total);
      const int n_kb = (it.kb1 - it.kb0) * p.n_seg;
      if (n_kb <= 0) continue;   // tri_b: this K split lies entirely below the triangle (all roles skip it alike)
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
            const uint64_t ad = A_KMAJOR ? tc::smem_desc_at(desc_k, sa + k * 32) : tc::smem_desc_at(desc_mn, sa + k * 2048);
            // MN-major: 16 k-rows = 2 swizzle atoms of 1024 B
            const uint64_t bd = B_KMAJOR ? tc::smem_desc_at(desc_k, sb + k * 32) : tc::smem_desc_at(desc_mn, sb + k * 2048);
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
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
  } else {
    // ===== epilogue warps =====
    const int q = warp & 3;  // TMEM lane quarter this warp may access
    uint8_t* my_epi = epi + (warp - 2) * 2 * EPI_BUF;
    int acc = 0;
    uint32_t acc_phase = 0;
    int buf = 0;
    double mse_sum = 0.0;
    for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
      const Item it = decode_item<SYRK>(p, item, kb_total);
      const int ti = it.ti, tj = it.tj;
      if (it.kb1 <= it.kb0) continue;
      tc::mbar_wait(&tfull[acc], acc_phase);
      tc::tc_fence_after();
      const int row0 = ti * BM + q * 32;
      float tile_sum = 0.f;
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        const int col0 = tj * BN + c * 32;
        if (SYRK && (col0 >> 7) < ti) continue;            // warp-uniform: 128-column block below the diagonal
        if (col0 >= p.N || row0 >= p.M) continue;          // warp-uniform: out of range
        uint32_t v[32];
        tc::tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BN + c * 32), v);
        tc::tmem_ld_wait();
        if (EPI == EPI_MSE || EPI == EPI_ROWSSE || EPI == EPI_STORE16) {
          // TMEM -> registers -> round to the 16-bit output dtype -> squared error against org_out (or store)
          const int row = row0 + lane;
          if (row < p.M) {
            uint8_t* org_row = reinterpret_cast<uint8_t*>(const_cast<void*>(p.org)) +
                               ((int64_t)it.g * p.org_grp_stride + (int64_t)row * p.ld_org + col0) * 2;
            const int ncols = min(32, p.N - col0);
            if (EPI == EPI_MSE) tile_sum += p.org_fmt == 1 ? mse_row32<1>(v, org_row, ncols) : mse_row32<0>(v, org_row, ncols);
            if (EPI == EPI_ROWSSE) tile_sum += p.org_fmt == 1 ? sse2_row32<1>(v, org_row, ncols) : sse2_row32<0>(v, org_row, ncols);
            if (EPI == EPI_STORE16) {
              if (p.bias) {
                const uint8_t* bp = reinterpret_cast<const uint8_t*>(p.bias) + (int64_t)col0 * 2;
#pragma unroll
                for (int c2 = 0; c2 < 16; ++c2) {
                  if (2 * c2 < ncols) {
                    const float2 b2 = p.org_fmt == 1 ? unpack16x2<1>(*reinterpret_cast<const uint32_t*>(bp + c2 * 4))
                                                     : unpack16x2<0>(*reinterpret_cast<const uint32_t*>(bp + c2 * 4));
                    v[2 * c2] = __float_as_uint(__fadd_rn(__uint_as_float(v[2 * c2]), b2.x));
                    v[2 * c2 + 1] = __float_as_uint(__fadd_rn(__uint_as_float(v[2 * c2 + 1]), b2.y));
                  }
                }
              }
              if (p.org_fmt == 1) store16_row32<1>(v, org_row, ncols); else store16_row32<0>(v, org_row, ncols);
            }
          }
        } else {
          // TMEM -> registers -> 128B-swizzled smem -> TMA reduce-add into D.
          // staging buffer `buf` was handed to TMA two chunks ago: wait until that read is done
          if (lane == 0) tc::bulk_wait_read<1>();
          __syncwarp();
          uint8_t* dst = my_epi + buf * EPI_BUF + lane * 128;
#pragma unroll
          for (int ch = 0; ch < 8; ++ch) {
            const int pch = ch ^ (lane & 7);  // 128B swizzle: 16-byte chunk index xor (row mod 8)
            const uint32_t sg = NEG ? 0x80000000u : 0u;
            *reinterpret_cast<uint4*>(dst + pch * 16) =
                make_uint4(v[4 * ch] ^ sg, v[4 * ch + 1] ^ sg, v[4 * ch + 2] ^ sg, v[4 * ch + 3] ^ sg);
          }
          tc::fence_async_smem();
          __syncwarp();
          if (lane == 0) {
            tc::tma_reduce_add_2d(&maps.d, my_epi + buf * EPI_BUF, col0, row0);
            tc::bulk_commit();
          }
          buf ^= 1;
        }
      }
      if (EPI == EPI_ROWSSE) {   // lane = output channel: one atomic per (row, group, tile)
        if (row0 + lane < p.M) atomicAdd(p.rowsse + (int64_t)(row0 + lane) * p.n_grp + it.g, (double)tile_sum);
      } else {
        mse_sum += (double)tile_sum;
      }
      tc::tc_fence_before();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&tempty[acc]);
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
    if (EPI == EPI_MSE) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) mse_sum += __shfl_xor_sync(0xffffffffu, mse_sum, o);
      if (lane == 0 && mse_sum != 0.0) atomicAdd(p.loss, mse_sum);
    } else if (EPI == EPI_REDUCE_ADD) {
      if (lane == 0) tc::bulk_wait_all();
    }
  }

  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc::tc_fence_after();
    tc::tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace tcg

// Host launcher (defined in tc_gemm.cu). 16-bit operands; a[s] / b[s] are the segment base pointers.
//   a_kmajor: A_seg is [M, K] row-major with leading dimension lda (elements); else [K, M] row-major.
//   B_seg is [K, N] row-major, leading dimension ldb. D is fp32 [M, N], leading dimension ldd.
//   fmt: 0 = fp16, 1 = bf16. syrk requires M == N. neg: D -= instead of D +=.
int tc_gemm_launch(bool a_kmajor, bool syrk, bool neg, int fmt, const void* const* a, const void* const* b, int n_seg,
                   float* D, int64_t M, int64_t N, int64_t K, int64_t lda, int64_t ldb, int64_t ldd, cudaStream_t st);
// Same with: b_kmajor (B_seg given as [N, K] row-major), tri_b (B is upper triangular, B[k, n] = 0 for k > n: the K loop
// of a column block stops at its last column; requires !b_kmajor).
int tc_gemm_launch_ex(bool a_kmajor, bool b_kmajor, bool syrk, bool neg, bool tri_b, int fmt, const void* const* a,
                      const void* const* b, int n_seg, float* D, int64_t M, int64_t N, int64_t K, int64_t lda, int64_t ldb,
                      int64_t ldd, cudaStream_t st);

// loss[0] += sum_{t, r} rd16(org[t, r] - rd16(sum_k x[t, k] * w[r, k]))^2 — F.linear + the squared-error part of
// Awq.calculate_loss (awq.py:134-145) without writing the output. x [M, K], w [N, K], org [M, N], all 16-bit row-major.
int tc_linear_mse_launch(int fmt, const void* x, const void* w, const void* org, double* loss, int64_t M, int64_t N,
                         int64_t K, int64_t ldx, int64_t ldw, int64_t ldo, cudaStream_t st);

// out[t, r] = rd16(sum_k x[t, k] * w[r, k] (+ bias[r])) — F.linear on 16-bit tensors (the projections inside the inspect
// modules of Awq.search_scale_subset, awq.py:110-126). x [M, K], w [N, K], out [M, N] row-major; bias NULL or [N].
int tc_linear_launch(int fmt, const void* x, const void* w, const void* bias, void* out, int64_t M, int64_t N, int64_t K,
                     int64_t ldx, int64_t ldw, int64_t ldo, cudaStream_t st);

// Grouped product for AutoClip (auto_clip.py:133-156): for every group g of `gs` input channels,
//   out_g[r, t] = sum_{c in g} w[r, c] * x[t, c]          w [R, C], x [T, C] 16-bit row-major, C = n_grp * gs, gs % 64 == 0
// store != 0: org[g][r][t] = rd16(out_g)   (org is [n_grp, R, T], T % 8 == 0)
// store == 0: rowsse[r * n_grp + g] += sum_t rd16(rd16(rd16(out_g[r, t]) - org[g][r][t])^2)
int tc_grouped_rowsse_launch(int fmt, int store, const void* w, const void* x, void* org, double* rowsse, int64_t R,
                             int64_t T, int64_t C, int64_t gs, cudaStream_t st);

}  // namespace lcb
