// fp8_infer.cu — the block-wise FP8 inference step next to the quantization path:
//   lcb_act_quant_fp8  = act_quant  (kernel.py:7-54):    per-128 activation groups -> e4m3 codes + fp32 scales, HBM bound
//   lcb_fp8_gemm       = fp8_gemm   (kernel.py:140-242) (+ the bias add of block_wise_fp8_forward_func,
//                        module_utils.py:41-46): C = sum_kb (A8[:, kb] . B8[:, kb]^T) * a_s[:, kb] * b_s[n / 128, kb]
//
// The GEMM is a persistent warp-specialised tcgen05 kernel. fp32 block scales cannot ride the hardware's block-scaled MMA
// (that takes UE8M0 / UE4M3 scale factors), so every 128-deep K block is contracted into a FRESH TMEM accumulator stage by
// tcgen05.mma.kind::f8f6f4 and promoted by the epilogue warps: acc_reg = fma(tmem, a_s * b_s, acc_reg) — the tensor
// pipe runs K block k + 1 while the CUDA cores promote K block k (2 TMEM stages of 256 columns).
//   warp 0    : TMA producer, 4-stage ring of (A 128 x 128 B, B 256 x 128 B) e4m3 tiles, 128-byte swizzle, K-major both
//   warp 1    : MMA issuer (one thread): 4 x (M = 128, N = 256, K = 32) per K block, accumulate flag cleared on the first
//   warps 4-11: promotion + epilogue: thread = output row (TMEM lane) x 128-column half, 128 fp32 accumulators in
//               registers (setmaxnreg: 232 for these two warpgroups, 40 for the producer / issuer warpgroup)
#include <algorithm>
#include <cstdlib>
#include "tc.cuh"

namespace lcb {
namespace f8 {

// ----------------------------------------------------------------------------------------------------------------------
// act_quant: y = (x / s).to(e4m3), s = max|x| / 448 over each group of `G` consecutive elements (kernel.py:7-29)
// 8 elements per lane, LANES = G / 8 lanes per group, UNROLL independent groups per lane in flight.
// ----------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t cvt_e4m3x2(float lo, float hi) {
  uint16_t r;
  asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(r) : "f"(hi), "f"(lo));
  return r;
}

// The reference's only definition of act_quant is its Triton kernel (kernel.py:7-29), and Triton lowers fp32 `/` to
// PTX div.full.f32 (seen in the PTX it generates on this B200: `div.full.f32 %s, %amax, 448.0; div.full.f32 %y, %x, %s`):
// a full-range approximate division, <= 2 ulp, NOT the IEEE quotient. bf16 inputs put many x / s exactly on an e4m3
// rounding tie, where the last bit of the quotient decides the code — so the same instruction is used here, for the scale
// and for the elements (0 / 0 and inf / inf give NaN codes as in the reference, no special cases).
__device__ __forceinline__ float div_full(float a, float b) {
  float r;
  asm("div.full.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
  return r;
}

// max |x| over the 8 values of a slot. 16-bit inputs: on the packed words (|x| by clearing the sign bits, NaN-propagating
// packed max), one unpack at the end.
template <typename T>
__device__ __forceinline__ float absmax8(const uint4& raw, const float* f) {
  if (sizeof(T) == 4) {
    float a = fabsf(f[0]);
#pragma unroll
    for (int i = 1; i < 8; ++i) a = max_nan(a, fabsf(f[i]));
    return a;
  } else if (TI<T>::ID == LCB_BF16) {
    const uint32_t m = 0x7fff7fffu;
    uint32_t x0 = raw.x & m, x1 = raw.y & m, x2 = raw.z & m, x3 = raw.w & m, r0, r1, r2;
    asm("max.NaN.bf16x2 %0, %1, %2;" : "=r"(r0) : "r"(x0), "r"(x1));
    asm("max.NaN.bf16x2 %0, %1, %2;" : "=r"(r1) : "r"(x2), "r"(x3));
    asm("max.NaN.bf16x2 %0, %1, %2;" : "=r"(r2) : "r"(r0), "r"(r1));
    return max_nan(__uint_as_float(r2 << 16), __uint_as_float(r2 & 0xffff0000u));
  } else {
    const uint32_t m = 0x7fff7fffu;
    uint32_t x0 = raw.x & m, x1 = raw.y & m, x2 = raw.z & m, x3 = raw.w & m, r0, r1, r2;
    asm("max.NaN.f16x2 %0, %1, %2;" : "=r"(r0) : "r"(x0), "r"(x1));
    asm("max.NaN.f16x2 %0, %1, %2;" : "=r"(r1) : "r"(x2), "r"(x3));
    asm("max.NaN.f16x2 %0, %1, %2;" : "=r"(r2) : "r"(r0), "r"(r1));
    const float2 t = __half22float2(*reinterpret_cast<const __half2*>(&r2));
    return max_nan(t.x, t.y);
  }
}

// One warp-chunk = UNROLL rows of 32 consecutive slots (slot = 8 elements; slot v belongs to group v / LANES), so that
// every access of a chunk is the chunk pointer + lane offset + a compile-time immediate. FULL = false: the last, partial
// chunk (per-slot guards).
template <typename T, int LANES, int UNROLL, bool FULL>
__device__ __forceinline__ void act_quant_chunk(const uint8_t* __restrict__ xp, uint8_t* __restrict__ yp,
                                                float* __restrict__ sp, int lane, int n_live) {
  constexpr int rs = 32;   // slots between the rows of a chunk
  constexpr int SLOT_BYTES = 8 * (int)sizeof(T);
  constexpr int NR = sizeof(T) == 4 ? 2 : 1;
  uint4 raw[UNROLL][NR];
#pragma unroll
  for (int u = 0; u < UNROLL; ++u) {
    if (FULL || u * rs + lane < n_live) {
#pragma unroll
      for (int h = 0; h < NR; ++h) raw[u][h] = ld_stream(xp + (u * rs) * SLOT_BYTES + h * 16);
    } else {
#pragma unroll
      for (int h = 0; h < NR; ++h) raw[u][h] = make_uint4(0, 0, 0, 0);
    }
  }
#pragma unroll
  for (int u = 0; u < UNROLL; ++u) {
    float f[8];
    if (sizeof(T) == 4) {
      TI<float>::unpack(raw[u][0], f);
      TI<float>::unpack(raw[u][NR - 1], f + 4);
    } else {
      TI<T>::unpack(raw[u][0], f);
    }
    float a = absmax8<T>(raw[u][0], f);
#pragma unroll
    for (int o = LANES / 2; o > 0; o >>= 1) a = max_nan(a, __shfl_xor_sync(0xffffffffu, a, o));
    uint32_t w[2];
    const float s = div_full(a, 448.0f);
    float qv[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) qv[i] = div_full(f[i], s);
    w[0] = cvt_e4m3x2(qv[0], qv[1]) | (cvt_e4m3x2(qv[2], qv[3]) << 16);
    w[1] = cvt_e4m3x2(qv[4], qv[5]) | (cvt_e4m3x2(qv[6], qv[7]) << 16);
    if (FULL || u * rs + lane < n_live) {
      *reinterpret_cast<uint2*>(yp + (u * rs) * 8) = make_uint2(w[0], w[1]);
      if ((lane % LANES) == 0) sp[u * (rs / LANES)] = s;
    }
  }
}

template <typename T, int LANES, int UNROLL>
__global__ void __launch_bounds__(256) act_quant_kernel(const T* __restrict__ x, uint8_t* __restrict__ y,
                                                        float* __restrict__ s_out, int64_t n_groups) {
  constexpr int SLOT_BYTES = 8 * (int)sizeof(T);
  constexpr int CHUNK = 32 * UNROLL;   // slots per warp-chunk: UNROLL adjacent rows of 32 slots (2 KB of bf16 per warp)
  const int64_t n_slots = n_groups * LANES;
  const int64_t n_chunks = (n_slots + CHUNK - 1) / CHUNK;
  const int lane = threadIdx.x & 31;
  const int64_t n_warps = (int64_t)gridDim.x * (blockDim.x >> 5);
  for (int64_t c = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); c < n_chunks; c += n_warps) {
    const int64_t v0 = c * CHUNK;
    const uint8_t* xp = reinterpret_cast<const uint8_t*>(x) + (v0 + lane) * SLOT_BYTES;
    uint8_t* yp = y + (v0 + lane) * 8;
    float* sp = s_out + (v0 + lane) / LANES;
    if (v0 + CHUNK <= n_slots) act_quant_chunk<T, LANES, UNROLL, true>(xp, yp, sp, lane, CHUNK);
    else act_quant_chunk<T, LANES, UNROLL, false>(xp, yp, sp, lane, (int)(n_slots - v0));
  }
}

template <typename T, int LANES>
int launch_act_quant(const void* x, void* y, float* s, int64_t n_groups, cudaStream_t st) {
  constexpr int UNROLL = sizeof(T) == 4 ? 2 : 4;   // 64 B of loads in flight per lane (sweep: profiles/r1_fp8_bench.jsonl)
  const int64_t n_slots = n_groups * LANES;
  int64_t blocks = (n_slots + 256 * UNROLL - 1) / (256 * UNROLL);
  const int64_t cap = (int64_t)kNumSMs * 8;
  const int grid = (int)(blocks < 1 ? 1 : (blocks > cap ? cap : blocks));
  act_quant_kernel<T, LANES, UNROLL><<<grid, 256, 0, st>>>(reinterpret_cast<const T*>(x), reinterpret_cast<uint8_t*>(y), s, n_groups);
  LCB_CUDA_CHECK_LAUNCH("act_quant_fp8");
  return LCB_OK;
}

template <typename T>
int dispatch_act_quant(const void* x, void* y, float* s, int64_t n_groups, int64_t group, cudaStream_t st) {
  switch (group) {
    case 8: return launch_act_quant<T, 1>(x, y, s, n_groups, st);
    case 16: return launch_act_quant<T, 2>(x, y, s, n_groups, st);
    case 32: return launch_act_quant<T, 4>(x, y, s, n_groups, st);
    case 64: return launch_act_quant<T, 8>(x, y, s, n_groups, st);
    case 128: return launch_act_quant<T, 16>(x, y, s, n_groups, st);
    case 256: return launch_act_quant<T, 32>(x, y, s, n_groups, st);
    default: return LCB_E_UNSUPPORTED;
  }
}

// ----------------------------------------------------------------------------------------------------------------------
// fp8_gemm
// ----------------------------------------------------------------------------------------------------------------------
constexpr int BM = 128, BN = 256, BKB = 128;          // BKB: K block in elements == bytes == one 128B swizzle row
constexpr int STAGES = 4;
constexpr int A_BYTES = BM * BKB, B_BYTES = BN * BKB, STAGE_BYTES = A_BYTES + B_BYTES;   // 48 KB
constexpr int NACC = 512 / BN;                        // TMEM accumulator stages
constexpr int EPI_WARPS = 8;                          // two per TMEM lane quarter: one waits on tcgen05.ld, one runs FMAs
constexpr int EPI_COLS = BN / (EPI_WARPS / 4);        // accumulator columns per epilogue thread
constexpr int FIRST_EPI_WARP = 4;
constexpr int NTHREADS = (FIRST_EPI_WARP + EPI_WARPS) * 32;
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 256 + 1024;

struct Params {
  int M, N, K;
  int nblk_i, nblk_j, n_tiles, nkb, nsb;   // nsb = ceil(N / 128) rows of b_s
  const float* a_s;   // [M, nkb]
  const float* b_s;   // [ceil(N / 128), nkb]
  void* c;            // [M, N], leading dimension ldc
  int64_t ldc;
  const void* bias;   // NULL or [N]
  int bias_dtype;
  uint32_t idesc;
  // split-K (few output tiles: M <= 128 decode shapes): tile t -> (output tile t / ksplit, K slice t % ksplit of kbs blocks);
  // slices add their fp32 partial sums into `ws` [M, N]; the slice that arrives last reads the sums back, rounds, adds the
  // bias, stores the tile and leaves sums and counter zero for the next call (the workspace is zero between calls)
  int ksplit, kbs;
  float* ws;
  int* cnt;           // arrivals per output tile (behind the sums in the workspace); zero between calls
  int col_fast;       // tile order: column blocks fastest (a wave covers few row blocks x all column blocks)
};
struct Maps {
  CUtensorMap a, b;
};

__device__ __forceinline__ void mma_f8_ss(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// kind::f8f6f4 instruction descriptor: D fp32 (bit 4), A / B e4m3 (0) or e5m2 (1), both K-major
__host__ __device__ constexpr uint32_t idesc_f8(int afmt, int bfmt, int M_, int N_) {
  return (1u << 4) | ((uint32_t)afmt << 7) | ((uint32_t)bfmt << 10) | ((uint32_t)(N_ >> 3) << 17) | ((uint32_t)(M_ >> 4) << 24);
}

template <typename TO>
__global__ void __launch_bounds__(NTHREADS, 1) fp8_gemm_kernel(const __grid_constant__ Maps maps, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
  uint64_t* empty = full + STAGES;
  uint64_t* tfull = empty + STAGES;
  uint64_t* tempty = tfull + NACC;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + NACC);
  volatile int* last_flag = reinterpret_cast<volatile int*>(tmem_slot + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    tc::tma_prefetch_desc(&maps.a);
    tc::tma_prefetch_desc(&maps.b);
    for (int s = 0; s < STAGES; ++s) {
      tc::mbar_init(&full[s], 1);
      tc::mbar_init(&empty[s], 1);
    }
    for (int a = 0; a < NACC; ++a) {
      tc::mbar_init(&tfull[a], 1);
      tc::mbar_init(&tempty[a], EPI_WARPS);
    }
    tc::fence_barrier_init();
  }
  if (warp == 1) tc::tmem_alloc(tmem_slot, 512);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int nkb = p.nkb;

  if (warp < FIRST_EPI_WARP) {
  asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = blockIdx.x; t < p.n_tiles; t += gridDim.x) {
        // tile index runs over column blocks fastest (p.col_fast): consecutive CTAs share the A row block
        const int to = t / p.ksplit, ks = t % p.ksplit;
        const int ti = p.col_fast ? to / p.nblk_j : to % p.nblk_i, tj = p.col_fast ? to % p.nblk_j : to / p.nblk_i;
        const int kb0 = ks * p.kbs, kb1 = min(nkb, kb0 + p.kbs);
        for (int kb = kb0; kb < kb1; ++kb) {
          tc::mbar_wait(&empty[stage], phase ^ 1);
          uint8_t* sa = smem + stage * STAGE_BYTES;
          tc::mbar_expect_tx(&full[stage], STAGE_BYTES);
          tc::tma_load_2d(sa, &maps.a, &full[stage], kb * BKB, ti * BM);
          tc::tma_load_2d(sa + A_BYTES, &maps.b, &full[stage], kb * BKB, tj * BN);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    int stage = 0, acc = 0;
    uint32_t phase = 0, acc_phase = 0;
    const uint64_t desc_k = tc::smem_desc_sw128(0, 1024);   // K-major SWIZZLE_128B: 8-row groups 1024 B apart
    for (int t = blockIdx.x; t < p.n_tiles; t += gridDim.x) {
      const int kb0 = (t % p.ksplit) * p.kbs, kb1 = min(nkb, kb0 + p.kbs);
      for (int kb = kb0; kb < kb1; ++kb) {
        tc::mbar_wait(&tempty[acc], acc_phase ^ 1);
        tc::mbar_wait(&full[stage], phase);
        tc::tc_fence_after();
        if (lane == 0) {
          const uint32_t sa = tc::smem_u32(smem + stage * STAGE_BYTES);
          const uint32_t sb = sa + A_BYTES;
          const uint32_t d_tmem = tmem_base + (uint32_t)acc * BN;
#pragma unroll
          for (int k = 0; k < BKB / 32; ++k)   // K = 32 e4m3 per instruction = 32 B along the swizzled row
            mma_f8_ss(d_tmem, tc::smem_desc_at(desc_k, sa + k * 32), tc::smem_desc_at(desc_k, sb + k * 32), p.idesc, k != 0 ? 1u : 0u);
          tc::mma_commit(&empty[stage]);
          tc::mma_commit(&tfull[acc]);
        }
        __syncwarp();
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1;
        }
        if (++acc == NACC) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
  }
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
    // ===== promotion + epilogue =====
    const int q = warp & 3;   // TMEM lane quarter of this warp
    const int ch = ((warp - FIRST_EPI_WARP) >> 2) * EPI_COLS;   // first accumulator column of this warp
    int acc = 0;
    uint32_t acc_phase = 0;
    float accv[EPI_COLS];
    for (int t = blockIdx.x; t < p.n_tiles; t += gridDim.x) {
      const int to = t / p.ksplit, ks = t % p.ksplit;
      const int ti = p.col_fast ? to / p.nblk_j : to % p.nblk_i, tj = p.col_fast ? to % p.nblk_j : to / p.nblk_i;
      const int kb0 = ks * p.kbs, kb1 = min(nkb, kb0 + p.kbs);
      const int row = ti * BM + q * 32 + lane;
      const float* asp = p.a_s + (int64_t)min(row, p.M - 1) * nkb;
      // weight scale block of this warp's 128 columns (clamped for the half tile past the ragged edge: nothing is stored)
      const float* bsp = p.b_s + (int64_t)min((tj * BN + ch) >> 7, p.nsb - 1) * nkb;
#pragma unroll
      for (int j = 0; j < EPI_COLS; ++j) accv[j] = 0.f;
      float sc = __fmul_rn(asp[kb0], bsp[kb0]);
#pragma unroll 1
      for (int kb = kb0; kb < kb1; ++kb) {
        const int kn = min(kb + 1, nkb - 1);
        const float sc_next = __fmul_rn(__ldg(asp + kn), __ldg(bsp + kn));
        tc::mbar_wait(&tfull[acc], acc_phase);
        tc::tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BN + ch);
        // software pipeline over 32-column chunks: the tcgen05.ld of chunk c + 1 is in flight while chunk c is promoted
        uint32_t v0[32], v1[32];
        tc::tmem_ld_32x32(taddr, v0);
        tc::tmem_ld_wait();
#pragma unroll
        for (int c = 0; c < EPI_COLS / 32; c += 2) {
          tc::tmem_ld_32x32(taddr + (c + 1) * 32, v1);
#pragma unroll
          for (int j = 0; j < 32; ++j) accv[c * 32 + j] = __fmaf_rn(__uint_as_float(v0[j]), sc, accv[c * 32 + j]);
          tc::tmem_ld_wait();
          if (c + 2 < EPI_COLS / 32) tc::tmem_ld_32x32(taddr + (c + 2) * 32, v0);
#pragma unroll
          for (int j = 0; j < 32; ++j) accv[(c + 1) * 32 + j] = __fmaf_rn(__uint_as_float(v1[j]), sc, accv[(c + 1) * 32 + j]);
          if (c + 2 < EPI_COLS / 32) tc::tmem_ld_wait();
        }
        tc::tc_fence_before();
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(&tempty[acc]);
        if (++acc == NACC) {
          acc = 0;
          acc_phase ^= 1;
        }
        sc = sc_next;
      }
      if (p.ksplit > 1) {
        // ---- split-K: fp32 partial sums of this K slice into the workspace (L2 reductions, 16 bytes each) ----
        const int col0 = tj * BN + ch;
        float* wrow = p.ws + (int64_t)row * p.N + col0;
        const bool vec_ok = (p.N % 4 == 0);
        if (row < p.M) {
#pragma unroll
          for (int c = 0; c < EPI_COLS; c += 4) {
            if (vec_ok && col0 + c + 4 <= p.N) {
              asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(wrow + c), "f"(accv[c]), "f"(accv[c + 1]),
                           "f"(accv[c + 2]), "f"(accv[c + 3])
                           : "memory");
            } else {
#pragma unroll
              for (int i = 0; i < 4; ++i)
                if (col0 + c + i < p.N) atomicAdd(wrow + c + i, accv[c + i]);
            }
          }
        }
        // the slice that arrives last owns the tile (threadFenceReduction pattern: reductions, fence, barrier, one counter
        // increment; the owner fences again and reads the sums from L2)
        __threadfence();
        asm volatile("bar.sync 1, %0;" ::"n"(EPI_WARPS * 32) : "memory");
        if (threadIdx.x == FIRST_EPI_WARP * 32) *last_flag = (atomicAdd(p.cnt + to, 1) == p.ksplit - 1) ? 1 : 0;
        asm volatile("bar.sync 1, %0;" ::"n"(EPI_WARPS * 32) : "memory");
        if (*last_flag == 0) continue;
        __threadfence();
        // the owner's tail is a ROLLED loop over the live part of the tile, coalesced along the rows of the workspace
        // (a second unrolled 128-accumulator block would be cold code: its instruction fetches cost more than the data):
        // read the sums, put the zeros back, round, add the bias, store
        {
          const int et = threadIdx.x - FIRST_EPI_WARP * 32;
          const int r0 = ti * BM, c0 = tj * BN;
          const int rows = min(BM, p.M - r0), cols = min(BN, p.N - c0);
          TO* cb = reinterpret_cast<TO*>(p.c);
          if (vec_ok && (reinterpret_cast<uintptr_t>(p.c) & 15) == 0) {
            const int nq = cols >> 2, total = rows * nq;
            constexpr int U = 8, NT = EPI_WARPS * 32;       // U loads in flight per thread: the loop is L2-latency bound
#pragma unroll 1
            for (int base = et; base < total; base += U * NT) {
              float4 v[U];
              int64_t off[U];
#pragma unroll
              for (int u = 0; u < U; ++u) {
                const int idx = base + u * NT;
                const int r = idx / nq, c = (idx - r * nq) << 2;
                off[u] = (int64_t)(r0 + r) * p.N + c0 + c;
                if (idx < total) v[u] = __ldcg(reinterpret_cast<const float4*>(p.ws + off[u]));
              }
#pragma unroll
              for (int u = 0; u < U; ++u) {
                if (base + u * NT >= total) break;
                __stcg(reinterpret_cast<float4*>(p.ws + off[u]), make_float4(0.f, 0.f, 0.f, 0.f));
                float o[4] = {TI<TO>::rd(v[u].x), TI<TO>::rd(v[u].y), TI<TO>::rd(v[u].z), TI<TO>::rd(v[u].w)};
                if (p.bias) {
                  const int c = (int)(off[u] - (off[u] / p.N) * p.N);
#pragma unroll
                  for (int i = 0; i < 4; ++i) o[i] = TI<TO>::rd(__fadd_rn(o[i], load_rt(p.bias, c + i, p.bias_dtype)));
                }
                if (sizeof(TO) == 4) {
                  *reinterpret_cast<float4*>(cb + off[u]) = make_float4(o[0], o[1], o[2], o[3]);
                } else {
                  TO t4[4];
#pragma unroll
                  for (int i = 0; i < 4; ++i) TI<TO>::store1(t4 + i, o[i]);
                  *reinterpret_cast<uint2*>(cb + off[u]) = *reinterpret_cast<const uint2*>(t4);
                }
              }
            }
          } else {
#pragma unroll 1
            for (int idx = et; idx < rows * cols; idx += EPI_WARPS * 32) {
              const int r = idx / cols, c = idx - r * cols;
              const int64_t off = (int64_t)(r0 + r) * p.N + c0 + c;
              float o = TI<TO>::rd(__ldcg(p.ws + off));
              __stcg(p.ws + off, 0.f);
              if (p.bias) o = TI<TO>::rd(__fadd_rn(o, load_rt(p.bias, c0 + c, p.bias_dtype)));
              TI<TO>::store1(cb + off, o);
            }
          }
        }
        if (threadIdx.x == FIRST_EPI_WARP * 32) p.cnt[to] = 0;
        continue;
      }
      // ---- store: round to the output dtype, then the bias add of block_wise_fp8_forward_func (y += bias) ----
      if (row < p.M) {
        const int col0 = tj * BN + ch;
        TO* crow = reinterpret_cast<TO*>(p.c) + (int64_t)row * p.ldc + col0;
        constexpr int V = TI<TO>::VEC;
        const bool vec_ok = (p.ldc % V == 0) && ((reinterpret_cast<uintptr_t>(p.c) & 15) == 0);
#pragma unroll
        for (int c = 0; c < EPI_COLS; c += V) {
          float o[V];
#pragma unroll
          for (int i = 0; i < V; ++i) o[i] = TI<TO>::rd(accv[c + i]);
          if (p.bias) {
#pragma unroll
            for (int i = 0; i < V; ++i)
              if (col0 + c + i < p.N) o[i] = TI<TO>::rd(__fadd_rn(o[i], load_rt(p.bias, col0 + c + i, p.bias_dtype)));
          }
          if (vec_ok && col0 + c + V <= p.N) {
            *reinterpret_cast<uint4*>(crow + c) = TI<TO>::pack(o);
          } else {
#pragma unroll
            for (int i = 0; i < V; ++i)
              if (col0 + c + i < p.N) TI<TO>::store1(crow + c + i, o[i]);
          }
        }
      }
    }
  }

  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc::tc_fence_after();
    tc::tmem_dealloc(tmem_base, 512);
  }
}

template <typename TO>
int launch_gemm(const Maps& maps, const Params& p, int grid, cudaStream_t st) {
  static DeviceOnce attr_once;
  if (attr_once.needed()) {
    if (cudaFuncSetAttribute(fp8_gemm_kernel<TO>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES) != cudaSuccess) {
      set_last_error("cudaFuncSetAttribute(fp8_gemm)", cudaGetLastError());
      return LCB_E_CUDA;
    }
    attr_once.done();
  }
  fp8_gemm_kernel<TO><<<grid, NTHREADS, SMEM_BYTES, st>>>(maps, p);
  LCB_CUDA_CHECK_LAUNCH("fp8_gemm");
  return LCB_OK;
}

}  // namespace f8
}  // namespace lcb

extern "C" int lcb_act_quant_fp8(const void* x, void* y, float* scales, int64_t n_groups, int64_t group, int32_t dtype,
                                 void* stream) {
  using namespace lcb;
  if (n_groups == 0) return LCB_OK;
  if (!x || !y || !scales || n_groups < 0) return LCB_E_INVALID;
  if ((reinterpret_cast<uintptr_t>(x) & 15) || (reinterpret_cast<uintptr_t>(y) & 7)) return LCB_E_ALIGN;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (dtype == LCB_BF16) return f8::dispatch_act_quant<__nv_bfloat16>(x, y, scales, n_groups, group, st);
  if (dtype == LCB_F16) return f8::dispatch_act_quant<__half>(x, y, scales, n_groups, group, st);
  if (dtype == LCB_F32) return f8::dispatch_act_quant<float>(x, y, scales, n_groups, group, st);
  return LCB_E_UNSUPPORTED;
}

namespace lcb {
namespace f8 {
// K slices per output tile when the tiles alone leave most SMs idle. Model fitted on the B200 (tools/fp8_decode_bench.py,
// profiles/r2_fp8_ksplit_sweep.txt): a call costs ~6.3 us + 0.5 us per K block of its longest CTA; a split tile adds
// ~2.7 us + 0.065 us per live row (16-byte L2 reductions of every slice, the owner's read-back and store). One wave only:
// slices x tiles <= SMs. LCB_FP8_KSPLIT=<n> forces n slices (experiments).
static int pick_ksplit(int n_out_tiles, int nkb, int64_t M) {
  if (n_out_tiles * 2 > kNumSMs || nkb < 4) return 1;
  static const int forced = [] { const char* e = getenv("LCB_FP8_KSPLIT"); return e ? atoi(e) : 0; }();
  int ks = forced > 0 ? forced : kNumSMs / n_out_tiles;
  ks = std::max(1, std::min(ks, nkb / 2));
  if (forced > 0 || ks == 1) return ks;
  const int kbs = (nkb + ks - 1) / ks;
  const float saving = 0.5f * (float)(nkb - kbs);
  const float tail = 2.7f + 0.065f * (float)std::min<int64_t>(M, BM);
  return saving > 1.2f * tail ? ks : 1;
}
}  // namespace f8
}  // namespace lcb

extern "C" int64_t lcb_fp8_gemm_workspace_bytes(int64_t M, int64_t N, int64_t K) {
  using namespace lcb::f8;
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  const int n_out = (int)((M + BM - 1) / BM) * (int)((N + BN - 1) / BN);
  if (pick_ksplit(n_out, (int)((K + BKB - 1) / BKB), M) <= 1) return 0;
  return ((M * N + 3) / 4 * 4 + (n_out + 3) / 4 * 4) * (int64_t)sizeof(float);   // fp32 sums [M, N] + one counter per tile
}

static int fp8_gemm_impl(const void* a, const float* a_s, const void* b, const float* b_s, void* c, const void* bias,
                         int64_t M, int64_t N, int64_t K, int32_t out_dtype, int32_t bias_dtype, void* workspace,
                         void* stream);

extern "C" int lcb_fp8_gemm(const void* a, const float* a_s, const void* b, const float* b_s, void* c, const void* bias,
                            int64_t M, int64_t N, int64_t K, int32_t out_dtype, int32_t bias_dtype, void* stream) {
  return fp8_gemm_impl(a, a_s, b, b_s, c, bias, M, N, K, out_dtype, bias_dtype, nullptr, stream);
}

extern "C" int lcb_fp8_gemm_ws(const void* a, const float* a_s, const void* b, const float* b_s, void* c, const void* bias,
                               int64_t M, int64_t N, int64_t K, int32_t out_dtype, int32_t bias_dtype, void* workspace,
                               void* stream) {
  return fp8_gemm_impl(a, a_s, b, b_s, c, bias, M, N, K, out_dtype, bias_dtype, workspace, stream);
}

static int fp8_gemm_impl(const void* a, const float* a_s, const void* b, const float* b_s, void* c, const void* bias,
                         int64_t M, int64_t N, int64_t K, int32_t out_dtype, int32_t bias_dtype, void* workspace,
                         void* stream) {
  using namespace lcb;
  using namespace lcb::f8;
  if (M == 0 || N == 0) return LCB_OK;
  if (!a || !a_s || !b || !b_s || !c || M < 0 || N < 0 || K <= 0) return LCB_E_INVALID;
  if (M > (1 << 30) || N > (1 << 30) || K > (1 << 30)) return LCB_E_SHAPE;
  if ((K % 16) || (reinterpret_cast<uintptr_t>(a) & 15) || (reinterpret_cast<uintptr_t>(b) & 15)) return LCB_E_ALIGN;   // TMA pitch
  if (bias && bias_dtype != LCB_F32 && bias_dtype != LCB_F16 && bias_dtype != LCB_BF16) return LCB_E_UNSUPPORTED;
  Maps maps;
  Params p;
  p.M = (int)M;
  p.N = (int)N;
  p.K = (int)K;
  p.nblk_i = (int)((M + BM - 1) / BM);
  p.nblk_j = (int)((N + BN - 1) / BN);
  p.nkb = (int)((K + BKB - 1) / BKB);
  p.ksplit = 1;
  if (workspace) {
    if (reinterpret_cast<uintptr_t>(workspace) & 15) return LCB_E_ALIGN;
    p.ksplit = pick_ksplit(p.nblk_i * p.nblk_j, p.nkb, M);
  }
  p.kbs = (p.nkb + p.ksplit - 1) / p.ksplit;
  p.ksplit = (p.nkb + p.kbs - 1) / p.kbs;          // no empty slice
  p.ws = reinterpret_cast<float*>(workspace);
  // a wave of 148 tiles covers a few row blocks x all column blocks: an A row block is read from HBM once (shared by the
  // CTAs of its row through the L2) and the weight stays L2-resident across waves. ncu, 8192 x 2048 x 7168: dram read
  // 132 -> 75 MB (= the operands), L2 hit 80 -> 87 %, +2-13 % TFLOP/s. LCB_FP8_COL_FAST=0 restores row blocks fastest.
  p.col_fast = getenv("LCB_FP8_COL_FAST") ? atoi(getenv("LCB_FP8_COL_FAST")) : 1;
  p.cnt = workspace ? reinterpret_cast<int*>(p.ws + (M * N + 3) / 4 * 4) : nullptr;
  p.n_tiles = p.nblk_i * p.nblk_j * p.ksplit;
  p.nsb = (int)((N + 127) / 128);
  p.a_s = a_s;
  p.b_s = b_s;
  p.c = c;
  p.ldc = N;
  p.bias = bias;
  p.bias_dtype = bias_dtype;
  p.idesc = idesc_f8(0, 0, BM, BN);
  bool ok = make_tmap_2d(&maps.a, a, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, M, K, K, BM, BKB) &&
            make_tmap_2d(&maps.b, b, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, N, K, K, BN, BKB);
  if (!ok) {
    set_last_error("cuTensorMapEncodeTiled", cudaErrorInvalidValue);
    return LCB_E_CUDA;
  }
  const int grid = p.n_tiles < kNumSMs ? p.n_tiles : kNumSMs;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (out_dtype == LCB_BF16) return launch_gemm<__nv_bfloat16>(maps, p, grid, st);
  if (out_dtype == LCB_F16) return launch_gemm<__half>(maps, p, grid, st);
  if (out_dtype == LCB_F32) return launch_gemm<float>(maps, p, grid, st);
  return LCB_E_UNSUPPORTED;
}
