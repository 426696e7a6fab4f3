// autoclip.cu — AutoClip v1 search (auto_clip.py:84-191): for every (output channel, group) pick the shrink step
// i_s in [0, n_steps) whose clamped + fake-quantized weights reproduce the layer output best on the sampled tokens.
//
//   ac_init_kernel    : org_max / org_min per (row, group) (clip_sym: org_max = |w|.amax), best = org, min_err = 1e9
//   ac_literal_kernel : out[r, g, t] = rd( sum_j rd(x[t, g, j] * w[r, g, j]) )  — the reference's `(x * w).sum(dim=-1)`
//                       (auto_clip.py:137, 148) LITERALLY: every product is rounded to the tensor dtype (one packed
//                       HMUL2 per two products), the sum runs in fp32 and is rounded once. ORG mode stores it
//                       (org_out); STEP mode compares it with org_out: err = rd(mean_t rd(rd(out - org)^2)), and keeps
//                       the step when err < min_err (strict: the first minimum wins, like the reference loop).
//   per step          : lcb_fakequant with spec.clip_ratio = +/- (1 - i_s / n_grid)   (clamp fused into the quant kernel)
//                       then ac_literal_kernel<STEP>.
// Round 1 ran the products on tcgen05 (exact products summed in fp32): 3.6x faster, but the per-product rounding of the
// reference moves ~30 % of the outputs by one ulp of the tensor dtype, which moves the bf16-rounded error mean and picked
// another shrink step for 3-15 % of the (channel, group) pairs. The selected index is the contract (north star: "same
// selected clip index"), so the arithmetic is the reference's. The reference materialises x * w as a [256, tok, C] tensor
// per step (1 GiB at Llama-3-8B shapes); here a thread owns one (row, group), keeps the 128 weights of its row in 64
// packed registers and streams token tiles from shared memory (broadcast LDS.128): no intermediate tensor.
// What cannot be pinned: the ORDER of torch's fp32 sum over the 128 rounded products (CPU: vectorised cascade, CUDA: a
// tree of per-thread partials). Here: even elements in order, odd elements in order, then one add. A different order
// changes the fp32 sum by ~1e-7 relative, which matters only when that sum lies within that distance of a rounding
// boundary of the tensor dtype (probability ~2^-15 per output).
#include <stdlib.h>
#include "tc_gemm.cuh"

namespace lcb {
namespace {

template <typename T>
__global__ void __launch_bounds__(256) ac_init_kernel(const T* __restrict__ W, T* __restrict__ best_max, T* __restrict__ best_min,
                                                      float* __restrict__ omax, float* __restrict__ omin,
                                                      float* __restrict__ min_err, int64_t n_units, int gs, int clip_sym) {
  const int lane = threadIdx.x & 31;
  const int64_t unit = ((int64_t)blockIdx.x * 256 + threadIdx.x) >> 5;
  if (unit >= n_units) return;
  const T* p = W + unit * gs;
  float mx = -INFINITY, mn = INFINITY;
  for (int c = lane; c < gs; c += 32) {
    const float v = TI<T>::load1(p + c);
    mx = max_nan(mx, clip_sym ? fabsf(v) : v);
    mn = min_nan(mn, v);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    mx = max_nan(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    mn = min_nan(mn, __shfl_xor_sync(0xffffffffu, mn, o));
  }
  if (lane == 0) {
    omax[unit] = mx;
    omin[unit] = mn;
    TI<T>::store1(best_max + unit, mx);
    TI<T>::store1(best_min + unit, mn);
    min_err[unit] = TI<T>::rd(1e9f);   // torch.ones_like(org_max_val) * 1e9 in the tensor dtype (auto_clip.py:121)
  }
}

// packed pair product rounded to the tensor dtype (IEEE RN, subnormals kept), accumulated in fp32
template <typename T>
struct P2;
template <>
struct P2<__nv_bfloat16> {
  __device__ __forceinline__ static void mul_acc(uint32_t a, uint32_t b, float& even, float& odd) {
    __nv_bfloat162 p = __hmul2(*reinterpret_cast<const __nv_bfloat162*>(&a), *reinterpret_cast<const __nv_bfloat162*>(&b));
    const uint32_t u = *reinterpret_cast<const uint32_t*>(&p);
    even = __fadd_rn(even, __uint_as_float(u << 16));
    odd = __fadd_rn(odd, __uint_as_float(u & 0xffff0000u));
  }
};
template <>
struct P2<__half> {
  __device__ __forceinline__ static void mul_acc(uint32_t a, uint32_t b, float& even, float& odd) {
    __half2 p = __hmul2(*reinterpret_cast<const __half2*>(&a), *reinterpret_cast<const __half2*>(&b));
    const float2 f = __half22float2(p);
    even = __fadd_rn(even, f.x);
    odd = __fadd_rn(odd, f.y);
  }
};

constexpr int AC_ROWS = 128;   // rows (threads) per CTA
constexpr int AC_TOK = 16;     // tokens per shared-memory tile
constexpr int AC_COLS = 128;   // columns per tile (= one group for group_size 128)

// grid (ceil(R / 128), C / gs). ORG: writes org[(g * R + r) * T_rows + t]. STEP: reads it, updates best_* / min_err.
template <typename T, bool ORG>
__global__ void __launch_bounds__(AC_ROWS) ac_literal_kernel(const T* __restrict__ Wq, const T* __restrict__ x, T* __restrict__ org,
                                                            const float* __restrict__ omax, const float* __restrict__ omin,
                                                            float* __restrict__ min_err, T* __restrict__ best_max,
                                                            T* __restrict__ best_min, int64_t R, int64_t C, int64_t T_rows,
                                                            int gs, float ratio, float tokens, int clip_sym) {
  __shared__ __align__(16) uint4 xs[2][AC_TOK * AC_COLS / 8];
  const int64_t r = (int64_t)blockIdx.x * AC_ROWS + threadIdx.x;
  const int g = blockIdx.y;
  const bool live = r < R;
  const int njc = gs / AC_COLS > 0 ? gs / AC_COLS : 1;
  const int cols = gs < AC_COLS ? gs : AC_COLS;            // gs: a multiple of 8 up to 128, or k * 128
  const int nv = cols / 8;                                  // uint4 per tile row
  const T* wrow = Wq + (live ? r : 0) * C + (int64_t)g * gs;
  const T* xcol = x + (int64_t)g * gs;
  T* orow = org + ((int64_t)g * R + (live ? r : 0)) * T_rows;
  uint32_t w[AC_COLS / 2];
  auto load_w = [&](int jc) {
#pragma unroll
    for (int k = 0; k < AC_COLS / 8; ++k) {
      uint4 v = make_uint4(0, 0, 0, 0);
      if (k < nv) v = *reinterpret_cast<const uint4*>(wrow + (int64_t)jc * AC_COLS + k * 8);
      w[4 * k] = v.x;
      w[4 * k + 1] = v.y;
      w[4 * k + 2] = v.z;
      w[4 * k + 3] = v.w;
    }
  };
  auto stage = [&](int buf, int64_t t0, int jc) {             // x[t0 .. t0+16, g*gs + jc*128 .. +cols] -> shared memory
    for (int i = threadIdx.x; i < AC_TOK * nv; i += AC_ROWS) {
      const int t = i / nv, k = i - t * nv;
      uint4 v = make_uint4(0, 0, 0, 0);
      if (t0 + t < T_rows) v = *reinterpret_cast<const uint4*>(xcol + (t0 + t) * C + (int64_t)jc * AC_COLS + k * 8);
      xs[buf][t * (AC_COLS / 8) + k] = v;
    }
  };
  if (njc == 1) load_w(0);
  double sse = 0.0;
  const int64_t ntile = (T_rows + AC_TOK - 1) / AC_TOK * njc;
  stage(0, 0, 0);
  __syncthreads();
  float even[AC_TOK], odd[AC_TOK];
  for (int64_t it = 0; it < ntile; ++it) {
    const int64_t t0 = it / njc * AC_TOK;
    const int jc = (int)(it % njc);
    const int buf = (int)(it & 1);
    if (it + 1 < ntile) stage(buf ^ 1, (it + 1) / njc * AC_TOK, (int)((it + 1) % njc));   // the other buffer: free since the last barrier
    if (jc == 0) {
#pragma unroll
      for (int t = 0; t < AC_TOK; ++t) even[t] = odd[t] = 0.f;
    }
    if (njc > 1) load_w(jc);
    if (live) {
#pragma unroll
      for (int t = 0; t < AC_TOK; ++t) {
#pragma unroll
        for (int k = 0; k < AC_COLS / 8; ++k) {
          if (k < nv) {
            const uint4 xv = xs[buf][t * (AC_COLS / 8) + k];
            P2<T>::mul_acc(xv.x, w[4 * k], even[t], odd[t]);
            P2<T>::mul_acc(xv.y, w[4 * k + 1], even[t], odd[t]);
            P2<T>::mul_acc(xv.z, w[4 * k + 2], even[t], odd[t]);
            P2<T>::mul_acc(xv.w, w[4 * k + 3], even[t], odd[t]);
          }
        }
      }
      if (jc == njc - 1) {
        float out[AC_TOK];
#pragma unroll
        for (int t = 0; t < AC_TOK; ++t) out[t] = TI<T>::rd(__fadd_rn(even[t], odd[t]));
#pragma unroll
        for (int h = 0; h < AC_TOK / 8; ++h) {
          if (t0 + h * 8 < T_rows) {                       // T_rows % 8 == 0
            if (ORG) {
              *reinterpret_cast<uint4*>(orow + t0 + h * 8) = TI<T>::pack(out + h * 8);
            } else {
              float o[8];
              TI<T>::unpack(*reinterpret_cast<const uint4*>(orow + t0 + h * 8), o);
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                const float d = TI<T>::rd(__fsub_rn(out[h * 8 + i], o[i]));       // cur_out - org_out
                sse += (double)TI<T>::rd(__fmul_rn(d, d));                          // .pow(2)
              }
            }
          }
        }
      }
    }
    __syncthreads();
  }
  if (!ORG && live) {
    const int64_t u = r * gridDim.y + g;
    // .mean(dim=1): fp32 accumulation, one division, rounded to the tensor dtype (auto_clip.py:150)
    const float err = TI<T>::rd(__fdiv_rn((float)sse, tokens));
    if (err < min_err[u]) {
      const float mxv = TI<T>::rd(__fmul_rn(omax[u], ratio));                       // org_max_val * (1 - i_s / n_grid)
      const float mnv = clip_sym ? -mxv : TI<T>::rd(__fmul_rn(omin[u], ratio));
      min_err[u] = err;
      TI<T>::store1(best_max + u, mxv);
      TI<T>::store1(best_min + u, mnv);
    }
  }
}

// ---- the same arithmetic with the fp32 SUM on the tensor pipe -----------------------------------------------------------
// The products still come from HMUL2 (rounded to the tensor dtype, as the reference rounds them); they are formed directly
// in the A-fragment layout of mma.sync.m16n8k16 (rows = 16 weight rows, k = 16 columns of the group) and multiplied by a
// B fragment of ones: D[m][*] += sum_k p[m][k], fp32 accumulation. That replaces, per packed pair, two unpack operations and
// two FADDs by 1/4 of a warp-wide MMA: ~0.65 instead of ~2.5 issued instructions per product. The tensor core's fp32 sum
// is one more summation ORDER (see the header: which order torch uses is not pinned either).
// Warp: 16 rows x one group (<= 128 columns, a multiple of 16), weights held as A-side fragments in 32 registers; CTA: 8
// warps = 128 rows sharing the token tiles. Token tiles sit in shared memory PERMUTED so that the 16 packed words a lane
// needs for a token (columns 2t, 2t+1 and 8+2t, 8+2t+1 of every 16-column block; t = lane % 4) are 64 contiguous bytes.
// Four tokens are in flight per iteration (four independent MMA chains); lane t of a quad finishes token t0 + t.
template <typename T>
struct MmaOnes;
template <>
struct MmaOnes<__nv_bfloat16> {
  __device__ __forceinline__ static uint32_t mul(uint32_t a, uint32_t b) {
    __nv_bfloat162 p = __hmul2(*reinterpret_cast<const __nv_bfloat162*>(&a), *reinterpret_cast<const __nv_bfloat162*>(&b));
    return *reinterpret_cast<const uint32_t*>(&p);
  }
  __device__ __forceinline__ static void mma(float* c, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3) {
    const uint32_t one = 0x3f803f80u;
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(one), "r"(one));
  }
};
template <>
struct MmaOnes<__half> {
  __device__ __forceinline__ static uint32_t mul(uint32_t a, uint32_t b) {
    __half2 p = __hmul2(*reinterpret_cast<const __half2*>(&a), *reinterpret_cast<const __half2*>(&b));
    return *reinterpret_cast<const uint32_t*>(&p);
  }
  __device__ __forceinline__ static void mma(float* c, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3) {
    const uint32_t one = 0x3c003c00u;
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(one), "r"(one));
  }
};

constexpr int ACM_TOK = 32;   // tokens per shared-memory tile (8 KB at 128 columns)

// grid (ceil(R / 128), C / gs), 256 threads. NKB = gs / 16 (compile time: 8 for group 128). Same contract as ac_literal_kernel.
template <typename T, bool ORG, int NKB>
__global__ void __launch_bounds__(256, 4) ac_mma_kernel(const T* __restrict__ Wq, const T* __restrict__ x, T* __restrict__ org,
                                                     const float* __restrict__ omax, const float* __restrict__ omin,
                                                     float* __restrict__ min_err, T* __restrict__ best_max, T* __restrict__ best_min,
                                                     int64_t R, int64_t C, int64_t T_rows, float ratio, float tokens, int clip_sym) {
  constexpr int GS = NKB * 16, WPT = GS / 2;      // packed words per token row of the group
  __shared__ __align__(16) uint32_t xs[2][ACM_TOK * WPT];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int gid = lane >> 2, tid = lane & 3;
  const int g = blockIdx.y;
  const int64_t r0 = (int64_t)blockIdx.x * 128 + warp * 16;
  const int64_t ra = r0 + gid, rb = r0 + gid + 8;
  const uint32_t* Wp = reinterpret_cast<const uint32_t*>(Wq);
  const uint32_t* xp = reinterpret_cast<const uint32_t*>(x);
  // A-side weight fragments: (row ra | rb) x (words tid, 4 + tid of every 16-column block)
  uint32_t wa0[NKB], wa1[NKB], wb0[NKB], wb1[NKB];
#pragma unroll
  for (int kb = 0; kb < NKB; ++kb) {
    const int64_t cw = ((int64_t)g * GS + kb * 16) / 2 + tid;
    wa0[kb] = ra < R ? Wp[ra * (C / 2) + cw] : 0u;
    wa1[kb] = ra < R ? Wp[ra * (C / 2) + cw + 4] : 0u;
    wb0[kb] = rb < R ? Wp[rb * (C / 2) + cw] : 0u;
    wb1[kb] = rb < R ? Wp[rb * (C / 2) + cw + 4] : 0u;
  }
  auto stage = [&](int buf, int64_t t0) {
    for (int i = threadIdx.x; i < ACM_TOK * WPT; i += 256) {
      const int t = i / WPT, wd = i - t * WPT;         // word wd of the group's row: block kb = wd / 8, j = wd % 8
      const int kb = wd >> 3, j = wd & 7;
      uint32_t v = 0u;
      if (t0 + t < T_rows) v = xp[(t0 + t) * (C / 2) + ((int64_t)g * GS) / 2 + wd];
      xs[buf][t * WPT + (j & 3) * (2 * NKB) + kb * 2 + (j >> 2)] = v;
    }
  };
  double sse_a = 0.0, sse_b = 0.0;
  const int64_t ntile = (T_rows + ACM_TOK - 1) / ACM_TOK;
  stage(0, 0);
  __syncthreads();
  for (int64_t it = 0; it < ntile; ++it) {
    const int buf = (int)(it & 1);
    const int64_t t0 = it * ACM_TOK;
    if (it + 1 < ntile) stage(buf ^ 1, t0 + ACM_TOK);
#pragma unroll 1
    for (int q = 0; q < ACM_TOK; q += 4) {
      if (t0 + q >= T_rows) break;                       // T_rows % 8 == 0: whole quads
      float c[4][4];
#pragma unroll
      for (int u = 0; u < 4; ++u) c[u][0] = c[u][1] = c[u][2] = c[u][3] = 0.f;
      uint32_t xf[4][2 * NKB];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const uint32_t* src = &xs[buf][(q + u) * WPT + tid * (2 * NKB)];
#pragma unroll
        for (int k = 0; k < 2 * NKB; k += 4) {
          if (2 * NKB - k >= 4) {
            const uint4 v = *reinterpret_cast<const uint4*>(src + k);
            xf[u][k] = v.x; xf[u][k + 1] = v.y; xf[u][k + 2] = v.z; xf[u][k + 3] = v.w;
          } else {
            const uint2 v = *reinterpret_cast<const uint2*>(src + k);
            xf[u][k] = v.x; xf[u][k + 1] = v.y;
          }
        }
      }
#pragma unroll
      for (int kb = 0; kb < NKB; ++kb) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const uint32_t x0 = xf[u][2 * kb], x1 = xf[u][2 * kb + 1];
          MmaOnes<T>::mma(c[u], MmaOnes<T>::mul(wa0[kb], x0), MmaOnes<T>::mul(wb0[kb], x0), MmaOnes<T>::mul(wa1[kb], x1),
                          MmaOnes<T>::mul(wb1[kb], x1));
        }
      }
      // every column of D holds the row sum: lane `tid` of a quad finishes token t0 + q + tid for rows ra and rb
      float oa = c[0][0], ob = c[0][2];
      if (tid == 1) { oa = c[1][0]; ob = c[1][2]; }
      if (tid == 2) { oa = c[2][0]; ob = c[2][2]; }
      if (tid == 3) { oa = c[3][0]; ob = c[3][2]; }
      oa = TI<T>::rd(oa);
      ob = TI<T>::rd(ob);
      const int64_t t = t0 + q + tid;
      if (ORG) {
        if (ra < R) TI<T>::store1(org + ((int64_t)g * R + ra) * T_rows + t, oa);
        if (rb < R) TI<T>::store1(org + ((int64_t)g * R + rb) * T_rows + t, ob);
      } else {
        if (ra < R) {
          const float d = TI<T>::rd(__fsub_rn(oa, TI<T>::load1(org + ((int64_t)g * R + ra) * T_rows + t)));
          sse_a += (double)TI<T>::rd(__fmul_rn(d, d));
        }
        if (rb < R) {
          const float d = TI<T>::rd(__fsub_rn(ob, TI<T>::load1(org + ((int64_t)g * R + rb) * T_rows + t)));
          sse_b += (double)TI<T>::rd(__fmul_rn(d, d));
        }
      }
    }
    __syncthreads();
  }
  if (!ORG) {
    sse_a += __shfl_xor_sync(0xffffffffu, sse_a, 1);
    sse_a += __shfl_xor_sync(0xffffffffu, sse_a, 2);
    sse_b += __shfl_xor_sync(0xffffffffu, sse_b, 1);
    sse_b += __shfl_xor_sync(0xffffffffu, sse_b, 2);
    if (tid == 0) {
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int64_t r = h ? rb : ra;
        if (r >= R) continue;
        const int64_t u = r * gridDim.y + g;
        const float err = TI<T>::rd(__fdiv_rn((float)(h ? sse_b : sse_a), tokens));
        if (err < min_err[u]) {
          const float mxv = TI<T>::rd(__fmul_rn(omax[u], ratio));
          const float mnv = clip_sym ? -mxv : TI<T>::rd(__fmul_rn(omin[u], ratio));
          min_err[u] = err;
          TI<T>::store1(best_max + u, mxv);
          TI<T>::store1(best_min + u, mnv);
        }
      }
    }
  }
}

template <typename T, bool ORG>
int launch_ac(const T* Wq, const T* x, T* org, const float* omax, const float* omin, float* min_err, T* bmax, T* bmin, int64_t R,
              int64_t C, int64_t T_rows, int gs, float ratio, float tokens, int clip_sym, cudaStream_t st) {
  const unsigned ng = (unsigned)(C / gs);
  const dim3 grid((unsigned)((R + 127) / 128), ng);
  const bool mma_ok = (C % 2 == 0) && (T_rows % 8 == 0) && (getenv("LCB_AUTOCLIP_NO_MMA") == nullptr);
  if (mma_ok && gs == 128) {
    ac_mma_kernel<T, ORG, 8><<<grid, 256, 0, st>>>(Wq, x, org, omax, omin, min_err, bmax, bmin, R, C, T_rows, ratio, tokens, clip_sym);
  } else if (mma_ok && gs == 64) {
    ac_mma_kernel<T, ORG, 4><<<grid, 256, 0, st>>>(Wq, x, org, omax, omin, min_err, bmax, bmin, R, C, T_rows, ratio, tokens, clip_sym);
  } else if (mma_ok && gs == 32) {
    ac_mma_kernel<T, ORG, 2><<<grid, 256, 0, st>>>(Wq, x, org, omax, omin, min_err, bmax, bmin, R, C, T_rows, ratio, tokens, clip_sym);
  } else {   // other group sizes (per_channel rows, groups of 8 / 16 / 24 ...): the CUDA-core sum, same arithmetic
    ac_literal_kernel<T, ORG><<<grid, AC_ROWS, 0, st>>>(Wq, x, org, omax, omin, min_err, bmax, bmin, R, C, T_rows, gs, ratio, tokens,
                                                        clip_sym);
  }
  LCB_CUDA_CHECK_LAUNCH("autoclip_literal");
  return LCB_OK;
}

inline int64_t align_up(int64_t x, int64_t a) { return (x + a - 1) / a * a; }

template <typename T>
int autoclip_typed(const void* W, const void* x, void* best_max, void* best_min, int64_t R, int64_t C, int64_t T_rows,
                   int64_t tokens, int64_t gs, int n_grid, int n_steps, int clip_sym, const lcb_quant_spec* spec,
                   uint8_t* ws, cudaStream_t st) {
  const int64_t ng = C / gs, nu = R * ng;
  void* Q = ws;
  uint8_t* q = ws + align_up(R * C * 2, 256);
  void* org = q;
  q += align_up(ng * R * T_rows * 2, 256);
  double* sse = reinterpret_cast<double*>(q);
  q += align_up(nu * 8, 256);
  float* omax = reinterpret_cast<float*>(q);
  q += align_up(nu * 4, 256);
  float* omin = reinterpret_cast<float*>(q);
  q += align_up(nu * 4, 256);
  float* min_err = reinterpret_cast<float*>(q);
  q += align_up(nu * 4, 256);
  void* fqws = q;
  (void)sse;
  ac_init_kernel<T><<<(int)((nu * 32 + 255) / 256), 256, 0, st>>>((const T*)W, (T*)best_max, (T*)best_min, omax, omin, min_err,
                                                                  nu, (int)gs, clip_sym);
  LCB_CUDA_CHECK_LAUNCH("ac_init");
  int rc = launch_ac<T, true>((const T*)W, (const T*)x, (T*)org, omax, omin, min_err, (T*)best_max, (T*)best_min, R, C, T_rows,
                              (int)gs, 1.f, (float)tokens, clip_sym, st);
  if (rc != LCB_OK) return rc;
  lcb_quant_spec sp = *spec;
  for (int i_s = 0; i_s < n_steps; ++i_s) {
    const float ratio = (float)(1.0 - (double)i_s / (double)n_grid);
    sp.clip_ratio = clip_sym ? -ratio : ratio;
    rc = lcb_fakequant(W, Q, nullptr, nullptr, R, C, 1, gs, &sp, 1, fqws, st);
    if (rc != LCB_OK) return rc;
    rc = launch_ac<T, false>((const T*)Q, (const T*)x, (T*)org, omax, omin, min_err, (T*)best_max, (T*)best_min, R, C, T_rows,
                             (int)gs, ratio, (float)tokens, clip_sym, st);
    if (rc != LCB_OK) return rc;
  }
  return LCB_OK;
}

}  // namespace
}  // namespace lcb

extern "C" int64_t lcb_autoclip_workspace_bytes(int64_t R, int64_t C, int64_t T_rows, int64_t group_size) {
  using namespace lcb;
  const int64_t gs = group_size > 0 ? group_size : C;
  if (gs < 1 || C % gs) return 0;
  const int64_t ng = C / gs, nu = R * ng;
  return align_up(R * C * 2, 256) + align_up(ng * R * T_rows * 2, 256) + align_up(nu * 8, 256) + 3 * align_up(nu * 4, 256) +
         lcb_fakequant_workspace_bytes() + 1024;
}

extern "C" int lcb_autoclip_search(const void* W, const void* x, void* best_max, void* best_min, int64_t R, int64_t C,
                                   int64_t T_rows, int64_t tokens, int32_t group_size, int32_t n_grid, int32_t n_steps,
                                   int32_t clip_sym, const lcb_quant_spec* spec, void* workspace, void* stream) {
  using namespace lcb;
  if (!W || !x || !best_max || !best_min || !spec || !workspace) return LCB_E_INVALID;
  if (R < 1 || C < 1 || T_rows < 1 || tokens < 1 || tokens > T_rows || n_grid < 1 || n_steps < 0 || n_steps > n_grid)
    return LCB_E_INVALID;
  if (spec->dtype != LCB_BF16 && spec->dtype != LCB_F16) return LCB_E_UNSUPPORTED;
  if (spec->out_mode != LCB_OUT_FAKE || spec->math_dtype != spec->dtype) return LCB_E_INVALID;
  const int64_t gs = group_size > 0 ? group_size : C;
  if (C % gs) return LCB_E_SHAPE;
  if (gs % 8 || (gs > 128 && gs % 128) || T_rows % 8) return LCB_E_UNSUPPORTED;   // tiles of <= 128 or k * 128 columns, 16-byte vectors
  uint8_t* ws = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~uintptr_t(255));
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (spec->dtype == LCB_BF16)
    return autoclip_typed<__nv_bfloat16>(W, x, best_max, best_min, R, C, T_rows, tokens, gs, n_grid, n_steps, clip_sym, spec, ws, st);
  return autoclip_typed<__half>(W, x, best_max, best_min, R, C, T_rows, tokens, gs, n_grid, n_steps, clip_sym, spec, ws, st);
}

// ---- the search taken apart (clip_version 'v2' and w+a configs: the host composes the per-step weights itself) ---------------
// lcb_autoclip_init : org_max / org_min per (row, group) in fp32, best_* = org_*, min_err = 1e9 (auto_clip.py:116-125)
// lcb_autoclip_step : is_org != 0: org[g, r, t] = (x * w).sum(-1) (auto_clip.py:151); else the candidate pass: err of
//                     (xq * Wq).sum(-1) against org, best_* / min_err updated where err < min_err (auto_clip.py:154-179).
extern "C" int lcb_autoclip_init(const void* W, void* best_max, void* best_min, float* org_max, float* org_min, float* min_err,
                                 int64_t R, int64_t C, int32_t group_size, int32_t clip_sym, int32_t dtype, void* stream) {
  using namespace lcb;
  if (!W || !best_max || !best_min || !org_max || !org_min || !min_err) return LCB_E_INVALID;
  if (R < 1 || C < 1) return LCB_E_INVALID;
  if (dtype != LCB_BF16 && dtype != LCB_F16) return LCB_E_UNSUPPORTED;
  const int64_t gs = group_size > 0 ? group_size : C;
  if (C % gs) return LCB_E_SHAPE;
  const int64_t nu = R * (C / gs);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int grid = (int)((nu * 32 + 255) / 256);
  if (dtype == LCB_BF16)
    ac_init_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)W, (__nv_bfloat16*)best_max, (__nv_bfloat16*)best_min,
                                                        org_max, org_min, min_err, nu, (int)gs, clip_sym);
  else
    ac_init_kernel<__half><<<grid, 256, 0, st>>>((const __half*)W, (__half*)best_max, (__half*)best_min, org_max, org_min, min_err,
                                                 nu, (int)gs, clip_sym);
  LCB_CUDA_CHECK_LAUNCH("ac_init");
  return LCB_OK;
}

extern "C" int lcb_autoclip_step(const void* Wq, const void* x, void* org, const float* org_max, const float* org_min,
                                 float* min_err, void* best_max, void* best_min, int64_t R, int64_t C, int64_t T_rows,
                                 int64_t tokens, int32_t group_size, float ratio, int32_t clip_sym, int32_t is_org, int32_t dtype,
                                 void* stream) {
  using namespace lcb;
  if (!Wq || !x || !org) return LCB_E_INVALID;
  if (!is_org && (!org_max || !org_min || !min_err || !best_max || !best_min)) return LCB_E_INVALID;
  if (R < 1 || C < 1 || T_rows < 1 || tokens < 1 || tokens > T_rows) return LCB_E_INVALID;
  if (dtype != LCB_BF16 && dtype != LCB_F16) return LCB_E_UNSUPPORTED;
  const int64_t gs = group_size > 0 ? group_size : C;
  if (C % gs) return LCB_E_SHAPE;
  if (gs % 8 || (gs > 128 && gs % 128) || T_rows % 8) return LCB_E_UNSUPPORTED;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (dtype == LCB_BF16) {
    using T = __nv_bfloat16;
    return is_org ? launch_ac<T, true>((const T*)Wq, (const T*)x, (T*)org, org_max, org_min, min_err, (T*)best_max, (T*)best_min, R, C,
                                       T_rows, (int)gs, 1.f, (float)tokens, clip_sym, st)
                  : launch_ac<T, false>((const T*)Wq, (const T*)x, (T*)org, org_max, org_min, min_err, (T*)best_max, (T*)best_min, R,
                                        C, T_rows, (int)gs, ratio, (float)tokens, clip_sym, st);
  }
  using T = __half;
  return is_org ? launch_ac<T, true>((const T*)Wq, (const T*)x, (T*)org, org_max, org_min, min_err, (T*)best_max, (T*)best_min, R, C,
                                     T_rows, (int)gs, 1.f, (float)tokens, clip_sym, st)
                : launch_ac<T, false>((const T*)Wq, (const T*)x, (T*)org, org_max, org_min, min_err, (T*)best_max, (T*)best_min, R, C,
                                      T_rows, (int)gs, ratio, (float)tokens, clip_sym, st);
}
