// gptq.cu — GPTQ blocked column update (gptq.py:199-244) as three kernels per column block:
//
//   gptq_group_qparams_kernel : per-row (min, max) -> (scale, zero) of every quantization group that STARTS inside the
//       block, searched on the outer W exactly where the reference does (gptq.py:215-223: `W[:, i1+i : i1+i+gs]`, i.e.
//       columns that have received every earlier block's lazy update but none of this block's in-block updates).
//   gptq_block_kernel : the serial part. 8 lanes own one weight row; lane l keeps columns l, l+8, ... of the 128-column
//       block in registers. Step i: broadcast w = W1[:, i] inside the 8-lane group, quant-dequant it (fp32, the literal
//       reference chain), err = (w - q) / d, then every lane updates its remaining columns
//       W1[:, j] -= err * Hinv1[i, j] (separate multiply and subtract, as the reference's outer-product + sub).
//       With identical inputs this kernel is bit-identical to the reference's in-block loop. Err leaves as bf16 hi + lo.
//   lazy-batch updates on tensor cores (tc_gemm.cuh, D -= A * B with bf16 hi/lo splits of Err and Hinv):
//       level 1, after each block : W[:, i2 : s2] -= Err[:, i1:i2] * Hinv[i1:i2, i2:s2]     (inside a 1024-wide super-block)
//       level 2, after each super-block : W[:, s2 : C] -= Err[:, s1:s2] * Hinv[s1:s2, s2:C]  (K = 1024)
//   The reference applies every block to ALL trailing columns (R*C^2/256*8 bytes of W traffic, 26 GB for Llama-3-8B
//   down_proj); the two-level schedule is the same sum with K = 1024 for the far part (5 GB).
#include <algorithm>
#include "fq_math.cuh"
#include "tc_gemm.cuh"

namespace lcb {
namespace {

struct GSpec {
  float qmin, qmax;
  int sym, round_zp;
};

// ---- per-row group qparams (fp32 literal chain: quant.py:132-143, 545-559) --------------------------------------------
// one warp per (row, group); groups g0 .. g0+ng-1 start at column g * gs and span min(gs, n_nonout - g*gs) columns
__global__ void __launch_bounds__(256)
    gptq_group_qparams_kernel(const float* __restrict__ W, float* __restrict__ gscales, float* __restrict__ gzeros,
                              int64_t R, int64_t C, int64_t n_nonout, int gs, int g0, int ng, SpecD sp) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (warp >= R * ng) return;
  const int64_t row = warp / ng;
  const int g = g0 + (int)(warp % ng);
  const int64_t c0 = (int64_t)g * gs;
  const int64_t span = min((int64_t)gs, n_nonout - c0);
  float mx = -INFINITY, mn = INFINITY;
  for (int64_t c = lane; c < span; c += 32) {
    const float x = W[row * C + c0 + c];
    mx = max_nan(mx, x);
    mn = min_nan(mn, x);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    mx = max_nan(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    mn = min_nan(mn, __shfl_xor_sync(0xffffffffu, mn, o));
  }
  if (lane == 0) {
    UnitQ u;
    qparams_from_range(mn, mx, sp, u);
    gscales[(int64_t)g * R + row] = u.s;
    if (gzeros) gzeros[(int64_t)g * R + row] = u.z;
  }
}

// ---- IntegerQuantizer.quant_dequant of one fp32 value (quant.py:699-717), math dtype fp32 ------------------------------
// x / s is RN(x / s): by a Markstein step on r = RN(1 / s) when (s, x) are inside the window where that step is exact
// (tests/test_fq_math_identity.py), by the IEEE division otherwise. `fast` is per (row, group).
__device__ __noinline__ float div_ieee(float x, float s) { return __fdiv_rn(x, s); }   // cold: kept out of the step body
__device__ __forceinline__ float div_exact(float x, float s, float r, bool fast) {
  if (fast && (fabsf(x) >= 1e-30f || x == 0.f) && fabsf(x) <= 1e30f) {
    const float q0 = __fmul_rn(x, r);
    return __fmaf_rn(__fmaf_rn(-s, q0, x), r, q0);
  }
  return div_ieee(x, s);
}
__device__ __forceinline__ float qdq_f32(float x, float s, float r, bool fast, float z, const GSpec& g) {
  float q;
  if (g.round_zp) {
    q = clamp_nan(__fadd_rn(rintf(div_exact(x, s, r, fast)), z), g.qmin, g.qmax);
  } else {
    q = clamp_nan(rintf(__fadd_rn(div_ieee(x, max_nan(s, 1e-9f)), z)), g.qmin, g.qmax);
  }
  return __fmul_rn(__fsub_rn(q, z), s);
}

constexpr int LPR = 8;    // lanes per weight row
constexpr int NK = 16;    // columns per lane (block of up to 128 columns)
constexpr int BSMAX = LPR * NK;
// Hinv1 in shared memory, laid out for the update loop: row i holds, for every lane l, the NK values Hinv1[i][kk * LPR + l]
// (kk = 0 .. NK-1) contiguously, lane blocks HLANE floats apart (20: the 8 lanes of a row hit distinct bank groups with
// 128-bit loads). One step then reads its remaining columns with (NK - k) / 4 LDS.128 instead of NK - k scalar loads.
constexpr int HLANE = NK + 4, HROW = LPR * HLANE;
constexpr int BLOCK_SMEM = BSMAX * HROW * 4 + BSMAX * 12;  // Hinv1 + diag + 1/diag + per-step group index

// group switch (rare, CTA-uniform): scale / zero / reciprocal of the new group; out of line to keep the step body small
struct GroupQ {
  float s, z, rs;
  int fast;
};
__device__ __noinline__ GroupQ load_group(const float* gscales, const float* gzeros, int64_t off, int sym) {
  GroupQ q;
  q.s = gscales[off];
  q.z = (gzeros && !sym) ? gzeros[off] : 0.f;
  q.fast = (q.s >= 1e-20f) && (q.s <= 1e20f);
  q.rs = q.fast ? __frcp_rn(q.s) : 1.f;
  return q;
}
__device__ __noinline__ GroupQ make_group(float s, float z) {
  GroupQ q;
  q.s = s;
  q.z = z;
  q.fast = (s >= 1e-20f) && (s <= 1e20f);
  q.rs = q.fast ? __frcp_rn(s) : 1.f;
  return q;
}

// The rows of one warp (4 rows x 8 lanes) through the 128 column steps of a block.
// SPEC = true : speculative pass. Both divisions are Markstein steps without guards; a flag records whether any operand
//               left the window in which that step equals the IEEE division (or a scale / diagonal had no usable
//               reciprocal). If the flag is set anywhere in the warp nothing is stored and `true` is returned.
// SPEC = false: the literal pass (guards + IEEE fallbacks), always stores.
template <bool FUSED_QP, bool SPEC>
__device__ __forceinline__ bool gptq_rows(const float* __restrict__ W, float* __restrict__ Out, float* __restrict__ Losses,
                                          __nv_bfloat16* __restrict__ err_hi, __nv_bfloat16* __restrict__ err_lo,
                                          float* __restrict__ gscales, float* __restrict__ gzeros, int64_t R, int64_t C,
                                          int i1, int count, int err_col0, int err_ld, int group_size, const GSpec& gsp,
                                          const SpecD& sp, const float* Hs, const float* dg, const float* rdiag,
                                          const int* gidx) {
  const int lane = threadIdx.x & 31;
  const int l = lane & (LPR - 1);
  const int64_t row_raw = (int64_t)blockIdx.x * 32 + (threadIdx.x / LPR);
  const bool row_ok = row_raw < R;
  const int64_t row = row_ok ? row_raw : R - 1;  // keep every lane in the shuffles; stores are masked

  float reg[NK], tmpv[NK], diffv[NK], errv[NK];
#pragma unroll
  for (int k = 0; k < NK; ++k) {
    const int j = k * LPR + l;
    reg[k] = j < count ? W[row * C + i1 + j] : 0.f;
    tmpv[k] = diffv[k] = errv[k] = 0.f;
  }
  float sq[4] = {1.f, 1.f, 1.f, 1.f}, zq[4] = {0.f, 0.f, 0.f, 0.f};  // FUSED_QP: qparams of the (<= 4) groups of this block
  if (FUSED_QP) {
    // groups of gk = group_size / 8 register slots; after the xor-shuffles every lane of the row holds the group's
    // range and computes the qparams itself (registers only: no global round trip inside the kernel)
    const int gk = group_size / LPR;
#pragma unroll
    for (int gi = 0; gi < 4; ++gi) {
      const int k0 = gi * gk;
      if (k0 * LPR >= count) break;  // uniform
      float mx = -INFINITY, mn = INFINITY;
#pragma unroll
      for (int k = 0; k < NK; ++k) {
        if (k >= k0 && k < k0 + gk && k * LPR + l < count) {
          mx = max_nan(mx, reg[k]);
          mn = min_nan(mn, reg[k]);
        }
      }
#pragma unroll
      for (int o = LPR / 2; o > 0; o >>= 1) {
        mx = max_nan(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        mn = min_nan(mn, __shfl_xor_sync(0xffffffffu, mn, o));
      }
      UnitQ u;
      qparams_from_range(mn, mx, sp, u);
      sq[gi] = u.s;
      zq[gi] = u.z;
      if (l == 0 && row_ok) {
        const int g = gidx[k0 * LPR];
        gscales[(int64_t)g * R + row] = u.s;
        if (gzeros && !gsp.sym) gzeros[(int64_t)g * R + row] = u.z;
      }
    }
  }
  int cur_g = -1;
  const int g_first = gidx[0];
  float s = 1.f, z = 0.f, rs = 1.f;
  bool fast = false;
  bool bad = false;
#pragma unroll
  for (int k = 0; k < NK; ++k) {
#pragma unroll 1
    for (int l0 = 0; l0 < LPR; ++l0) {
      const int i = k * LPR + l0;
      if (i >= count) break;  // uniform
      const float w = __shfl_sync(0xffffffffu, reg[k], (lane & ~(LPR - 1)) | l0);
      const int g = gidx[i];
      if (g != cur_g) {  // uniform across the CTA, rare
        cur_g = g;
        GroupQ gq;
        if (FUSED_QP) {
          const int gi = g - g_first;   // groups of a fused block are consecutive
          gq = make_group(gi == 0 ? sq[0] : gi == 1 ? sq[1] : gi == 2 ? sq[2] : sq[3],
                          gi == 0 ? zq[0] : gi == 1 ? zq[1] : gi == 2 ? zq[2] : zq[3]);
        } else {
          gq = load_group(gscales, gzeros, (int64_t)g * R + row, gsp.sym);
        }
        s = gq.s;
        z = gq.z;
        rs = gq.rs;
        fast = gq.fast != 0;
        if (SPEC) bad |= !fast;
      }
      const float d = dg[i];
      const float rd = rdiag[i];
      float q, diff, err;
      if (SPEC) {
        const float q0 = __fmul_rn(w, rs);
        float t = __fmaf_rn(__fmaf_rn(-s, q0, w), rs, q0);                       // RN(w / s) inside the window
        t = clamp_nan(__fadd_rn(rintf(t), z), gsp.qmin, gsp.qmax);
        q = __fmul_rn(__fsub_rn(t, z), s);
        diff = __fsub_rn(w, q);
        const float e0 = __fmul_rn(diff, rd);
        err = __fmaf_rn(__fmaf_rn(-d, e0, diff), rd, e0);                          // RN(diff / d) inside the window
        const float aw = fabsf(w), ad = fabsf(diff);
        bad |= !((aw >= 1e-30f || w == 0.f) && aw <= 1e30f) | !((ad >= 1e-30f || diff == 0.f) && ad <= 1e30f) | (rd == 0.f);
      } else {
        q = qdq_f32(w, s, rs, fast, z, gsp);
        diff = __fsub_rn(w, q);
        err = div_exact(diff, d, rd, rd != 0.f);   // (w - q) / d
      }
      if (l == l0) {
        tmpv[k] = w;
        diffv[k] = diff;
        errv[k] = err;
      }
      // W1[:, j] -= err * Hinv1[i, j] for the columns j > i this lane owns: separate multiply and subtract (the
      // reference's outer product + sub). 128-bit loads of the lane's contiguous Hinv1 slice.
      const float4* h4 = reinterpret_cast<const float4*>(Hs + i * HROW + l * HLANE);
#pragma unroll
      for (int c4 = k / 4; c4 < NK / 4; ++c4) {
        const float4 h = h4[c4];
        const float hv[4] = {h.x, h.y, h.z, h.w};
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          const int kk = c4 * 4 + t;
          if (kk > k || (kk == k && l > l0)) reg[kk] = __fsub_rn(reg[kk], __fmul_rn(err, hv[t]));
        }
      }
    }
  }
  if (SPEC && __any_sync(0xffffffffu, bad)) return true;
  if (!row_ok) return false;
#pragma unroll
  for (int k = 0; k < NK; ++k) {
    const int j = k * LPR + l;
    if (j < count) {
      const float d = dg[j];
      Out[row * C + i1 + j] = tmpv[k];
      Losses[row * C + i1 + j] = __fdiv_rn(__fmul_rn(diffv[k], diffv[k]), __fmul_rn(2.f, __fmul_rn(d, d)));  // (w-q)^2 / (2 d^2)
      const __nv_bfloat16 h = __float2bfloat16_rn(errv[k]);
      err_hi[row * err_ld + err_col0 + j] = h;
      err_lo[row * err_ld + err_col0 + j] = __float2bfloat16_rn(__fsub_rn(errv[k], __bfloat162float(h)));
    }
  }
  return false;
}

// grid: ceil(R / 32) CTAs of 256 threads (32 rows); dynamic smem: BLOCK_SMEM
// FUSED_QP: the quantization groups of this block start inside it, are aligned to multiples of 8 columns and end inside
// it, so their (min, max) are taken from the block's own (not yet updated) register copy — exactly the outer W the
// reference searches (gptq.py:215-223) — and no separate qparams kernel runs.
template <bool FUSED_QP>
__global__ void __launch_bounds__(256)
    gptq_block_kernel(const float* __restrict__ W, const float* __restrict__ Hinv, float* __restrict__ Out,
                      float* __restrict__ Losses, __nv_bfloat16* __restrict__ err_hi, __nv_bfloat16* __restrict__ err_lo,
                      float* __restrict__ gscales, float* __restrict__ gzeros,
                      const int32_t* __restrict__ col_group, int64_t R, int64_t C, int i1, int count, int err_col0,
                      int err_ld, int group_size, GSpec gsp, SpecD sp) {
  extern __shared__ float Hs[];  // [BSMAX][HROW] (see HLANE), rows/cols >= count are zero
  float* dg = Hs + BSMAX * HROW;                          // Hinv1[i, i]
  float* rdiag = dg + BSMAX;                              // RN(1 / Hinv1[i, i])
  int* gidx = reinterpret_cast<int*>(rdiag + BSMAX);      // group of column i1 + i
  {  // Hinv1 tile -> shared memory: 16 loads in flight per thread before the first store (a plain copy loop pays the L2
     // latency once per element)
    constexpr int PER = BSMAX * BSMAX / 256, B = 16;
    const float* hbase = Hinv + (int64_t)i1 * C + i1;
#pragma unroll 1
    for (int b0 = 0; b0 < PER; b0 += B) {
      float v[B];
#pragma unroll
      for (int u = 0; u < B; ++u) {
        const int idx = (b0 + u) * 256 + threadIdx.x;
        const int r = idx / BSMAX, c = idx % BSMAX;
        v[u] = (r < count && c < count) ? hbase[(int64_t)r * C + c] : 0.f;
      }
#pragma unroll
      for (int u = 0; u < B; ++u) {
        const int idx = (b0 + u) * 256 + threadIdx.x;
        const int r = idx / BSMAX, c = idx % BSMAX;
        Hs[r * HROW + (c % LPR) * HLANE + c / LPR] = v[u];
      }
    }
  }
  if (threadIdx.x < BSMAX) {
    const int i = threadIdx.x;
    const float d = i < count ? Hinv[(int64_t)(i1 + i) * C + i1 + i] : 1.f;
    dg[i] = d;
    rdiag[i] = (fabsf(d) >= 1e-20f && fabsf(d) <= 1e20f) ? __frcp_rn(d) : 0.f;  // 0: take the IEEE division
    const int col = i1 + (i < count ? i : count - 1);
    gidx[i] = col_group ? col_group[col] : (group_size > 0 ? col / group_size : 0);
  }
  __syncthreads();
  bool redo = true;
  if (gsp.round_zp)   // speculative pass: branch-free arithmetic, valid when every operand stays inside the exactness windows
    redo = gptq_rows<FUSED_QP, true>(W, Out, Losses, err_hi, err_lo, gscales, gzeros, R, C, i1, count, err_col0, err_ld,
                                     group_size, gsp, sp, Hs, dg, rdiag, gidx);
  if (redo)           // warp-uniform, rare: the literal pass (IEEE divisions where a window is left)
    gptq_rows<FUSED_QP, false>(W, Out, Losses, err_hi, err_lo, gscales, gzeros, R, C, i1, count, err_col0, err_ld,
                               group_size, gsp, sp, Hs, dg, rdiag, gidx);
}

__global__ void __launch_bounds__(256) split_f32_2d_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ hi,
                                                           __nv_bfloat16* __restrict__ lo, int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float v = x[i];
    const __nv_bfloat16 h = __float2bfloat16_rn(v);
    hi[i] = h;
    lo[i] = __float2bfloat16_rn(v - __bfloat162float(h));
  }
}

inline int64_t align_up(int64_t x, int64_t a) { return (x + a - 1) / a * a; }
inline int super_block(int blocksize) { return blocksize * std::max(1, 1024 / blocksize); }

}  // namespace
}  // namespace lcb

extern "C" int64_t lcb_gptq_workspace_bytes(int64_t R, int64_t C, int32_t blocksize) {
  if (blocksize < 1) return 0;
  const int64_t sb = lcb::super_block(blocksize);
  return 2 * lcb::align_up(C * C * 2, 256) + 2 * lcb::align_up(R * sb * 2, 256) + 512;
}

extern "C" int lcb_gptq_weight_transform(float* W, const float* Hinv, float* Out, float* Losses, float* gscales,
                                         float* gzeros, const int32_t* col_group, int64_t R, int64_t C, int64_t n_nonout,
                                         int32_t blocksize, int32_t group_size, int32_t static_groups,
                                         const lcb_quant_spec* spec, void* workspace, void* stream) {
  using namespace lcb;
  if (!W || !Hinv || !Out || !Losses || !gscales || !spec || !workspace) return LCB_E_INVALID;
  if (R < 1 || C < 1 || n_nonout < 0 || n_nonout > C) return LCB_E_INVALID;
  if (spec->kind != LCB_KIND_INT) return LCB_E_UNSUPPORTED;        // FloatQuantizer inside GPTQ: not on the CUDA path
  if (blocksize < 8 || blocksize > BSMAX || blocksize % 8) return LCB_E_UNSUPPORTED;
  if (!(spec->qmax > spec->qmin)) return LCB_E_INVALID;
  if (!spec->sym && !gzeros) return LCB_E_INVALID;
  if (C % 8 || (reinterpret_cast<uintptr_t>(W) & 15) || (reinterpret_cast<uintptr_t>(Hinv) & 15)) return LCB_E_ALIGN;
  const bool dynamic_groups = group_size > 0 && !static_groups;
  if (dynamic_groups && col_group) return LCB_E_INVALID;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);

  const int sb = super_block(blocksize);
  uint8_t* ws = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~uintptr_t(255));
  __nv_bfloat16* hinv_hi = reinterpret_cast<__nv_bfloat16*>(ws);
  __nv_bfloat16* hinv_lo = reinterpret_cast<__nv_bfloat16*>(ws + align_up(C * C * 2, 256));
  __nv_bfloat16* err_hi = reinterpret_cast<__nv_bfloat16*>(ws + 2 * align_up(C * C * 2, 256));
  __nv_bfloat16* err_lo = reinterpret_cast<__nv_bfloat16*>(ws + 2 * align_up(C * C * 2, 256) + align_up(R * sb * 2, 256));

  static DeviceOnce attr_once;
  if (attr_once.needed()) {
    if (cudaFuncSetAttribute(gptq_block_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, BLOCK_SMEM) != cudaSuccess ||
        cudaFuncSetAttribute(gptq_block_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, BLOCK_SMEM) != cudaSuccess) {
      set_last_error("cudaFuncSetAttribute(gptq_block)", cudaGetLastError());
      return LCB_E_CUDA;
    }
    attr_once.done();
  }
  {
    const int64_t n = C * C;
    split_f32_2d_kernel<<<(int)std::min<int64_t>((n + 255) / 256, kNumSMs * 16), 256, 0, st>>>(Hinv, hinv_hi, hinv_lo, n);
    LCB_CUDA_CHECK_LAUNCH("gptq_split_hinv");
  }
  GSpec gsp;
  gsp.qmin = spec->qmin;
  gsp.qmax = spec->qmax;
  gsp.sym = spec->sym;
  gsp.round_zp = spec->round_zp;
  SpecD sp;  // fp32 math everywhere: W is fp32 (gptq.py:137)
  sp.dtype = sp.math_dtype = sp.qparam_dtype = sp.qp_math = LCB_F32;
  sp.kind = LCB_KIND_INT;
  sp.sym = spec->sym;
  sp.round_zp = spec->round_zp;
  sp.out_mode = LCB_OUT_NONE;
  sp.fast_ok = 0;
  sp.qmin = spec->qmin;
  sp.qmax = spec->qmax;
  sp.osf = 1.0f;
  sp.clip = 1.0f;

  const int row_ctas = (int)((R + 31) / 32);
  for (int64_t s1 = 0; s1 < n_nonout; s1 += sb) {
    const int64_t s2 = std::min<int64_t>(s1 + sb, n_nonout);
    for (int64_t i1 = s1; i1 < s2; i1 += blocksize) {
      const int64_t i2 = std::min<int64_t>(i1 + blocksize, s2);
      const int count = (int)(i2 - i1);
      // groups that start in this block; fused into the block kernel when they also end in it and are 8-aligned
      bool fused = false;
      if (dynamic_groups) {
        const int g0 = (int)((i1 + group_size - 1) / group_size);       // first group starting at or after i1
        const int g1 = (int)((i2 + group_size - 1) / group_size);       // first group starting at or after i2
        const int ng = g1 - g0;
        fused = (group_size % LPR == 0) && (i1 % group_size == 0) && (count % group_size == 0 || i2 == n_nonout) &&
                group_size <= count && (count + group_size - 1) / group_size <= 4;
        if (ng > 0 && !fused) {
          const int64_t warps = R * ng;
          gptq_group_qparams_kernel<<<(int)((warps + 7) / 8), 256, 0, st>>>(W, gscales, spec->sym ? nullptr : gzeros, R, C,
                                                                           n_nonout, group_size, g0, ng, sp);
          LCB_CUDA_CHECK_LAUNCH("gptq_group_qparams");
        }
      }
      if (fused)
        gptq_block_kernel<true><<<row_ctas, 256, BLOCK_SMEM, st>>>(W, Hinv, Out, Losses, err_hi, err_lo, gscales, gzeros,
                                                                  col_group, R, C, (int)i1, count, (int)(i1 - s1), sb,
                                                                  group_size, gsp, sp);
      else
        gptq_block_kernel<false><<<row_ctas, 256, BLOCK_SMEM, st>>>(W, Hinv, Out, Losses, err_hi, err_lo, gscales, gzeros,
                                                                   col_group, R, C, (int)i1, count, (int)(i1 - s1), sb,
                                                                   group_size, gsp, sp);
      LCB_CUDA_CHECK_LAUNCH("gptq_block");
      if (i2 < s2) {  // level 1: rest of the super-block
        const void* aa[3] = {err_hi + (i1 - s1), err_hi + (i1 - s1), err_lo + (i1 - s1)};
        const void* bb[3] = {hinv_hi + i1 * C + i2, hinv_lo + i1 * C + i2, hinv_hi + i1 * C + i2};
        int rc = tc_gemm_launch(true, false, true, 1, aa, bb, 3, W + i2, R, s2 - i2, count, sb, C, C, st);
        if (rc != LCB_OK) return rc;
      }
    }
    if (s2 < C) {  // level 2: everything right of the super-block (including OWQ outlier columns)
      const void* aa[3] = {err_hi, err_hi, err_lo};
      const void* bb[3] = {hinv_hi + s1 * C + s2, hinv_lo + s1 * C + s2, hinv_hi + s1 * C + s2};
      int rc = tc_gemm_launch(true, false, true, 1, aa, bb, 3, W + s2, R, C - s2, s2 - s1, sb, C, C, st);
      if (rc != LCB_OK) return rc;
    }
  }
  return LCB_OK;
}
