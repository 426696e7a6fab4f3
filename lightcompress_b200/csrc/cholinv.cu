// cholinv.cu — U = cholesky(inverse(H + damp I), upper): the factor GPTQ.process_hessian_and_weights builds with three
// cuSOLVER calls (gptq.py:165-174: H = cholesky(H); H = cholesky_inverse(H); H = cholesky(H, upper=True)).
//
// Identity used: let H = R R^T with R UPPER triangular (the "reverse" Cholesky factor, obtained by running the
// factorisation from the last row/column upwards). Then inverse(H) = R^-T R^-1 = (R^-1)^T (R^-1) and R^-1 is upper
// triangular with a positive diagonal, so by uniqueness of the Cholesky factor U = R^-1. One factorisation (n^3/3) and
// one triangular inverse (n^3/3) replace potrf + potri + potrf (4 n^3/3), and every O(n^3) term is a tcgen05 GEMM.
//
// Phase 1, block column k = nb-1 .. 0 (128 wide), in place on the upper triangle of H:
//   chol_diag_kernel : R_kk = reverse Cholesky of the 128 x 128 diagonal block (one CTA, fp32)
//   chol_trsm_kernel : R[0:k, k] = H[0:k, k] R_kk^-T  (8 lanes per row, fp32 substitution in registers) + bf16 splits
//   tc_gemm (SYRK)   : H[0:k, 0:k] -= R[0:k, k] R[0:k, k]^T        (upper 128-blocks only; K = 128)
// Phase 2, block row i = nb-1 .. 0:
//   tri_inv_diag_kernel : U_ii = R_ii^-1 for all i at once (one CTA per block)
//   tc_gemm             : T = R[i, i+1:] U[i+1:, i+1:]                (M = 128, split-K, triangular K range)
//   tri_row_kernel      : U[i, i+1:] = -U_ii T, U[i, i] = U_ii        (fp32) + bf16 splits, in place over R's row
// fp32 operands enter the tensor cores as NS bf16 pieces (NS = 2: hi + lo, 3 products, ~2^-17 per product; NS = 3:
// hi + mid + lo, 6 products, ~fp32). Everything else (diagonal blocks, substitutions, accumulators) is fp32.
#include <algorithm>
#include "tc_gemm.cuh"

namespace lcb {
namespace {

constexpr int NB = 128;
constexpr int LDS_ = NB + 1;  // padded smem row

struct Splits {
  __nv_bfloat16* s[3];
  int ns;
};

__device__ __forceinline__ void store_splits(const Splits& sp, int64_t idx, float v) {
  const __nv_bfloat16 h = __float2bfloat16_rn(v);
  sp.s[0][idx] = h;
  const float r1 = v - __bfloat162float(h);
  const __nv_bfloat16 m = __float2bfloat16_rn(r1);
  sp.s[1][idx] = m;
  if (sp.ns > 2) sp.s[2][idx] = __float2bfloat16_rn(r1 - __bfloat162float(m));
}

// 128 x 128 fp32 block global -> shared, NT threads, loads issued in batches of 16 before any store (a plain copy loop
// serialises on the load latency). dst index = f(r, c).
template <int NT, typename F>
__device__ __forceinline__ void load_block(const float* __restrict__ base, int64_t ld, F&& put) {
  constexpr int PER = NB * NB / NT;
  static_assert(PER % 16 == 0 || PER < 16, "batching");
  constexpr int B = PER < 16 ? PER : 16;
#pragma unroll 1
  for (int b0 = 0; b0 < PER; b0 += B) {
    float v[B];
#pragma unroll
    for (int u = 0; u < B; ++u) {
      const int idx = (b0 + u) * NT + threadIdx.x;
      v[u] = base[(int64_t)(idx >> 7) * ld + (idx & 127)];
    }
#pragma unroll
    for (int u = 0; u < B; ++u) {
      const int idx = (b0 + u) * NT + threadIdx.x;
      put(idx >> 7, idx & 127, v[u]);
    }
  }
}

// ---- H[i, i] += percdamp * mean(diag H)   (gptq.py:165-167); one CTA -------------------------------------------------
__global__ void __launch_bounds__(1024) chol_damp_kernel(float* __restrict__ H, int64_t n, int64_t ld, float percdamp) {
  __shared__ float red[32];
  __shared__ float damp_s;
  float s = 0.f;
  for (int64_t i = threadIdx.x; i < n; i += 1024) s += H[i * ld + i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x < 32) {
    s = red[threadIdx.x];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (threadIdx.x == 0) damp_s = percdamp * (s / (float)n);
  }
  __syncthreads();
  const float d = damp_s;
  for (int64_t i = threadIdx.x; i < n; i += 1024) H[i * ld + i] += d;
}

// ---- padded copy in: Hp = [[H, 0], [0, I]] (upper triangle is all that is read) ---------------------------------------
__global__ void __launch_bounds__(256) chol_pad_in_kernel(const float* __restrict__ H, float* __restrict__ Hp, int64_t n,
                                                          int64_t ld, int64_t np) {
  const int64_t total = np * np;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (int64_t)gridDim.x * 256) {
    const int64_t r = i / np, c = i % np;
    Hp[i] = (r < n && c < n) ? H[r * ld + c] : (r == c ? 1.f : 0.f);
  }
}
__global__ void __launch_bounds__(256) chol_pad_out_kernel(float* __restrict__ H, const float* __restrict__ Hp, int64_t n,
                                                           int64_t ld, int64_t np) {
  const int64_t total = n * n;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (int64_t)gridDim.x * 256) {
    const int64_t r = i / n, c = i % n;
    H[r * ld + c] = c >= r ? Hp[r * np + c] : 0.f;
  }
}
// zero the strict lower triangle (the reference's factor is exactly upper triangular)
__global__ void __launch_bounds__(256) chol_zero_lower_kernel(float* __restrict__ H, int64_t n, int64_t ld) {
  const int64_t r = blockIdx.y;
  for (int64_t c = (int64_t)blockIdx.x * 256 + threadIdx.x; c < r; c += (int64_t)gridDim.x * 256) H[r * ld + c] = 0.f;
}

// ---- reverse Cholesky of one 128 x 128 diagonal block: A = R R^T, R upper -------------------------------------------
// One CTA of 512 threads, the block in shared memory, four 32-column panels from the last to the first:
//   (a) warp 0 factors the 32 x 32 diagonal sub-block in registers (lane i = row i; step j = 31 .. 0 broadcasts the
//       pivot and the column by shuffles): R[j][j] = sqrt(a[j][j]), R[i][j] = a[i][j] / R[j][j], a[i][i'] -= R[i][j] R[i'][j];
//   (b) one thread per row above it solves x Rpp^T = a[row, panel] by back substitution (32 registers, Rpp broadcast);
//   (c) all threads retire the rank-32 term from the leading block, 4 x 4 register tiles over interleaved rows.
constexpr int PW = 32, XLD = PW + 1;
__global__ void __launch_bounds__(512) chol_diag_kernel(float* __restrict__ H, Splits sp, int64_t ld, int kb,
                                                        int* __restrict__ info) {
  extern __shared__ float smf[];
  float* A = smf;                        // [NB][LDS_]
  float* X = smf + NB * LDS_;            // [NB - PW][XLD]  solved panel rows
  float* Rp = X + (NB - PW) * XLD;       // [PW][XLD]       diagonal sub-block factor
  float* base = H + (int64_t)kb * NB * ld + (int64_t)kb * NB;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  load_block<512>(base, ld, [&](int r, int c, float v) { A[r * LDS_ + c] = v; });
  __syncthreads();
#pragma unroll 1
  for (int p = NB / PW - 1; p >= 0; --p) {
    const int c0 = p * PW;
    if (warp == 0) {
      float d[PW];
#pragma unroll
      for (int j = 0; j < PW; ++j) d[j] = A[(c0 + lane) * LDS_ + c0 + j];
#pragma unroll
      for (int j = PW - 1; j >= 0; --j) {
        const float ajj = __shfl_sync(0xffffffffu, d[j], j);
        if (lane == 0 && !(ajj > 0.f)) atomicCAS(info, 0, kb * NB + c0 + j + 1);
        // 1 / sqrt(a) by MUFU.RSQ + one Newton step (<= 1 ulp), so the serial chain has no IEEE sqrt / division
        float y = rsqrtf(ajj);
        y = __fmaf_rn(y, __fmaf_rn(-0.5f * ajj * y, y, 0.5f), y);
        const float rjj = __fmaf_rn(__fmaf_rn(-ajj * y, ajj * y, ajj), 0.5f * y, ajj * y);   // sqrt(a) = a y, refined
        const float c = lane < j ? d[j] * y : 0.f;
        d[j] = lane == j ? rjj : c;
        // entries below the diagonal (ip < lane) receive junk updates; they are never read
#pragma unroll
        for (int ip = 0; ip < j; ++ip) d[ip] = __fmaf_rn(-c, __shfl_sync(0xffffffffu, c, ip), d[ip]);
      }
#pragma unroll
      for (int j = 0; j < PW; ++j) {
        const float v = j >= lane ? d[j] : 0.f;
        A[(c0 + lane) * LDS_ + c0 + j] = v;
        Rp[lane * XLD + j] = v;
      }
    }
    __syncthreads();
    if (p == 0) break;
    if (tid < c0) {
      float h[PW];
#pragma unroll
      for (int j = 0; j < PW; ++j) h[j] = A[tid * LDS_ + c0 + j];
#pragma unroll
      for (int j = PW - 1; j >= 0; --j) {
        const float x = __fdiv_rn(h[j], Rp[j * XLD + j]);
        h[j] = x;
#pragma unroll
        for (int i = 0; i < j; ++i) h[i] = __fmaf_rn(-x, Rp[i * XLD + j], h[i]);
      }
#pragma unroll
      for (int j = 0; j < PW; ++j) {
        A[tid * LDS_ + c0 + j] = h[j];
        X[tid * XLD + j] = h[j];
      }
    }
    __syncthreads();
    const int q = c0 / 4;  // rows of one interleave class
    for (int t = tid; t < q * q; t += 512) {
      const int ti = t / q, tj = t % q;
      float acc[4][4];
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;
#pragma unroll 8
      for (int k = 0; k < PW; ++k) {
        float xi[4], xj[4];
#pragma unroll
        for (int a = 0; a < 4; ++a) {
          xi[a] = X[(ti + a * q) * XLD + k];
          xj[a] = X[(tj + a * q) * XLD + k];
        }
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
          for (int b = 0; b < 4; ++b) acc[a][b] = __fmaf_rn(xi[a], xj[b], acc[a][b]);
      }
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) {
          const int i = ti + a * q, ip = tj + b * q;
          if (ip >= i) A[i * LDS_ + ip] -= acc[a][b];
        }
    }
    __syncthreads();
  }
  const int64_t sbase = (int64_t)kb * NB * ld + (int64_t)kb * NB;
  for (int idx = tid; idx < NB * NB; idx += 512) {
    const int r = idx >> 7, c = idx & 127;
    const float v = c >= r ? A[r * LDS_ + c] : 0.f;
    base[(int64_t)r * ld + c] = v;
    store_splits(sp, sbase + (int64_t)r * ld + c, v);
  }
}

// ---- R[r, kblock] = H[r, kblock] R_kk^-T for the rows above the diagonal block ---------------------------------------
// x R_kk^T = h  <=>  sum_{j >= i} x[j] R[i][j] = h[i]: back substitution from j = 127. Eight lanes own one row; lane l
// keeps columns l, l + 8, ... in registers; step j broadcasts h[j], divides, and every lane retires x[j] R[i][j] from its
// remaining columns i < j (RT[j][i] = R[i][j] in shared memory, padded rows).
constexpr int TR_LPR = 8, TR_NK = NB / TR_LPR;
__global__ void __launch_bounds__(256) chol_trsm_kernel(float* __restrict__ H, Splits sp, int64_t ld, int kb) {
  extern __shared__ float smf[];
  float* RT = smf;  // [NB][LDS_]
  const float* diag = H + (int64_t)kb * NB * ld + (int64_t)kb * NB;
  load_block<256>(diag, ld, [&](int i, int j, float v) { RT[j * LDS_ + i] = v; });
  float* rinv = smf + NB * LDS_;  // [NB] 1 / R[j][j]: the chain multiplies instead of dividing
  if (threadIdx.x < NB) rinv[threadIdx.x] = __fdiv_rn(1.f, diag[(int64_t)threadIdx.x * ld + threadIdx.x]);
  __syncthreads();
  const int lane = threadIdx.x & 31, l = lane & (TR_LPR - 1);
  const int64_t row = (int64_t)blockIdx.x * 32 + (threadIdx.x / TR_LPR);
  const int64_t off = row * ld + (int64_t)kb * NB;
  float reg[TR_NK], xv[TR_NK];
#pragma unroll
  for (int k = 0; k < TR_NK; ++k) {
    reg[k] = H[off + k * TR_LPR + l];
    xv[k] = 0.f;
  }
#pragma unroll
  for (int k = TR_NK - 1; k >= 0; --k) {
#pragma unroll 4
    for (int l0 = TR_LPR - 1; l0 >= 0; --l0) {
      const int j = k * TR_LPR + l0;
      const float hj = __shfl_sync(0xffffffffu, reg[k], (lane & ~(TR_LPR - 1)) | l0);
      const float x = __fmul_rn(hj, rinv[j]);
      if (l == l0) xv[k] = x;
      const float* rt = RT + j * LDS_ + l;
#pragma unroll
      for (int kk = 0; kk <= k; ++kk)
        if (kk < k || l < l0) reg[kk] = __fmaf_rn(-x, rt[kk * TR_LPR], reg[kk]);
    }
  }
#pragma unroll
  for (int k = 0; k < TR_NK; ++k) {
    H[off + k * TR_LPR + l] = xv[k];
    store_splits(sp, off + k * TR_LPR + l, xv[k]);
  }
}

// ---- U_ii = R_ii^-1 for every diagonal block (grid = nb CTAs of 128 threads; thread c solves column c) --------------
__global__ void __launch_bounds__(128) tri_inv_diag_kernel(const float* __restrict__ H, float* __restrict__ Dinv, int64_t ld) {
  extern __shared__ float smf[];
  float* R = smf;               // [NB][LDS_]
  float* U = smf + NB * LDS_;   // [NB][NB]  (U[k][c])
  const int kb = blockIdx.x;
  const float* base = H + (int64_t)kb * NB * ld + (int64_t)kb * NB;
  load_block<128>(base, ld, [&](int r, int c, float v) {
    R[r * LDS_ + c] = v;
    U[r * NB + c] = 0.f;
  });
  __syncthreads();
  const int c = threadIdx.x;
  const int cmax = (c | 31);  // warp-uniform upper bound of this warp's columns
  U[c * NB + c] = __fdiv_rn(1.f, R[c * LDS_ + c]);
  __syncwarp();
  for (int i = cmax - 1; i >= 0; --i) {
    float a0 = 0.f, a1 = 0.f;
    int k = i + 1;
    for (; k + 1 <= cmax; k += 2) {   // U[k][c] is zero for k > c, so over-running to cmax is harmless
      a0 = __fmaf_rn(R[i * LDS_ + k], U[k * NB + c], a0);
      a1 = __fmaf_rn(R[i * LDS_ + k + 1], U[(k + 1) * NB + c], a1);
    }
    if (k <= cmax) a0 = __fmaf_rn(R[i * LDS_ + k], U[k * NB + c], a0);
    if (i < c) U[i * NB + c] = __fdiv_rn(-(a0 + a1), R[i * LDS_ + i]);
  }
  __syncthreads();
  float* out = Dinv + (int64_t)kb * NB * NB;
  for (int idx = threadIdx.x; idx < NB * NB; idx += 128) out[idx] = U[idx];
}

// ---- U[i, cols] = -U_ii T[:, cols]; T is zeroed after it has been read; the last CTA writes U[i, i] = U_ii ----------
// CTA = 32 columns x 256 threads: lane = column, warp w = rows 16 w .. 16 w + 15. The T tile goes through shared memory
// (coalesced), every thread keeps its column of T in registers and reads U_ii rows as broadcast float4.
constexpr int RW_COLS = 32, RW_THREADS = 256, RW_ROWS = NB / (RW_THREADS / 32);
__global__ void __launch_bounds__(RW_THREADS) tri_row_kernel(float* __restrict__ H, Splits sp, float* __restrict__ T,
                                                             const float* __restrict__ Dinv, int64_t ld, int64_t n, int ib) {
  extern __shared__ float smf[];
  float* Uii = smf;             // [NB][NB]
  float* Ts = smf + NB * NB;    // [NB][RW_COLS]
  const float* d = Dinv + (int64_t)ib * NB * NB;
  {
    float4 v[NB * NB / 4 / RW_THREADS];
#pragma unroll
    for (int u = 0; u < NB * NB / 4 / RW_THREADS; ++u) v[u] = reinterpret_cast<const float4*>(d)[u * RW_THREADS + threadIdx.x];
#pragma unroll
    for (int u = 0; u < NB * NB / 4 / RW_THREADS; ++u) reinterpret_cast<float4*>(Uii)[u * RW_THREADS + threadIdx.x] = v[u];
  }
  const int64_t row0 = (int64_t)ib * NB;
  if (blockIdx.x == gridDim.x - 1) {   // the diagonal block itself
    __syncthreads();
    for (int idx = threadIdx.x; idx < NB * NB; idx += RW_THREADS) {
      const int r = idx >> 7, c = idx & 127;
      const int64_t g = (row0 + r) * ld + row0 + c;
      H[g] = Uii[idx];
      store_splits(sp, g, Uii[idx]);
    }
    return;
  }
  const int64_t col0 = row0 + NB + (int64_t)blockIdx.x * RW_COLS;
  {
    constexpr int PER = NB * RW_COLS / RW_THREADS;
    float v[PER];
#pragma unroll
    for (int u = 0; u < PER; ++u) {
      const int idx = u * RW_THREADS + threadIdx.x;
      v[u] = T[(int64_t)(idx >> 5) * ld + col0 + (idx & 31)];
    }
#pragma unroll
    for (int u = 0; u < PER; ++u) {
      const int idx = u * RW_THREADS + threadIdx.x;
      Ts[idx] = v[u];
      T[(int64_t)(idx >> 5) * ld + col0 + (idx & 31)] = 0.f;
    }
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float t[NB];
#pragma unroll
  for (int k = 0; k < NB; ++k) t[k] = Ts[k * RW_COLS + lane];
#pragma unroll 1
  for (int rr = 0; rr < RW_ROWS; ++rr) {
    const int r = warp * RW_ROWS + rr;
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    const float4* ur = reinterpret_cast<const float4*>(Uii + r * NB);
#pragma unroll
    for (int kc = 0; kc < NB / 16; ++kc) {
      if (kc * 16 + 15 >= r) {  // warp-uniform: U_ii[r][k] = 0 for k < r
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float4 u = ur[kc * 4 + q];
          a0 = __fmaf_rn(u.x, t[kc * 16 + q * 4 + 0], a0);
          a1 = __fmaf_rn(u.y, t[kc * 16 + q * 4 + 1], a1);
          a2 = __fmaf_rn(u.z, t[kc * 16 + q * 4 + 2], a2);
          a3 = __fmaf_rn(u.w, t[kc * 16 + q * 4 + 3], a3);
        }
      }
    }
    const float v = -((a0 + a1) + (a2 + a3));
    const int64_t g = (row0 + r) * ld + col0 + lane;
    H[g] = v;
    store_splits(sp, g, v);
  }
}

inline int64_t align_up(int64_t x, int64_t a) { return (x + a - 1) / a * a; }

struct Layout {
  int64_t np, off_split[3], off_dinv, off_T, off_pad, off_info, total;
};
Layout make_layout(int64_t n, int ns) {
  Layout L;
  L.np = align_up(n, NB);
  int64_t o = 0;
  for (int s = 0; s < 3; ++s) {
    L.off_split[s] = o;
    if (s < ns) o += align_up(L.np * L.np * 2, 256);
  }
  L.off_dinv = o;
  o += align_up((L.np / NB) * NB * NB * 4, 256);
  L.off_T = o;
  o += align_up((int64_t)NB * L.np * 4, 256);
  L.off_pad = o;
  if (L.np != n) o += align_up(L.np * L.np * 4, 256);
  L.off_info = o;
  o += 256;
  L.total = o + 256;
  return L;
}

}  // namespace
}  // namespace lcb

extern "C" int64_t lcb_cholinv_workspace_bytes(int64_t n, int32_t splits) {
  if (n < 1 || splits < 2 || splits > 3) return 0;
  return lcb::make_layout(n, splits).total;
}

extern "C" int lcb_cholinv_upper(float* H, int64_t n, int64_t ld, float percdamp, int32_t splits, void* workspace,
                                 int32_t* info, void* stream) {
  using namespace lcb;
  if (!H || !workspace || !info || n < 1 || ld < n) return LCB_E_INVALID;
  if (splits < 2 || splits > 3) return LCB_E_INVALID;
  if (n > (1 << 20)) return LCB_E_SHAPE;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const Layout L = make_layout(n, splits);
  uint8_t* ws = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~uintptr_t(255));
  const int64_t np = L.np;
  const int nb = (int)(np / NB);
  Splits sp;
  sp.ns = splits;
  for (int s = 0; s < 3; ++s) sp.s[s] = reinterpret_cast<__nv_bfloat16*>(ws + L.off_split[s < splits ? s : 0]);
  float* Dinv = reinterpret_cast<float*>(ws + L.off_dinv);
  float* T = reinterpret_cast<float*>(ws + L.off_T);

  constexpr int SM_DIAG = (NB * LDS_ + NB * XLD) * 4, SM_TRSM = (NB * LDS_ + NB) * 4, SM_INV = (NB * LDS_ + NB * NB) * 4, SM_ROW = (NB * NB + NB * RW_COLS) * 4;
  static DeviceOnce attr_once;
  if (attr_once.needed()) {
    if (cudaFuncSetAttribute(chol_diag_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_DIAG) != cudaSuccess ||
        cudaFuncSetAttribute(chol_trsm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_TRSM) != cudaSuccess ||
        cudaFuncSetAttribute(tri_inv_diag_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_INV) != cudaSuccess ||
        cudaFuncSetAttribute(tri_row_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_ROW) != cudaSuccess) {
      set_last_error("cudaFuncSetAttribute(cholinv)", cudaGetLastError());
      return LCB_E_CUDA;
    }
    attr_once.done();
  }
  if (cudaMemsetAsync(info, 0, sizeof(int32_t), st) != cudaSuccess) return LCB_E_CUDA;
  if (percdamp != 0.f) {
    chol_damp_kernel<<<1, 1024, 0, st>>>(H, n, ld, percdamp);
    LCB_CUDA_CHECK_LAUNCH("chol_damp");
  }
  float* A = H;       // the matrix the factorisation runs on
  int64_t lda = ld;
  if (np != n || (ld % 4) || (reinterpret_cast<uintptr_t>(H) & 15)) {
    if (np == n) return LCB_E_ALIGN;  // only the padded path owns a workspace copy
    A = reinterpret_cast<float*>(ws + L.off_pad);
    lda = np;
    chol_pad_in_kernel<<<kNumSMs * 8, 256, 0, st>>>(H, A, n, ld, np);
    LCB_CUDA_CHECK_LAUNCH("chol_pad_in");
  }
  // the bf16 pieces share A's indexing (leading dimension lda); rows below the diagonal must read as zero in phase 2
  if (lda != np) return LCB_E_ALIGN;  // pieces are [np, np]: an unpadded H must be contiguous
  for (int s = 0; s < splits; ++s)
    if (cudaMemsetAsync(sp.s[s], 0, (size_t)np * np * 2, st) != cudaSuccess) return LCB_E_CUDA;
  if (cudaMemsetAsync(T, 0, (size_t)NB * np * 4, st) != cudaSuccess) return LCB_E_CUDA;

  // products of the pieces: (0,0) (0,1) (1,0) [+ (1,1) (0,2) (2,0)]
  const int pa[6] = {0, 0, 1, 1, 0, 2}, pb[6] = {0, 1, 0, 1, 2, 0};
  const int n_seg = splits == 2 ? 3 : 6;

  // ---- phase 1: A = R R^T, R upper, from the last block column to the first ----
  // Two-level trailing update (same sums, a quarter of the fp32 read-modify-write traffic): inside a super-block of
  // SUPER panels a factored panel is retired only from the super-block's remaining columns (K = 128, narrow); when the
  // super-block is done, its SUPER * 128 columns are retired from everything to the left in one SYRK (K = 512).
  constexpr int SUPER = 4;
  for (int k = nb - 1; k >= 0; --k) {
    chol_diag_kernel<<<1, 512, SM_DIAG, st>>>(A, sp, lda, k, info);
    LCB_CUDA_CHECK_LAUNCH("chol_diag");
    if (k == 0) break;
    const int64_t n1 = (int64_t)k * NB;
    chol_trsm_kernel<<<(int)(n1 / 32), 256, SM_TRSM, st>>>(A, sp, lda, k);
    LCB_CUDA_CHECK_LAUNCH("chol_trsm");
    const void* aa[6];
    const void* bb[6];
    const int64_t s0 = (int64_t)(k / SUPER) * SUPER * NB;   // first column of this super-block
    if (n1 > s0) {   // near: H[0:n1, s0:n1] -= R[0:n1, k] R[s0:n1, k]^T
      for (int s = 0; s < n_seg; ++s) {
        aa[s] = sp.s[pa[s]] + n1;                // rows 0 .. n1-1, columns n1 .. n1+127
        bb[s] = sp.s[pb[s]] + s0 * lda + n1;     // rows s0 .. n1-1 of the same block column
      }
      const int rc = tc_gemm_launch_ex(true, true, false, true, false, 1, aa, bb, n_seg, A + s0, n1, n1 - s0, NB, lda, lda, lda, st);
      if (rc != LCB_OK) return rc;
    }
    if (n1 == s0) {   // far: the finished super-block [s0, s1) leaves everything to its left
      const int64_t s1 = std::min<int64_t>(s0 + (int64_t)SUPER * NB, np);
      for (int s = 0; s < n_seg; ++s) {
        aa[s] = sp.s[pa[s]] + s0;                // rows 0 .. s0-1, columns s0 .. s1-1
        bb[s] = sp.s[pb[s]] + s0;
      }
      const int rc = tc_gemm_launch_ex(true, true, true, true, false, 1, aa, bb, n_seg, A, s0, s0, s1 - s0, lda, lda, lda, st);
      if (rc != LCB_OK) return rc;
    }
  }
  // ---- phase 2: U = R^-1 by block rows from the bottom, in place ----
  tri_inv_diag_kernel<<<nb, 128, SM_INV, st>>>(A, Dinv, lda);
  LCB_CUDA_CHECK_LAUNCH("tri_inv_diag");
  for (int i = nb - 1; i >= 0; --i) {
    const int64_t c0 = (int64_t)(i + 1) * NB;
    const int64_t nt = np - c0;   // trailing columns
    if (nt > 0) {
      const void* aa[6];
      const void* bb[6];
      for (int s = 0; s < n_seg; ++s) {
        aa[s] = sp.s[pa[s]] + (int64_t)i * NB * lda + c0;   // R[i, c0:]
        bb[s] = sp.s[pb[s]] + c0 * lda + c0;                // U[c0:, c0:]
      }
      const int rc = tc_gemm_launch_ex(true, false, false, false, true, 1, aa, bb, n_seg, T + c0, NB, nt, nt, lda, lda, np, st);
      if (rc != LCB_OK) return rc;
    }
    tri_row_kernel<<<(int)(nt / RW_COLS) + 1, RW_THREADS, SM_ROW, st>>>(A, sp, T, Dinv, lda, np, i);
    LCB_CUDA_CHECK_LAUNCH("tri_row");
  }
  if (A != H) {
    chol_pad_out_kernel<<<kNumSMs * 8, 256, 0, st>>>(H, A, n, ld, np);
    LCB_CUDA_CHECK_LAUNCH("chol_pad_out");
  } else {
    dim3 grid((unsigned)std::min<int64_t>((n + 255) / 256, 64), (unsigned)n);
    chol_zero_lower_kernel<<<grid, 256, 0, st>>>(H, n, ld);
    LCB_CUDA_CHECK_LAUNCH("chol_zero_lower");
  }
  return LCB_OK;
}
