// autoclip.cu — AutoClip v1 search (auto_clip.py:84-191): for every (output channel, group) pick the shrink step
// i_s in [0, n_steps) whose clamped + fake-quantized weights reproduce the layer output best on the sampled tokens.
//
//   ac_init_kernel   : org_max / org_min per (row, group) (clip_sym: org_max = |w|.amax), best = org, min_err = 1e9
//   tc_gemm (grouped): org_out[g] = rd(W_g x_g^T)                          (EPI_STORE16, K = group_size per group)
//   per step         : lcb_fakequant with spec.clip_ratio = +/- (1 - i_s / n_grid)   (clamp fused into the quant kernel)
//                      tc_gemm (grouped, EPI_ROWSSE): sse[r, g] = sum_t rd(rd(rd(out) - org)^2)   (no output tensor)
//                      ac_update_kernel: err = rd(sse / T); strict < keeps the first minimum like the reference loop
// The reference materialises x * w as a [256, tok, C] tensor per step; here a step is one fake-quant pass over W and one
// tensor-core pass whose epilogue reduces over tokens. Products are exact and summed in fp32 (the reference rounds each
// product to the tensor dtype first), so near-tied steps can resolve differently: tests report the agreement.
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

template <typename T>
__global__ void __launch_bounds__(256) ac_update_kernel(const double* __restrict__ sse, const float* __restrict__ omax,
                                                        const float* __restrict__ omin, float* __restrict__ min_err,
                                                        T* __restrict__ best_max, T* __restrict__ best_min, int64_t n_units,
                                                        float ratio, float tokens, int clip_sym) {
  const int64_t u = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (u >= n_units) return;
  // .pow(2).mean(dim=1): fp32 accumulation, one division, rounded to the tensor dtype (auto_clip.py:150)
  const float err = TI<T>::rd(__fdiv_rn((float)sse[u], tokens));
  if (err < min_err[u]) {
    const float mxv = TI<T>::rd(__fmul_rn(omax[u], ratio));                         // org_max_val * (1 - i_s / n_grid)
    const float mnv = clip_sym ? -mxv : TI<T>::rd(__fmul_rn(omin[u], ratio));
    min_err[u] = err;
    TI<T>::store1(best_max + u, mxv);
    TI<T>::store1(best_min + u, mnv);
  }
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
  const int fmt = TI<T>::ID == LCB_F16 ? 0 : 1;
  ac_init_kernel<T><<<(int)((nu * 32 + 255) / 256), 256, 0, st>>>((const T*)W, (T*)best_max, (T*)best_min, omax, omin, min_err,
                                                                  nu, (int)gs, clip_sym);
  LCB_CUDA_CHECK_LAUNCH("ac_init");
  int rc = tc_grouped_rowsse_launch(fmt, 1, W, x, org, nullptr, R, T_rows, C, gs, st);
  if (rc != LCB_OK) return rc;
  lcb_quant_spec sp = *spec;
  for (int i_s = 0; i_s < n_steps; ++i_s) {
    const float ratio = (float)(1.0 - (double)i_s / (double)n_grid);
    sp.clip_ratio = clip_sym ? -ratio : ratio;
    rc = lcb_fakequant(W, Q, nullptr, nullptr, R, C, 1, gs, &sp, 1, fqws, st);
    if (rc != LCB_OK) return rc;
    if (cudaMemsetAsync(sse, 0, (size_t)nu * 8, st) != cudaSuccess) return LCB_E_CUDA;
    rc = tc_grouped_rowsse_launch(fmt, 0, Q, x, org, sse, R, T_rows, C, gs, st);
    if (rc != LCB_OK) return rc;
    ac_update_kernel<T><<<(int)((nu + 255) / 256), 256, 0, st>>>(sse, omax, omin, min_err, (T*)best_max, (T*)best_min, nu,
                                                                 ratio, (float)tokens, clip_sym);
    LCB_CUDA_CHECK_LAUNCH("ac_update");
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
  if (gs % 64 || T_rows % 8) return LCB_E_UNSUPPORTED;
  uint8_t* ws = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~uintptr_t(255));
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (spec->dtype == LCB_BF16)
    return autoclip_typed<__nv_bfloat16>(W, x, best_max, best_min, R, C, T_rows, tokens, gs, n_grid, n_steps, clip_sym, spec, ws, st);
  return autoclip_typed<__half>(W, x, best_max, best_min, R, C, T_rows, tokens, gs, n_grid, n_steps, clip_sym, spec, ws, st);
}
