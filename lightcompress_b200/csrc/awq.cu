// awq.cu — AWQ scale search on the GPU (subsystem 3 of the north star), awq.py:179-248.
//
// Per candidate ratio n the reference does, for every linear of the subset:
//   W.mul_(scales)                      (awq.py:41-47, in the weight dtype)       -> scale_cols_kernel (mul)
//   W = fake_quant_weight_dynamic(W)    (awq.py:147-164)                          -> lcb_fakequant (fq_group kernel)
//   x_tmp = x / scales                  (base_bq.py:877-889)                      -> scale_cols_kernel (div, IEEE)
//   out = inspect_module(x_tmp); loss = (org_out - out).float().pow(2).mean()     (awq.py:110-145)
// When the inspect module is the linear itself (o_proj, down_proj, fc2, ...) the last line is one tcgen05 GEMM whose
// epilogue rounds the accumulator to the 16-bit output dtype, subtracts it from org_out in that dtype and reduces the
// squared error (tc_gemm.cuh, EPI_MSE): `out` is never written and org_out is read once per candidate.
// The scaled + fake-quantised weight (<= 117 MB for Llama-3-8B) stays L2/HBM resident between the two kernels; its
// round trip is < 1 % of the candidate's GEMM time at T = 65536 calibration tokens.
#include "tc_gemm.cuh"

namespace lcb {
namespace {

// out[r, c] = rd(in[r, c] * s[c])  (op 0)   |   rd(in[r, c] / s[c])  (op 1, IEEE division)
template <typename T>
__global__ void __launch_bounds__(256)
    scale_cols_kernel(const T* __restrict__ in, const T* __restrict__ s, T* __restrict__ out, int64_t rows, int64_t cols, int op) {
  constexpr int VEC = TI<T>::VEC;
  const int64_t vpr = cols / VEC;
  const int64_t total = rows * vpr;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < total; v += stride) {
    const int64_t c0 = (v % vpr) * VEC;
    const uint4 raw = ld_stream(in + v * VEC);
    const uint4 sraw = *reinterpret_cast<const uint4*>(s + c0);
    float f[VEC], g[VEC];
    TI<T>::unpack(raw, f);
    TI<T>::unpack(sraw, g);
#pragma unroll
    for (int i = 0; i < VEC; ++i) f[i] = op == 0 ? __fmul_rn(f[i], g[i]) : __fdiv_rn(f[i], g[i]);
    st_stream(out + v * VEC, TI<T>::pack(f));
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
    scale_cols_scalar_kernel(const T* __restrict__ in, const T* __restrict__ s, T* __restrict__ out, int64_t rows, int64_t cols,
                             int op) {
  const int64_t total = rows * cols;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += stride) {
    const float a = TI<T>::load1(in + e), b = TI<T>::load1(s + e % cols);
    TI<T>::store1(out + e, op == 0 ? __fmul_rn(a, b) : __fdiv_rn(a, b));
  }
}

template <typename T>
int scale_cols_typed(const void* in, const void* s, void* out, int64_t rows, int64_t cols, int op, cudaStream_t st) {
  constexpr int VEC = TI<T>::VEC;
  const bool vec = (cols % VEC == 0) && !((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out) |
                                           reinterpret_cast<uintptr_t>(s)) & 15);
  const int64_t work = vec ? rows * (cols / VEC) : rows * cols;
  const int grid = (int)std::min<int64_t>((work + 255) / 256, (int64_t)kNumSMs * 16);
  if (vec)
    scale_cols_kernel<T><<<grid, 256, 0, st>>>((const T*)in, (const T*)s, (T*)out, rows, cols, op);
  else
    scale_cols_scalar_kernel<T><<<grid, 256, 0, st>>>((const T*)in, (const T*)s, (T*)out, rows, cols, op);
  LCB_CUDA_CHECK_LAUNCH("scale_cols");
  return LCB_OK;
}

inline int64_t align_up(int64_t x, int64_t a) { return (x + a - 1) / a * a; }

}  // namespace
}  // namespace lcb

extern "C" int lcb_scale_cols(const void* in, const void* scales, void* out, int64_t rows, int64_t cols, int32_t dtype,
                              int32_t op, void* stream) {
  using namespace lcb;
  if (!in || !scales || !out || rows < 0 || cols < 0 || (op != 0 && op != 1)) return LCB_E_INVALID;
  if (rows == 0 || cols == 0) return LCB_OK;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (dtype == LCB_BF16) return scale_cols_typed<__nv_bfloat16>(in, scales, out, rows, cols, op, st);
  if (dtype == LCB_F16) return scale_cols_typed<__half>(in, scales, out, rows, cols, op, st);
  if (dtype == LCB_F32) return scale_cols_typed<float>(in, scales, out, rows, cols, op, st);
  return LCB_E_UNSUPPORTED;
}

extern "C" int lcb_linear_mse(const void* x, const void* w, const void* org_out, double* loss, int64_t T, int64_t R,
                              int64_t C, int32_t dtype, void* stream) {
  using namespace lcb;
  if (!x || !w || !org_out || !loss) return LCB_E_INVALID;
  if (dtype != LCB_BF16 && dtype != LCB_F16) return LCB_E_UNSUPPORTED;
  return tc_linear_mse_launch(dtype == LCB_F16 ? 0 : 1, x, w, org_out, loss, T, R, C, C, C, R,
                              reinterpret_cast<cudaStream_t>(stream));
}

extern "C" int lcb_linear(const void* x, const void* w, const void* bias, void* out, int64_t T, int64_t R, int64_t C,
                          int32_t dtype, void* stream) {
  using namespace lcb;
  if (!x || !w || !out) return LCB_E_INVALID;
  if (dtype != LCB_BF16 && dtype != LCB_F16) return LCB_E_UNSUPPORTED;
  return tc_linear_launch(dtype == LCB_F16 ? 0 : 1, x, w, bias, out, T, R, C, C, C, R, reinterpret_cast<cudaStream_t>(stream));
}

extern "C" int64_t lcb_awq_workspace_bytes(int64_t T, int64_t R, int64_t C, int32_t n_cand) {
  (void)n_cand;
  return 2 * lcb::align_up(R * C * 2, 256) + lcb::align_up(T * C * 2, 256) + lcb_fakequant_workspace_bytes() + 1024;
}

extern "C" int lcb_awq_candidates(const void* W, const void* x, const void* org_out, const void* cand_scales, double* loss,
                                  int64_t T, int64_t R, int64_t C, int32_t n_cand, int32_t group_size,
                                  const lcb_quant_spec* spec, void* workspace, void* stream) {
  using namespace lcb;
  if (!W || !x || !org_out || !cand_scales || !loss || !spec || !workspace || n_cand < 0) return LCB_E_INVALID;
  if (spec->dtype != LCB_BF16 && spec->dtype != LCB_F16) return LCB_E_UNSUPPORTED;
  if (spec->out_mode != LCB_OUT_FAKE) return LCB_E_INVALID;
  uint8_t* ws = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~uintptr_t(255));
  void* wsc = ws;                                  // W * s
  void* wq = ws + align_up(R * C * 2, 256);        // fake_quant(W * s)
  void* xs = ws + 2 * align_up(R * C * 2, 256);    // x / s
  void* fqws = ws + 2 * align_up(R * C * 2, 256) + align_up(T * C * 2, 256);
  const int64_t ucols = group_size > 0 ? group_size : C;
  for (int n = 0; n < n_cand; ++n) {
    const void* s = reinterpret_cast<const uint8_t*>(cand_scales) + (int64_t)n * C * 2;
    int rc = lcb_scale_cols(W, s, wsc, R, C, spec->dtype, 0, stream);
    if (rc != LCB_OK) return rc;
    rc = lcb_fakequant(wsc, wq, nullptr, nullptr, R, C, 1, ucols, spec, 1, fqws, stream);
    if (rc != LCB_OK) return rc;
    rc = lcb_scale_cols(x, s, xs, T, C, spec->dtype, 1, stream);
    if (rc != LCB_OK) return rc;
    rc = lcb_linear_mse(xs, wq, org_out, loss + n, T, R, C, spec->dtype, stream);
    if (rc != LCB_OK) return rc;
  }
  return LCB_OK;
}
