// pack.cu — real-quant export packers: the reference packs on the host (numpy) or with a Python loop
// over columns (module_utils.py:931-955, 1099-1158); here each is one bandwidth-bound kernel.
#include <algorithm>
#include "common.cuh"

namespace lcb {
namespace {

// VllmRealQuantLinear.pack (module_utils.py:931-955): codes + 2^(b-1) as unsigned, `pf = 32 / b` consecutive columns per
// int32 word, column j * pf + i in bits [b i, b i + b); columns past the end pad with zero (before the offset is
// applied — numpy pads the already offset array with 0).
template <typename CT>
__global__ void __launch_bounds__(256) pack_rows_kernel(const CT* __restrict__ codes, int32_t* __restrict__ out, int64_t rows,
                                                        int64_t cols, int64_t packed_cols, int bit, int offset) {
  const int pf = 32 / bit;
  const int64_t total = rows * packed_cols;
  for (int64_t w = (int64_t)blockIdx.x * 256 + threadIdx.x; w < total; w += (int64_t)gridDim.x * 256) {
    const int64_t r = w / packed_cols, j = w % packed_cols;
    const CT* src = codes + r * cols + j * pf;
    uint32_t word = 0;
    for (int i = 0; i < pf; ++i) {
      if (j * pf + i < cols) {
        const uint32_t v = (uint32_t)(uint8_t)((int)src[i] + offset);   // (weight + offset).to(torch.uint8)
        word |= v << (bit * i);
      }
    }
    out[w] = (int32_t)word;
  }
}

// AutoawqRealQuantLinear.gemm_pack (module_utils.py:1099-1158), 4-bit:
//   intweight[r, c] = int(round((w[r, c] + zeros[g, r] * scales[g, r]) / scales[g, r])),  g = c / group
//   qweight[c, r / 8] |= intweight[r8 * 8 + order[i], c] << 4 i,   order = 0 2 4 6 1 3 5 7      (same packing for qzeros)
// Arithmetic dtype as torch promotes it: scales are fp16; an fp16 weight keeps everything in fp16 (one rounding per op),
// a bf16 / fp32 weight promotes the sum and the quotient to fp32. zeros * scales is an fp16 product either way.
// CTA = 32 input channels x 64 output channels; words go through shared memory so that every row of qweight is written
// as one 32-byte segment.
template <typename T>
__global__ void __launch_bounds__(256) awq_gemm_pack_kernel(const T* __restrict__ W, const __half* __restrict__ scales,
                                                            const int32_t* __restrict__ zeros, int32_t* __restrict__ qweight,
                                                            int64_t R, int64_t C, int group) {
  __shared__ int32_t tile[32][9];
  const int cl = threadIdx.x & 31, wl = threadIdx.x >> 5;
  const int64_t c = (int64_t)blockIdx.x * 32 + cl;
  const int64_t r8 = (int64_t)blockIdx.y * 8 + wl;
  const int64_t R8 = R / 8;
  const int order[8] = {0, 2, 4, 6, 1, 3, 5, 7};
  int32_t word = 0;
  if (c < C && r8 < R8) {
    const int64_t g = c / group;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int64_t r = r8 * 8 + order[i];
      const __half s = scales[g * R + r];
      const float sf = __half2float(s);
      const float sz = __half2float(__float2half_rn(__fmul_rn((float)zeros[g * R + r], sf)));   // int32 * fp16 -> fp16
      const float w = TI<T>::load1(W + r * C + c);
      float q;
      if (TI<T>::ID == LCB_F16) {
        const float t = __half2float(__float2half_rn(__fadd_rn(w, sz)));
        q = __half2float(__float2half_rn(__fdiv_rn(t, sf)));
      } else {
        q = __fdiv_rn(__fadd_rn(w, sz), sf);
      }
      word |= (int32_t)rintf(q) << (4 * i);
    }
  }
  tile[cl][wl] = word;
  __syncthreads();
  const int cl2 = threadIdx.x >> 3, w2 = threadIdx.x & 7;
  const int64_t c2 = (int64_t)blockIdx.x * 32 + cl2, r82 = (int64_t)blockIdx.y * 8 + w2;
  if (c2 < C && r82 < R8) qweight[c2 * R8 + r82] = tile[cl2][w2];
}

// qzeros[g, r / 8] from zeros[g, r] (int32), same nibble order
__global__ void __launch_bounds__(256) awq_pack_zeros_kernel(const int32_t* __restrict__ zeros, int32_t* __restrict__ qzeros,
                                                             int64_t G, int64_t R) {
  const int64_t R8 = R / 8;
  const int64_t w = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (w >= G * R8) return;
  const int64_t g = w / R8, r8 = w % R8;
  const int order[8] = {0, 2, 4, 6, 1, 3, 5, 7};
  int32_t word = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) word |= zeros[g * R + r8 * 8 + order[i]] << (4 * i);
  qzeros[w] = word;
}

}  // namespace
}  // namespace lcb

extern "C" int lcb_pack_rows(const void* codes, int32_t* out, int64_t rows, int64_t cols, int32_t code_bytes, int32_t bit,
                             void* stream) {
  using namespace lcb;
  if (!codes || !out || rows < 0 || cols < 0) return LCB_E_INVALID;
  if (bit < 1 || bit > 8 || (code_bytes != 1 && code_bytes != 4)) return LCB_E_UNSUPPORTED;
  if (rows == 0 || cols == 0) return LCB_OK;
  const int pf = 32 / bit;
  const int64_t packed_cols = (cols + pf - 1) / pf;
  const int offset = (1 << bit) / 2;
  const int64_t total = rows * packed_cols;
  const int grid = (int)std::min<int64_t>((total + 255) / 256, (int64_t)kNumSMs * 16);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (code_bytes == 1)
    pack_rows_kernel<int8_t><<<grid, 256, 0, st>>>((const int8_t*)codes, out, rows, cols, packed_cols, bit, offset);
  else
    pack_rows_kernel<int32_t><<<grid, 256, 0, st>>>((const int32_t*)codes, out, rows, cols, packed_cols, bit, offset);
  LCB_CUDA_CHECK_LAUNCH("pack_rows");
  return LCB_OK;
}

extern "C" int lcb_awq_gemm_pack(const void* W, const void* scales, const int32_t* zeros, int32_t* qweight, int32_t* qzeros,
                                 int64_t R, int64_t C, int32_t group_size, int32_t dtype, void* stream) {
  using namespace lcb;
  if (!W || !scales || !zeros || !qweight || !qzeros || R < 1 || C < 1 || group_size < 1) return LCB_E_INVALID;
  if (R % 32 || C % group_size) return LCB_E_SHAPE;   // the reference sizes qweight as R / 32 * bit words per row
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  dim3 grid((unsigned)((C + 31) / 32), (unsigned)((R / 8 + 7) / 8));
  const __half* sc = reinterpret_cast<const __half*>(scales);
  if (dtype == LCB_F16) awq_gemm_pack_kernel<__half><<<grid, 256, 0, st>>>((const __half*)W, sc, zeros, qweight, R, C, group_size);
  else if (dtype == LCB_BF16)
    awq_gemm_pack_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)W, sc, zeros, qweight, R, C, group_size);
  else if (dtype == LCB_F32) awq_gemm_pack_kernel<float><<<grid, 256, 0, st>>>((const float*)W, sc, zeros, qweight, R, C, group_size);
  else return LCB_E_UNSUPPORTED;
  LCB_CUDA_CHECK_LAUNCH("awq_gemm_pack");
  const int64_t G = C / group_size;
  awq_pack_zeros_kernel<<<(int)((G * (R / 8) + 255) / 256), 256, 0, st>>>(zeros, qzeros, G, R);
  LCB_CUDA_CHECK_LAUNCH("awq_pack_zeros");
  return LCB_OK;
}
