// tc_gemm.cu — host side of the tcgen05 contraction (tc_gemm.cuh): tensor-map encoding and launch.
#include "tc_gemm.cuh"

namespace lcb {

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

namespace {
template <bool AK, bool BK_, bool SY, bool NG, int EP>
int launch_variant(const tcg::Maps& maps, const tcg::Params& p, int grid, cudaStream_t st) {
  static DeviceOnce attr_once;
  if (attr_once.needed()) {
    if (cudaFuncSetAttribute(tcg::gemm_kernel<AK, BK_, SY, NG, EP>, cudaFuncAttributeMaxDynamicSharedMemorySize, tcg::SMEM_BYTES) !=
        cudaSuccess) {
      set_last_error("cudaFuncSetAttribute(tc_gemm)", cudaGetLastError());
      return LCB_E_CUDA;
    }
    attr_once.done();
  }
  tcg::gemm_kernel<AK, BK_, SY, NG, EP><<<grid, tcg::NTHREADS, tcg::SMEM_BYTES, st>>>(maps, p);
  LCB_CUDA_CHECK_LAUNCH("tc_gemm");
  return LCB_OK;
}
}  // namespace

int tc_gemm_launch(bool a_kmajor, bool syrk, bool neg, int fmt, const void* const* a, const void* const* b, int n_seg,
                   float* D, int64_t M, int64_t N, int64_t K, int64_t lda, int64_t ldb, int64_t ldd, cudaStream_t st) {
  return tc_gemm_launch_ex(a_kmajor, false, syrk, neg, false, fmt, a, b, n_seg, D, M, N, K, lda, ldb, ldd, st);
}

int tc_gemm_launch_ex(bool a_kmajor, bool b_kmajor, bool syrk, bool neg, bool tri_b, int fmt, const void* const* a,
                      const void* const* b, int n_seg, float* D, int64_t M, int64_t N, int64_t K, int64_t lda, int64_t ldb,
                      int64_t ldd, cudaStream_t st) {
  using namespace tcg;
  if (tri_b && b_kmajor) return LCB_E_INVALID;
  if (M <= 0 || N <= 0 || K <= 0) return LCB_OK;
  if (n_seg < 1 || n_seg > MAX_SEG) return LCB_E_INVALID;
  if (syrk && M != N) return LCB_E_INVALID;
  if (M > (1 << 30) || N > (1 << 30) || K > (1 << 30)) return LCB_E_SHAPE;
  // TMA: 16-byte aligned bases and row pitches
  if ((lda % 8) || (ldb % 8) || (ldd % 4) || (reinterpret_cast<uintptr_t>(D) & 15)) return LCB_E_ALIGN;
  for (int s = 0; s < n_seg; ++s)
    if ((reinterpret_cast<uintptr_t>(a[s]) & 15) || (reinterpret_cast<uintptr_t>(b[s]) & 15)) return LCB_E_ALIGN;

  Maps maps;
  Params p;
  p.M = (int)M;
  p.N = (int)N;
  p.K = (int)K;
  p.nblk_i = (int)((M + BM - 1) / BM);
  p.nblk_j = (int)((N + BN - 1) / BN);
  if (syrk) {
    p.n_tiles = 0;
    for (int i = 0; i < p.nblk_i; ++i) p.n_tiles += p.nblk_j - (i >> 1);
  } else {
    p.n_tiles = p.nblk_i * p.nblk_j;
  }
  const int kb_total = (int)((K + BK - 1) / BK);
  // split K until the grid is filled about twice, keeping >= 8 K blocks (512 deep) per split
  int nsplit = 1;
  while (p.n_tiles * nsplit < 2 * kNumSMs && kb_total / (nsplit * 2) >= 8) nsplit *= 2;
  p.kb_per_split = (kb_total + nsplit - 1) / nsplit;
  p.nsplit = (kb_total + p.kb_per_split - 1) / p.kb_per_split;
  p.n_seg = n_seg;
  p.n_grp = 1;
  p.kb_grp = 0;
  p.org_grp_stride = 0;
  p.rowsse = nullptr;
  p.tri_b = tri_b ? 1 : 0;
  p.idesc = tc::idesc_f16(fmt, BM, BN, a_kmajor ? 0 : 1, b_kmajor ? 0 : 1);
  p.org = nullptr;
  p.ld_org = 0;
  p.org_fmt = fmt;
  p.loss = nullptr;
  p.bias = nullptr;

  const CUtensorMapDataType dt16 = fmt == 0 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  bool ok = true;
  for (int s = 0; s < n_seg; ++s) {
    if (a_kmajor)
      ok = ok && make_tmap_2d(&maps.a[s], a[s], dt16, 2, M, K, lda * 2, BM, BK);   // [M, K]: box 128 rows x 64 k
    else
      ok = ok && make_tmap_2d(&maps.a[s], a[s], dt16, 2, K, M, lda * 2, BK, 64);   // [K, M]: box 64 k-rows x 64 m
    if (b_kmajor)
      ok = ok && make_tmap_2d(&maps.b[s], b[s], dt16, 2, N, K, ldb * 2, BN, BK);   // [N, K]: box 256 n-rows x 64 k
    else
      ok = ok && make_tmap_2d(&maps.b[s], b[s], dt16, 2, K, N, ldb * 2, BK, 64);   // [K, N]: box 64 k-rows x 64 n
  }
  for (int s = n_seg; s < MAX_SEG; ++s) {
    maps.a[s] = maps.a[0];
    maps.b[s] = maps.b[0];
  }
  ok = ok && make_tmap_2d(&maps.d, D, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, M, N, ldd * 4, 32, 32);
  if (!ok) {
    set_last_error("cuTensorMapEncodeTiled", cudaErrorInvalidValue);
    return LCB_E_CUDA;
  }
  const int n_items = p.n_tiles * p.nsplit;
  const int grid = n_items < kNumSMs ? n_items : kNumSMs;
  if (b_kmajor) {
    if (a_kmajor && syrk && neg) return launch_variant<true, true, true, true, EPI_REDUCE_ADD>(maps, p, grid, st);
    if (a_kmajor && !syrk && neg) return launch_variant<true, true, false, true, EPI_REDUCE_ADD>(maps, p, grid, st);
    return LCB_E_UNSUPPORTED;
  }
  if (!a_kmajor && syrk && !neg) return launch_variant<false, false, true, false, EPI_REDUCE_ADD>(maps, p, grid, st);
  if (a_kmajor && !syrk && neg) return launch_variant<true, false, false, true, EPI_REDUCE_ADD>(maps, p, grid, st);
  if (a_kmajor && !syrk && !neg) return launch_variant<true, false, false, false, EPI_REDUCE_ADD>(maps, p, grid, st);
  return LCB_E_UNSUPPORTED;
}

int tc_linear_mse_launch(int fmt, const void* x, const void* w, const void* org, double* loss, int64_t M, int64_t N,
                         int64_t K, int64_t ldx, int64_t ldw, int64_t ldo, cudaStream_t st) {
  using namespace tcg;
  if (M <= 0 || N <= 0) return LCB_OK;
  if (K <= 0) return LCB_E_INVALID;
  if (M > (1 << 30) || N > (1 << 30) || K > (1 << 30)) return LCB_E_SHAPE;
  if ((ldx % 8) || (ldw % 8) || (ldo % 8) || (N % 8) || (reinterpret_cast<uintptr_t>(x) & 15) ||
      (reinterpret_cast<uintptr_t>(w) & 15) || (reinterpret_cast<uintptr_t>(org) & 15))
    return LCB_E_ALIGN;
  Maps maps;
  Params p;
  p.M = (int)M;
  p.N = (int)N;
  p.K = (int)K;
  p.nblk_i = (int)((M + BM - 1) / BM);
  p.nblk_j = (int)((N + BN - 1) / BN);
  p.n_tiles = p.nblk_i * p.nblk_j;
  p.nsplit = 1;  // the output is rounded before it is compared: no split-K
  p.kb_per_split = (int)((K + BK - 1) / BK);
  p.n_seg = 1;
  p.n_grp = 1;
  p.kb_grp = 0;
  p.org_grp_stride = 0;
  p.rowsse = nullptr;
  p.tri_b = 0;
  p.idesc = tc::idesc_f16(fmt, BM, BN, 0, 0);
  p.org = org;
  p.ld_org = ldo;
  p.org_fmt = fmt;
  p.loss = loss;
  p.bias = nullptr;
  const CUtensorMapDataType dt16 = fmt == 0 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  bool ok = true;
  for (int s = 0; s < MAX_SEG; ++s) {
    ok = ok && make_tmap_2d(&maps.a[s], x, dt16, 2, M, K, ldx * 2, BM, BK);   // [M, K]: box 128 rows x 64 k
    ok = ok && make_tmap_2d(&maps.b[s], w, dt16, 2, N, K, ldw * 2, BN, BK);   // [N, K]: box 256 rows x 64 k
  }
  maps.d = maps.a[0];  // unused by the MSE epilogue
  if (!ok) {
    set_last_error("cuTensorMapEncodeTiled", cudaErrorInvalidValue);
    return LCB_E_CUDA;
  }
  const int grid = p.n_tiles < kNumSMs ? p.n_tiles : kNumSMs;
  return launch_variant<true, true, false, false, EPI_MSE>(maps, p, grid, st);
}

int tc_linear_launch(int fmt, const void* x, const void* w, const void* bias, void* out, int64_t M, int64_t N, int64_t K,
                     int64_t ldx, int64_t ldw, int64_t ldo, cudaStream_t st) {
  using namespace tcg;
  if (M <= 0 || N <= 0) return LCB_OK;
  if (K <= 0) return LCB_E_INVALID;
  if (M > (1 << 30) || N > (1 << 30) || K > (1 << 30)) return LCB_E_SHAPE;
  if ((ldx % 8) || (ldw % 8) || (ldo % 8) || (N % 8) || (reinterpret_cast<uintptr_t>(x) & 15) ||
      (reinterpret_cast<uintptr_t>(w) & 15) || (reinterpret_cast<uintptr_t>(out) & 15) ||
      (bias && (reinterpret_cast<uintptr_t>(bias) & 3)))
    return LCB_E_ALIGN;
  Maps maps;
  Params p;
  p.M = (int)M;
  p.N = (int)N;
  p.K = (int)K;
  p.nblk_i = (int)((M + BM - 1) / BM);
  p.nblk_j = (int)((N + BN - 1) / BN);
  p.n_tiles = p.nblk_i * p.nblk_j;
  p.nsplit = 1;  // the output is rounded in the epilogue: no split-K
  p.kb_per_split = (int)((K + BK - 1) / BK);
  p.n_seg = 1;
  p.n_grp = 1;
  p.kb_grp = 0;
  p.org_grp_stride = 0;
  p.rowsse = nullptr;
  p.tri_b = 0;
  p.idesc = tc::idesc_f16(fmt, BM, BN, 0, 0);
  p.org = out;
  p.ld_org = ldo;
  p.org_fmt = fmt;
  p.loss = nullptr;
  p.bias = bias;
  const CUtensorMapDataType dt16 = fmt == 0 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  bool ok = make_tmap_2d(&maps.a[0], x, dt16, 2, M, K, ldx * 2, BM, BK) &&   // [M, K]: box 128 rows x 64 k
            make_tmap_2d(&maps.b[0], w, dt16, 2, N, K, ldw * 2, BN, BK);     // [N, K]: box 256 rows x 64 k
  if (!ok) {
    set_last_error("cuTensorMapEncodeTiled", cudaErrorInvalidValue);
    return LCB_E_CUDA;
  }
  for (int s = 1; s < MAX_SEG; ++s) {
    maps.a[s] = maps.a[0];
    maps.b[s] = maps.b[0];
  }
  maps.d = maps.a[0];  // unused by this epilogue
  const int grid = p.n_tiles < kNumSMs ? p.n_tiles : kNumSMs;
  return launch_variant<true, true, false, false, EPI_STORE16>(maps, p, grid, st);
}

int tc_grouped_rowsse_launch(int fmt, int store, const void* w, const void* x, void* org, double* rowsse, int64_t R,
                             int64_t T, int64_t C, int64_t gs, cudaStream_t st) {
  using namespace tcg;
  if (R <= 0 || T <= 0) return LCB_OK;
  if (C <= 0 || gs <= 0 || C % gs) return LCB_E_SHAPE;
  if (gs % BK) return LCB_E_UNSUPPORTED;
  if (R > (1 << 30) || T > (1 << 30) || C > (1 << 30)) return LCB_E_SHAPE;
  if ((C % 8) || (T % 8) || (reinterpret_cast<uintptr_t>(x) & 15) || (reinterpret_cast<uintptr_t>(w) & 15) ||
      (reinterpret_cast<uintptr_t>(org) & 15))
    return LCB_E_ALIGN;
  if (!store && !rowsse) return LCB_E_INVALID;
  Maps maps;
  Params p;
  p.M = (int)R;
  p.N = (int)T;
  p.K = (int)C;
  p.nblk_i = (int)((R + BM - 1) / BM);
  p.nblk_j = (int)((T + BN - 1) / BN);
  p.n_tiles = p.nblk_i * p.nblk_j;
  p.nsplit = 1;
  p.n_grp = (int)(C / gs);
  p.kb_grp = (int)(gs / BK);
  p.kb_per_split = p.kb_grp;
  p.n_seg = 1;
  p.tri_b = 0;
  p.idesc = tc::idesc_f16(fmt, BM, BN, 0, 0);
  p.org = org;
  p.ld_org = T;
  p.org_grp_stride = R * T;
  p.org_fmt = fmt;
  p.loss = nullptr;
  p.bias = nullptr;
  p.rowsse = rowsse;
  const CUtensorMapDataType dt16 = fmt == 0 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  bool ok = make_tmap_2d(&maps.a[0], w, dt16, 2, R, C, C * 2, BM, BK) &&   // [R, C]: box 128 rows x 64 k
            make_tmap_2d(&maps.b[0], x, dt16, 2, T, C, C * 2, BN, BK);     // [T, C]: box 256 rows x 64 k
  if (!ok) {
    set_last_error("cuTensorMapEncodeTiled", cudaErrorInvalidValue);
    return LCB_E_CUDA;
  }
  for (int s = 1; s < MAX_SEG; ++s) {
    maps.a[s] = maps.a[0];
    maps.b[s] = maps.b[0];
  }
  maps.d = maps.a[0];  // unused by these epilogues
  const int64_t n_items = (int64_t)p.n_tiles * p.n_grp;
  if (n_items > (1ll << 30)) return LCB_E_SHAPE;
  const int grid = (int)(n_items < kNumSMs ? n_items : kNumSMs);
  if (store) return launch_variant<true, true, false, false, EPI_STORE16>(maps, p, grid, st);
  return launch_variant<true, true, false, false, EPI_ROWSSE>(maps, p, grid, st);
}

}  // namespace lcb
