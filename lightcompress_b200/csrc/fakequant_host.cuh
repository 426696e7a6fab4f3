// fakequant_host.cuh — host-side dispatch of lcb_fakequant (included by fakequant.cu)
#pragma once
#include <math.h>
#include <stdlib.h>

namespace lcb {

static bool encode_block_tmap(CUtensorMap* tm, const void* in, int dtype, int64_t rows, int64_t cols);
static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

static SpecD make_specd(const lcb_quant_spec* s) {
  SpecD d;
  d.dtype = s->dtype;
  d.math_dtype = s->math_dtype;
  d.qparam_dtype = s->qparam_dtype;
  d.kind = s->kind;
  d.sym = s->sym;
  d.round_zp = s->round_zp;
  d.out_mode = s->out_mode;
  d.qmin = s->qmin;
  d.qmax = s->qmax;
  d.osf = s->output_scale_factor;
  d.clip = s->clip_ratio;
  d.qp_math = s->qp_math_plus1 > 0 ? s->qp_math_plus1 - 1 : s->math_dtype;
  bool ints = (floorf(d.qmin) == d.qmin) && (floorf(d.qmax) == d.qmax) && fabsf(d.qmin) <= 256.f &&
              fabsf(d.qmax) <= 256.f;
  d.fast_ok = (d.kind == LCB_KIND_INT) ? (d.round_zp && ints) : 1;
  return d;
}

// <<<grid, block, 0, st>>> with the programmatic-stream-serialization attribute (see pdl_enter in fq_group.cuh)
template <typename... KArgs, typename... Args>
static void launch_pdl(void (*kernel)(KArgs...), int grid, int block, cudaStream_t st, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3((unsigned)block);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

static int grid_for(int64_t work_ctas, int ctas_per_sm) {
  int64_t cap = (int64_t)kNumSMs * ctas_per_sm;
  return (int)(work_ctas < cap ? (work_ctas < 1 ? 1 : work_ctas) : cap);
}

static inline float host_rd(float x, int dt) {
  if (dt == LCB_BF16) return __bfloat162float(__float2bfloat16_rn(x));
  if (dt == LCB_F16) return __half2float(__float2half_rn(x));
  return x;
}

// constants + chain selection of the group kernel (fq_group.cuh / fq_packed.cuh)
template <typename T>
static FastSpec make_fastspec(const SpecD& sp) {
  FastSpec fs;
  fs.qmin = sp.qmin;
  fs.qmax = sp.qmax;
  fs.div = (sp.sym || sp.kind != LCB_KIND_INT) ? sp.qmax : sp.qmax - sp.qmin;
  fs.rdiv = 1.0f / fs.div;
  fs.eps = host_rd(1e-5f, TI<T>::ID);
  const float span = sp.qmax - sp.qmin;
  fs.chain = CHAIN_F32;
  if (sp.kind == LCB_KIND_INT && sp.fast_ok) {
    if (TI<T>::ID == LCB_BF16) fs.chain = span <= 63.f ? CHAIN_BF16_SMALL : span <= 255.f ? CHAIN_BF16_WIDE : CHAIN_F32;
    if (TI<T>::ID == LCB_F16 && span <= 511.f) fs.chain = CHAIN_F16;
  }
  // bf16 qparams by FMUL: div must be an integer whose odd part is small (fq_packed.cuh)
  fs.cheap_qp = 0;
  if (TI<T>::ID == LCB_BF16 && sp.fast_ok && fs.div >= 1.f && fs.div <= 1e6f && floorf(fs.div) == fs.div) {
    long d = (long)fs.div;
    while ((d & 1) == 0) d >>= 1;
    fs.cheap_qp = d <= 1024 ? 1 : 0;
  }
  return fs;
}

template <typename T, int KINDC>
static int launch_dynamic_1d(const T* in, void* out, void* scales, void* zeros, int64_t n_units, int64_t unit_len,
                             const SpecD& sp, cudaStream_t st) {
  constexpr int VEC = TI<T>::VEC;
  const int64_t nvec = unit_len / VEC;
  static const int64_t group_max_nvec = [] {   // tuning knob (tools/fq_sweep.py): largest unit served by the sub-warp kernel
    const char* e = getenv("LCB_GROUP_MAX_NVEC");
    return e ? (int64_t)atoll(e) : (int64_t)256;
  }();
  if (nvec <= group_max_nvec) {
    // sub-warp per unit: lanes = pow2 >= nvec / 2 in [4, 32], NV vectors per lane
    int lanes = 4;
    while (lanes * 2 < nvec && lanes < 32) lanes <<= 1;
    const int nv = (int)((nvec + lanes - 1) / lanes);
    const int nvt = nv <= 1 ? 1 : nv <= 2 ? 2 : nv <= 4 ? 4 : 8;
    const int u = nvt >= 8 ? 1 : 8 / nvt;
    const FastSpec fs = make_fastspec<T>(sp);
    const int64_t units_per_cta = (int64_t)8 * u * (32 / lanes);
    // grid = 2 x the resident CTAs (launch bound: 64 registers x 256 threads = 4 CTAs / SM): a grid-stride loop over whole
    // waves has no partial last wave (148 x 16 CTAs on 3 resident = 5.33 waves cost 2.4 % on the Llama-3-8B block)
    static const int group_cps = [] { const char* e = getenv("LCB_GROUP_CTAS_PER_SM"); return e ? atoi(e) : 8; }();
    const int grid = grid_for((n_units + units_per_cta - 1) / units_per_cta, group_cps);
    // chain compiled into the kernel: the packed 16-bit chains and the inline fp32 chain serve LCB_OUT_FAKE only
    int chain = CHAIN_ANY_OUT;
    if (sp.out_mode == LCB_OUT_FAKE) chain = (KINDC == 0) ? fs.chain : CHAIN_F32;
#define LCB_LAUNCH_GROUP_C(NVc, CH)                                                                             \
  launch_pdl(fq_group_kernel<T, NVc, KINDC, CH>, grid, 256, st, in, out, scales, zeros, n_units, (int)unit_len, lanes, sp, fs)
#define LCB_LAUNCH_GROUP(NVc)                                                                                   \
  {                                                                                                             \
    if (chain == CHAIN_ANY_OUT) LCB_LAUNCH_GROUP_C(NVc, CHAIN_ANY_OUT);                                         \
    else if (chain == CHAIN_F32) LCB_LAUNCH_GROUP_C(NVc, CHAIN_F32);                                            \
    else if (TI<T>::ID == LCB_BF16 && KINDC == 0 && chain == CHAIN_BF16_SMALL) {                                \
      if constexpr (TI<T>::ID == LCB_BF16 && KINDC == 0) LCB_LAUNCH_GROUP_C(NVc, CHAIN_BF16_SMALL);             \
    } else if (TI<T>::ID == LCB_BF16 && KINDC == 0 && chain == CHAIN_BF16_WIDE) {                               \
      if constexpr (TI<T>::ID == LCB_BF16 && KINDC == 0) LCB_LAUNCH_GROUP_C(NVc, CHAIN_BF16_WIDE);              \
    } else if (TI<T>::ID == LCB_F16 && KINDC == 0 && chain == CHAIN_F16) {                                      \
      if constexpr (TI<T>::ID == LCB_F16 && KINDC == 0) LCB_LAUNCH_GROUP_C(NVc, CHAIN_F16);                     \
    } else LCB_LAUNCH_GROUP_C(NVc, CHAIN_ANY_OUT);                                                              \
  }
    switch (nvt) {
      case 1: LCB_LAUNCH_GROUP(1) break;
      case 2: LCB_LAUNCH_GROUP(2) break;
      case 4: LCB_LAUNCH_GROUP(4) break;
      default: LCB_LAUNCH_GROUP(8) break;
    }
#undef LCB_LAUNCH_GROUP_C
#undef LCB_LAUNCH_GROUP
  } else if (nvec <= 2048) {
    // W warps per row, 8 vectors per lane (fq_wrow_kernel)
    const int w = nvec <= 512 ? 2 : nvec <= 1024 ? 4 : 8;
    const FastSpec fs = make_fastspec<T>(sp);
    int chain = CHAIN_ANY_OUT;
    if (sp.out_mode == LCB_OUT_FAKE) chain = (KINDC == 0) ? fs.chain : CHAIN_F32;
    const int grid = grid_for((n_units + (8 / w) - 1) / (8 / w), w == 8 ? 9 : 8);   // whole waves: 3 (W = 8) / 2 CTAs per SM
#define LCB_LAUNCH_WROW_C(CH, Wc) \
  launch_pdl(fq_wrow_kernel<T, KINDC, CH, Wc>, grid, 256, st, in, out, scales, zeros, n_units, (int)unit_len, sp, fs)
  // 8 warps per row: 6 or 7 vectors per lane when that covers the row (14336 16-bit elements = 7 x 256 vectors: no idle
  // load slots; 8 per lane would leave 12.5 % of them empty)
  const int nv8 = (int)((nvec + 255) / 256);
#define LCB_LAUNCH_WROW(CH)                                                                                                     \
  {                                                                                                                             \
    if (w == 2) LCB_LAUNCH_WROW_C(CH, 2);                                                                                       \
    else if (w == 4) LCB_LAUNCH_WROW_C(CH, 4);                                                                                  \
    else if (nv8 == 7)                                                                                                          \
      launch_pdl(fq_wrow_kernel<T, KINDC, CH, 8, 7>, grid, 256, st, in, out, scales, zeros, n_units, (int)unit_len, sp, fs);   \
    else if (nv8 == 6)                                                                                                          \
      launch_pdl(fq_wrow_kernel<T, KINDC, CH, 8, 6>, grid, 256, st, in, out, scales, zeros, n_units, (int)unit_len, sp, fs);   \
    else LCB_LAUNCH_WROW_C(CH, 8);                                                                                              \
  }
    if (chain == CHAIN_F32) LCB_LAUNCH_WROW(CHAIN_F32)
    else if (TI<T>::ID == LCB_BF16 && KINDC == 0 && chain == CHAIN_BF16_SMALL) {
      if constexpr (TI<T>::ID == LCB_BF16 && KINDC == 0) LCB_LAUNCH_WROW(CHAIN_BF16_SMALL)
    } else if (TI<T>::ID == LCB_BF16 && KINDC == 0 && chain == CHAIN_BF16_WIDE) {
      if constexpr (TI<T>::ID == LCB_BF16 && KINDC == 0) LCB_LAUNCH_WROW(CHAIN_BF16_WIDE)
    } else if (TI<T>::ID == LCB_F16 && KINDC == 0 && chain == CHAIN_F16) {
      if constexpr (TI<T>::ID == LCB_F16 && KINDC == 0) LCB_LAUNCH_WROW(CHAIN_F16)
    } else LCB_LAUNCH_WROW(CHAIN_ANY_OUT)
#undef LCB_LAUNCH_WROW
#undef LCB_LAUNCH_WROW_C
  } else {
    int grid = grid_for(n_units, 4);
    fq_row_stream_kernel<T, KINDC><<<grid, 512, 0, st>>>(in, out, scales, zeros, n_units, unit_len, sp);
  }
  LCB_CUDA_CHECK_LAUNCH("fq_dynamic_1d");
  return LCB_OK;
}

template <typename T, int KINDC>
static int launch_static(const T* in, void* out, const void* scales, const void* zeros, int64_t rows, int64_t cols,
                         int64_t ur, int64_t uc, const SpecD& sp, cudaStream_t st) {
  constexpr int VEC = TI<T>::VEC;
  int64_t total = rows * (cols / VEC);
  int grid = grid_for((total + 1023) / 1024, 8);
  if (sp.math_dtype == LCB_F32)
    fq_static_kernel<T, float, KINDC><<<grid, 256, 0, st>>>(in, out, scales, zeros, rows, cols, ur, uc, sp);
  else
    fq_static_kernel<T, T, KINDC><<<grid, 256, 0, st>>>(in, out, scales, zeros, rows, cols, ur, uc, sp);
  LCB_CUDA_CHECK_LAUNCH("fq_static");
  return LCB_OK;
}

constexpr int kMaxParts = kNumSMs * 8;

template <typename T>
static int fakequant_typed(const void* in_, void* out, void* scales, void* zeros, int64_t rows, int64_t cols,
                           int64_t ur, int64_t uc, SpecD sp, int dynamic, void* workspace, cudaStream_t st) {
  constexpr int VEC = TI<T>::VEC;
  const T* in = reinterpret_cast<const T*>(in_);
  const int kindc = sp.kind == LCB_KIND_INT ? 0 : 1;
  if (ur > rows) ur = rows;
  if (uc > cols) uc = cols;
  const bool vec_ok = aligned16(in) && (out == nullptr || aligned16(out)) && (cols % VEC == 0) && (uc % VEC == 0);
  const bool whole = (ur == rows && uc == cols);
  const bool math_is_T = sp.math_dtype == sp.dtype;
#define LCB_KIND_SWITCH(...)                                  \
  if (kindc == 0) { constexpr int KINDC = 0; __VA_ARGS__; }   \
  else { constexpr int KINDC = 1; __VA_ARGS__; }

  if (!dynamic) {
    if (!scales) return LCB_E_INVALID;
    if (vec_ok && (math_is_T || sp.math_dtype == LCB_F32) && sp.out_mode != LCB_OUT_DEQUANT) {
      LCB_KIND_SWITCH(return launch_static<T, KINDC>(in, out, scales, zeros, rows, cols, ur, uc, sp, st));
    }
    int64_t units = ((rows + ur - 1) / ur) * ((cols + uc - 1) / uc);
    fq_generic_kernel<T><<<grid_for((units + 7) / 8, 8), 256, 0, st>>>(in, out, scales, zeros, rows, cols, ur, uc, 0, 0, sp);
    LCB_CUDA_CHECK_LAUNCH("fq_generic_static");
    return LCB_OK;
  }
  // dynamic
  if (vec_ok && math_is_T && !whole && (ur == 1 || rows == 1) && cols % uc == 0) {
    LCB_KIND_SWITCH(return launch_dynamic_1d<T, KINDC>(in, out, scales, zeros, rows * (cols / uc), uc, sp, st));
  }
  if (vec_ok && math_is_T && whole && rows * cols >= 4096) {
    if (!workspace) return LCB_E_INVALID;
    float* ws = reinterpret_cast<float*>(workspace);
    int64_t numel = rows * cols;
    int parts = grid_for((numel / VEC + 1023) / 1024, 8);
    minmax_partial_kernel<T><<<parts, 256, 0, st>>>(in, numel, ws);
    minmax_finalize_kernel<<<1, 256, 0, st>>>(ws, parts, scales, zeros, sp);
    LCB_CUDA_CHECK_LAUNCH("fq_minmax");
    if (sp.out_mode == LCB_OUT_NONE) return LCB_OK;
    SpecD sp2 = sp;
    sp2.qparam_dtype = LCB_F32;
    float* q = ws + 2 * parts;
    LCB_KIND_SWITCH(return launch_static<T, KINDC>(in, out, q, sp.sym ? nullptr : q + 1, rows, cols, rows, cols, sp2, st));
  }
  if (vec_ok && sp.math_dtype == LCB_F32 && ur == 128 && uc == 128 && sp.sym) {
    int64_t tiles = ((rows + 127) / 128) * ((cols + 127) / 128);
    if (tiles > 0x7fffffff) return LCB_E_SHAPE;
    if (kindc == 1 && sp.osf == 1.0f && (sp.out_mode == LCB_OUT_FAKE || sp.out_mode == LCB_OUT_FP8)) {
      const bool e5 = sp.kind == LCB_KIND_FP8_E5M2, codes = sp.out_mode == LCB_OUT_FP8;
      if (sizeof(T) == 2 && cols % 8 == 0 && rows < (1ll << 31) && cols < (1ll << 31)) {
        // TMA-staged tiles (16-byte row pitch): no swizzle, box 128 rows x 128 elements, out-of-bounds = zeros
        CUtensorMap tm;
        lcb_encode_tiled_fn enc = get_encode_tiled();
        const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
        const cuuint64_t strides[1] = {(cuuint64_t)cols * 2};
        const cuuint32_t box[2] = {128, 128};
        const cuuint32_t estr[2] = {1, 1};
        if (enc && enc(&tm, TI<T>::ID == LCB_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2,
                       const_cast<T*>(in), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                       CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS) {
          using T16 = typename std::conditional<sizeof(T) == 2, T, __nv_bfloat16>::type;
          void* tm_out = out;
          if (!e5 && !codes) fq_block2d_fp8_tma_kernel<T16, 0, 0><<<(int)tiles, 256, 0, st>>>(tm, tm_out, scales, rows, cols, sp);
          else if (!e5 && codes) fq_block2d_fp8_tma_kernel<T16, 0, 1><<<(int)tiles, 256, 0, st>>>(tm, tm_out, scales, rows, cols, sp);
          else if (e5 && !codes) fq_block2d_fp8_tma_kernel<T16, 1, 0><<<(int)tiles, 256, 0, st>>>(tm, tm_out, scales, rows, cols, sp);
          else fq_block2d_fp8_tma_kernel<T16, 1, 1><<<(int)tiles, 256, 0, st>>>(tm, tm_out, scales, rows, cols, sp);
          LCB_CUDA_CHECK_LAUNCH("fq_block2d_fp8_tma");
          return LCB_OK;
        }
      }
      if (!e5 && !codes) fq_block2d_fp8_kernel<T, 0, 0><<<(int)tiles, 256, 0, st>>>(in, out, scales, rows, cols, sp);
      else if (!e5 && codes) fq_block2d_fp8_kernel<T, 0, 1><<<(int)tiles, 256, 0, st>>>(in, out, scales, rows, cols, sp);
      else if (e5 && !codes) fq_block2d_fp8_kernel<T, 1, 0><<<(int)tiles, 256, 0, st>>>(in, out, scales, rows, cols, sp);
      else fq_block2d_fp8_kernel<T, 1, 1><<<(int)tiles, 256, 0, st>>>(in, out, scales, rows, cols, sp);
      LCB_CUDA_CHECK_LAUNCH("fq_block2d_fp8");
      return LCB_OK;
    }
    if (sizeof(T) == 2 && cols % 8 == 0 && rows < (1ll << 31) && cols < (1ll << 31) && aligned16(in)) {
      CUtensorMap tm;
      if (encode_block_tmap(&tm, in, TI<T>::ID, rows, cols)) {
        using T16 = typename std::conditional<sizeof(T) == 2, T, __nv_bfloat16>::type;
        LCB_KIND_SWITCH(fq_block2d_tma_kernel<T16, KINDC><<<(int)tiles, 256, 0, st>>>(tm, out, scales, rows, cols, sp));
        LCB_CUDA_CHECK_LAUNCH("fq_block2d_tma");
        return LCB_OK;
      }
    }
    LCB_KIND_SWITCH(fq_block2d_kernel<T, KINDC><<<(int)tiles, 256, 0, st>>>(in, out, scales, rows, cols, sp));
    LCB_CUDA_CHECK_LAUNCH("fq_block2d");
    return LCB_OK;
  }
  {
    int64_t units = ((rows + ur - 1) / ur) * ((cols + uc - 1) / uc);
    int abs_range = (ur > 1 && !whole) ? 1 : 0;  // per_block: range of |x| (quant.py:137-138)
    fq_generic_kernel<T><<<grid_for((units + 7) / 8, 8), 256, 0, st>>>(in, out, scales, zeros, rows, cols, ur, uc, 1, abs_range, sp);
    LCB_CUDA_CHECK_LAUNCH("fq_generic_dynamic");
    return LCB_OK;
  }
#undef LCB_KIND_SWITCH
}

}  // namespace lcb

namespace lcb {
// ---- several 16-bit weights, one launch: FP8 per_block (128 x 128) dynamic fake-quant / real-quant ------------------------
static bool encode_block_tmap(CUtensorMap* tm, const void* in, int dtype, int64_t rows, int64_t cols) {
  lcb_encode_tiled_fn enc = get_encode_tiled();
  if (!enc) return false;
  const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)cols * 2};
  const cuuint32_t box[2] = {128, 128};
  const cuuint32_t estr[2] = {1, 1};
  return enc(tm, dtype == LCB_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2,
             const_cast<void*>(in), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
             CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

template <typename T>
static int fakequant_block_fp8_multi_typed(const void* const* ins, void* const* outs, void* const* scales, const int64_t* rows,
                                           const int64_t* cols, int n, SpecD sp, cudaStream_t st) {
  BlockMultiDesc md;
  md.n = n;
  md.tile_start[0] = 0;
  for (int t = 0; t < n; ++t) {
    if (!ins[t] || !outs[t] || rows[t] < 0 || cols[t] < 0) return LCB_E_INVALID;
    if (cols[t] % 8 || rows[t] >= (1ll << 31) || cols[t] >= (1ll << 31)) return LCB_E_UNSUPPORTED;
    if (!aligned16(ins[t]) || !aligned16(outs[t])) return LCB_E_ALIGN;
    const int64_t tiles = ((rows[t] + 127) / 128) * ((cols[t] + 127) / 128);
    md.tile_start[t + 1] = md.tile_start[t] + tiles;
    md.out[t] = outs[t];
    md.scales[t] = scales ? scales[t] : nullptr;
    md.rows[t] = rows[t];
    md.cols[t] = cols[t];
    if (tiles > 0 && !encode_block_tmap(&md.tm[t], ins[t], TI<T>::ID, rows[t], cols[t])) {
      set_last_error("cuTensorMapEncodeTiled", cudaErrorInvalidValue);
      return LCB_E_CUDA;
    }
    if (tiles == 0) md.tm[t] = md.tm[0];
  }
  for (int t = n; t < kMaxBlockMulti; ++t) {
    md.tm[t] = md.tm[0];
    md.out[t] = md.scales[t] = nullptr;
    md.rows[t] = md.cols[t] = 0;
    md.tile_start[t + 1] = md.tile_start[n];
  }
  const int64_t tiles = md.tile_start[n];
  if (tiles == 0) return LCB_OK;
  if (tiles > 0x7fffffff) return LCB_E_SHAPE;
  const bool e5 = sp.kind == LCB_KIND_FP8_E5M2, codes = sp.out_mode == LCB_OUT_FP8;
  if (!e5 && !codes) fq_block2d_fp8_tma_multi_kernel<T, 0, 0><<<(int)tiles, 256, 0, st>>>(md, sp);
  else if (!e5 && codes) fq_block2d_fp8_tma_multi_kernel<T, 0, 1><<<(int)tiles, 256, 0, st>>>(md, sp);
  else if (e5 && !codes) fq_block2d_fp8_tma_multi_kernel<T, 1, 0><<<(int)tiles, 256, 0, st>>>(md, sp);
  else fq_block2d_fp8_tma_multi_kernel<T, 1, 1><<<(int)tiles, 256, 0, st>>>(md, sp);
  LCB_CUDA_CHECK_LAUNCH("fq_block2d_fp8_tma_multi");
  return LCB_OK;
}

// ---- several tensors, one launch (per_group / short-row dynamic fake-quant) ----------------------------------------------
template <typename T>
static int fakequant_multi_typed(const void* const* ins, void* const* outs, const int64_t* rows, const int64_t* cols, int n,
                                 int64_t uc, SpecD sp, cudaStream_t st) {
  constexpr int VEC = TI<T>::VEC;
  if (sp.kind != LCB_KIND_INT || sp.out_mode != LCB_OUT_FAKE || sp.math_dtype != sp.dtype) return LCB_E_UNSUPPORTED;
  if (uc % VEC) return LCB_E_UNSUPPORTED;
  const int64_t nvec = uc / VEC;
  if (nvec > 64) return LCB_E_UNSUPPORTED;
  int lanes = 4;
  while (lanes * 2 < nvec && lanes < 32) lanes <<= 1;
  const int nv = (int)((nvec + lanes - 1) / lanes);
  const int nvt = nv <= 1 ? 1 : 2;
  const int u = 8 / nvt;
  const int64_t upw = (int64_t)u * (32 / lanes);
  MultiDesc md;
  md.n = n;
  md.chunk_start[0] = 0;
  for (int t = 0; t < n; ++t) {
    if (!ins[t] || !outs[t] || rows[t] < 0 || cols[t] < 0) return LCB_E_INVALID;
    if (cols[t] % uc) return LCB_E_SHAPE;
    if (!aligned16(ins[t]) || !aligned16(outs[t])) return LCB_E_ALIGN;
    md.in[t] = ins[t];
    md.out[t] = outs[t];
    md.n_units[t] = rows[t] * (cols[t] / uc);
    md.chunk_start[t + 1] = md.chunk_start[t] + (md.n_units[t] + upw - 1) / upw;
  }
  for (int t = n; t < kMaxMulti; ++t) {
    md.in[t] = nullptr;
    md.out[t] = nullptr;
    md.n_units[t] = 0;
    md.chunk_start[t + 1] = md.chunk_start[n];
  }
  const int64_t chunks = md.chunk_start[n];
  if (chunks == 0) return LCB_OK;
  const FastSpec fs = make_fastspec<T>(sp);
  static const int multi_cps = [] { const char* e = getenv("LCB_GROUP_CTAS_PER_SM"); return e ? atoi(e) : 8; }();
  const int grid = grid_for((chunks + 7) / 8, multi_cps);
#define LCB_LAUNCH_MULTI(NVc, CH) \
  launch_pdl(fq_group_multi_kernel<T, NVc, 0, CH>, grid, 256, st, md, (int)uc, lanes, sp, fs)
#define LCB_LAUNCH_MULTI_NV(CH)              \
  {                                          \
    if (nvt == 1) LCB_LAUNCH_MULTI(1, CH);   \
    else LCB_LAUNCH_MULTI(2, CH);            \
  }
  if (fs.chain == CHAIN_F32) LCB_LAUNCH_MULTI_NV(CHAIN_F32)
  else if (TI<T>::ID == LCB_BF16 && fs.chain == CHAIN_BF16_SMALL) {
    if constexpr (TI<T>::ID == LCB_BF16) LCB_LAUNCH_MULTI_NV(CHAIN_BF16_SMALL)
  } else if (TI<T>::ID == LCB_BF16 && fs.chain == CHAIN_BF16_WIDE) {
    if constexpr (TI<T>::ID == LCB_BF16) LCB_LAUNCH_MULTI_NV(CHAIN_BF16_WIDE)
  } else if (TI<T>::ID == LCB_F16 && fs.chain == CHAIN_F16) {
    if constexpr (TI<T>::ID == LCB_F16) LCB_LAUNCH_MULTI_NV(CHAIN_F16)
  } else return LCB_E_UNSUPPORTED;
#undef LCB_LAUNCH_MULTI_NV
#undef LCB_LAUNCH_MULTI
  LCB_CUDA_CHECK_LAUNCH("fq_group_multi");
  return LCB_OK;
}
}  // namespace lcb

extern "C" int lcb_fakequant_multi(const void* const* ins, void* const* outs, const int64_t* rows, const int64_t* cols,
                                   int32_t n, int64_t unit_cols, const lcb_quant_spec* spec, void* stream) {
  using namespace lcb;
  if (!ins || !outs || !rows || !cols || !spec || n < 0 || unit_cols < 1) return LCB_E_INVALID;
  if (n == 0) return LCB_OK;
  if (n > kMaxMulti) return LCB_E_UNSUPPORTED;
  if (!(spec->qmax > spec->qmin)) return LCB_E_INVALID;
  SpecD sp = make_specd(spec);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  switch (spec->dtype) {
    case LCB_F32: return fakequant_multi_typed<float>(ins, outs, rows, cols, n, unit_cols, sp, st);
    case LCB_F16: return fakequant_multi_typed<__half>(ins, outs, rows, cols, n, unit_cols, sp, st);
    case LCB_BF16: return fakequant_multi_typed<__nv_bfloat16>(ins, outs, rows, cols, n, unit_cols, sp, st);
    default: return LCB_E_UNSUPPORTED;
  }
}

extern "C" int lcb_fakequant_block_fp8_multi(const void* const* ins, void* const* outs, void* const* scales,
                                             const int64_t* rows, const int64_t* cols, int32_t n, const lcb_quant_spec* spec,
                                             void* stream) {
  using namespace lcb;
  if (!ins || !outs || !rows || !cols || !spec || n < 0) return LCB_E_INVALID;
  if (n == 0) return LCB_OK;
  if (n > kMaxBlockMulti) return LCB_E_UNSUPPORTED;
  if (!(spec->qmax > spec->qmin)) return LCB_E_INVALID;
  if ((spec->kind != LCB_KIND_FP8_E4M3 && spec->kind != LCB_KIND_FP8_E5M2) || !spec->sym ||
      (spec->out_mode != LCB_OUT_FAKE && spec->out_mode != LCB_OUT_FP8) || spec->math_dtype != LCB_F32 ||
      spec->output_scale_factor != 1.0f || spec->clip_ratio != 1.0f)
    return LCB_E_UNSUPPORTED;
  SpecD sp = make_specd(spec);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  switch (spec->dtype) {
    case LCB_F16: return fakequant_block_fp8_multi_typed<__half>(ins, outs, scales, rows, cols, n, sp, st);
    case LCB_BF16: return fakequant_block_fp8_multi_typed<__nv_bfloat16>(ins, outs, scales, rows, cols, n, sp, st);
    default: return LCB_E_UNSUPPORTED;
  }
}

extern "C" int64_t lcb_fakequant_workspace_bytes(void) { return (int64_t)(2 * lcb::kMaxParts + 8) * sizeof(float); }

extern "C" int lcb_fakequant(const void* in, void* out, void* scales, void* zeros, int64_t rows, int64_t cols,
                             int64_t unit_rows, int64_t unit_cols, const lcb_quant_spec* spec, int dynamic,
                             void* workspace, void* stream) {
  using namespace lcb;
  if (!in || !spec || rows < 0 || cols < 0 || unit_rows < 1 || unit_cols < 1) return LCB_E_INVALID;
  if (spec->out_mode != LCB_OUT_NONE && !out) return LCB_E_INVALID;
  if (spec->kind < LCB_KIND_INT || spec->kind > LCB_KIND_FP8_E5M2) return LCB_E_UNSUPPORTED;
  if (spec->dtype < LCB_F32 || spec->dtype > LCB_BF16 || spec->math_dtype < LCB_F32 || spec->math_dtype > LCB_BF16 ||
      spec->qparam_dtype < LCB_F32 || spec->qparam_dtype > LCB_BF16)
    return LCB_E_UNSUPPORTED;
  if (spec->out_mode < LCB_OUT_FAKE || spec->out_mode > LCB_OUT_DEQUANT) return LCB_E_UNSUPPORTED;
  if (spec->out_mode == LCB_OUT_DEQUANT && dynamic) return LCB_E_INVALID;
  if (spec->kind != LCB_KIND_INT && !(spec->out_mode == LCB_OUT_FAKE || spec->out_mode == LCB_OUT_FP8 ||
                                      spec->out_mode == LCB_OUT_NONE || spec->out_mode == LCB_OUT_CODE_T ||
                                      spec->out_mode == LCB_OUT_DEQUANT))
    return LCB_E_INVALID;
  if (spec->kind == LCB_KIND_INT && spec->out_mode == LCB_OUT_FP8) return LCB_E_INVALID;
  if (!(spec->qmax > spec->qmin)) return LCB_E_INVALID;
  if (spec->clip_ratio != 1.0f && spec->clip_ratio != -1.0f) {   // AutoClip v1: dynamic 1-D units, math in the tensor dtype
    if (!(fabsf(spec->clip_ratio) > 0.f && fabsf(spec->clip_ratio) <= 1.f)) return LCB_E_INVALID;
    if (!dynamic || unit_rows != 1 || spec->math_dtype != spec->dtype) return LCB_E_UNSUPPORTED;
  }
  if (rows == 0 || cols == 0) return LCB_OK;
  // ragged units are only defined for per_block (zero padding); 1-D groups must divide (quant.py:627-630)
  if (unit_rows == 1 && unit_cols < cols && cols % unit_cols != 0) return LCB_E_SHAPE;
  SpecD sp = make_specd(spec);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  switch (spec->dtype) {
    case LCB_F32: return fakequant_typed<float>(in, out, scales, zeros, rows, cols, unit_rows, unit_cols, sp, dynamic, workspace, st);
    case LCB_F16: return fakequant_typed<__half>(in, out, scales, zeros, rows, cols, unit_rows, unit_cols, sp, dynamic, workspace, st);
    default: return fakequant_typed<__nv_bfloat16>(in, out, scales, zeros, rows, cols, unit_rows, unit_cols, sp, dynamic, workspace, st);
  }
}

// ---- call plans: the shape / spec part of an lcb_fakequant call, validated once ------------------------------------------
// A Python caller pays ~0.2 us per marshalled argument; a plan turns the 12-argument call into 7 pointers. Same kernels,
// same checks (they run on every call: a plan is only a bundle of arguments).
struct lcb_fq_plan_s {
  int64_t rows, cols, unit_rows, unit_cols;
  lcb_quant_spec spec;
  int dynamic;
};

extern "C" void* lcb_fakequant_plan_create(int64_t rows, int64_t cols, int64_t unit_rows, int64_t unit_cols,
                                           const lcb_quant_spec* spec, int dynamic) {
  if (!spec || rows < 0 || cols < 0 || unit_rows < 1 || unit_cols < 1) return nullptr;
  lcb_fq_plan_s* p = new (std::nothrow) lcb_fq_plan_s;
  if (!p) return nullptr;
  p->rows = rows;
  p->cols = cols;
  p->unit_rows = unit_rows;
  p->unit_cols = unit_cols;
  p->spec = *spec;
  p->dynamic = dynamic;
  return p;
}

extern "C" int lcb_fakequant_plan_run(const void* plan, const void* in, void* out, void* scales, void* zeros,
                                      void* workspace, void* stream) {
  if (!plan) return LCB_E_INVALID;
  const lcb_fq_plan_s* p = reinterpret_cast<const lcb_fq_plan_s*>(plan);
  return lcb_fakequant(in, out, scales, zeros, p->rows, p->cols, p->unit_rows, p->unit_cols, &p->spec, p->dynamic, workspace,
                       stream);
}

extern "C" void lcb_fakequant_plan_destroy(void* plan) { delete reinterpret_cast<lcb_fq_plan_s*>(plan); }
