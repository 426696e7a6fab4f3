/*
 * lcb200.h — C ABI of the B200-native LightCompress/LLMC quantization hot path.
 *
 * Every entry point takes plain device pointers, sizes and a CUDA stream handle
 * (passed as void* so that the header compiles without the CUDA toolkit). No
 * torch types cross this boundary. Every function returns LCB_OK (0) or a negative
 * LCB_E_* code; nothing allocates, nothing synchronises the host.
 *
 * The reference (zhangbilang/LightCompress, all-Python) has no FFI; each function
 * below names the reference Python interface it replaces (file:line under
 * /root/reference/llmc/compression/quantization/).
 */
#ifndef LCB200_H_
#define LCB200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LCB_ABI_VERSION 1

/* status codes (the Python shim maps them onto the reference's exception types) */
#define LCB_OK 0
#define LCB_E_INVALID (-1)     /* bad argument combination      -> AssertionError   */
#define LCB_E_SHAPE (-2)       /* non-divisible group etc.       -> ValueError       */
#define LCB_E_UNSUPPORTED (-3) /* unknown format                 -> NotImplementedError */
#define LCB_E_CUDA (-4)        /* launch failed                  -> RuntimeError     */
#define LCB_E_ALIGN (-5)       /* pointer not aligned as stated  -> ValueError       */

/* tensor storage dtypes */
#define LCB_F32 0
#define LCB_F16 1
#define LCB_BF16 2

/* quantization kind */
#define LCB_KIND_INT 0      /* IntegerQuantizer           quant.py:661  */
#define LCB_KIND_FP8_E4M3 1 /* FloatQuantizer bit='e4m3'  quant.py:963  */
#define LCB_KIND_FP8_E5M2 2 /* FloatQuantizer bit='e5m2'                */

/* what the kernel writes to `out` */
#define LCB_OUT_FAKE 0   /* quant-dequant value, storage dtype (fake_quant_*)          */
#define LCB_OUT_I8 1     /* integer code as int8   (real_quant, bit==8, qmin!=0)       */
#define LCB_OUT_U8 2     /* integer code as uint8  (real_quant, bit==8, qmin==0)       */
#define LCB_OUT_I32 3    /* integer code as int32  (real_quant, any other bit)         */
#define LCB_OUT_FP8 4    /* fp8 byte (e4m3fn / e5m2) (FloatQuantizer.real_quant_*)     */
#define LCB_OUT_NONE 5   /* nothing: qparams only  (get_tensor_qparams)                */
#define LCB_OUT_CODE_T 6 /* integer code kept in the math dtype (IntegerQuantizer.quant) */
#define LCB_OUT_DEQUANT 7 /* `in` holds codes in the math dtype; out = (q - z) * s (IntegerQuantizer.dequant, quant.py:710-712);
                             static calls only */

/*
 * Quantizer description. Mirrors the ctor kwargs of BaseQuantizer/IntegerQuantizer/
 * FloatQuantizer (quant.py:47-101, 662-678, 964-1003).
 *
 * Arithmetic model (SURVEY.md §8a "exact arithmetic spec"): every torch op of the
 * reference chain is "fp32 compute, round result to the math dtype". `math_dtype`
 * is that dtype: the tensor dtype for per_tensor/per_channel/per_group/per_token
 * dynamic quantization, LCB_F32 for per_block (quant.py:137-138 upcasts) and for
 * static calls whose scales are fp32 (type promotion).
 */
typedef struct lcb_quant_spec {
  int32_t dtype;       /* storage dtype of in / LCB_OUT_FAKE out                   */
  int32_t math_dtype;  /* dtype every intermediate is rounded to                   */
  int32_t qparam_dtype;/* storage dtype of the scales / zeros arrays               */
  int32_t kind;        /* LCB_KIND_*                                                */
  int32_t sym;         /* symmetric (zeros == 0)                                    */
  int32_t round_zp;    /* quant.py:73, :556-558                                     */
  int32_t out_mode;    /* LCB_OUT_*                                                 */
  int32_t qp_math_plus1; /* 0: scales/zeros are computed in math_dtype. d+1: the divisions of get_qparams round to dtype d
                            (per_tensor: 0-dim min/max divided by a 0-dim fp32 qmax promote to fp32, quant.py:552-556) */
  float qmin;          /* quant.py:666-677 (int) / :979-1003 (float)                */
  float qmax;
  float output_scale_factor; /* quant.py:716 (dequant only), 1.0f if unused        */
  float clip_ratio;    /* AutoClip v1 (auto_clip.py:139-146, 258-260): every unit is clamped to [min * r, max * r] before it is
                          quantized (r = |clip_ratio|; clip_ratio < 0 = clip_sym: [-absmax * r, absmax * r]); 1.0f = off.
                          Dynamic 1-D units only. */
} lcb_quant_spec;

/*
 * Grouped quant-dequant, the single-pass kernel.
 * The tensor is a row-major [rows, cols] matrix with leading dimension `cols`
 * (contiguous). A quantization unit is a `unit_rows x unit_cols` tile:
 *   per_group   : 1 x group_size        per_channel / per_token : 1 x cols
 *   per_tensor  : rows x cols            per_block               : bs x bs (ragged edges = zero padded)
 * dynamic != 0 : min/max/absmax -> qparams -> round -> clamp -> dequant in one pass;
 *                scales/zeros (may be NULL) receive the qparams, one per unit, row-major over the
 *                unit grid [ceil(rows/unit_rows), ceil(cols/unit_cols)].
 * dynamic == 0 : scales (and zeros unless sym) are inputs.
 * workspace    : >= lcb_fakequant_workspace_bytes() bytes, only used by per_tensor dynamic.
 *
 * Replaces IntegerQuantizer.get_tensor_qparams + quant_dequant / quant (quant.py:690-717),
 * fake_quant_{weight,act}_{dynamic,static} (quant.py:719-869), real_quant_weight_{dynamic,static}
 * (quant.py:871-953), FloatQuantizer equivalents (quant.py:1043-1221) and
 * weight_cast_to_fp8 (quant.py:34-43, kernel.py:56-83).
 */
int lcb_fakequant(const void* in, void* out, void* scales, void* zeros, int64_t rows, int64_t cols,
                  int64_t unit_rows, int64_t unit_cols, const lcb_quant_spec* spec, int dynamic,
                  void* workspace, void* stream);

int64_t lcb_fakequant_workspace_bytes(void);

/*
 * Call plans for lcb_fakequant: (rows, cols, unit_rows, unit_cols, spec, dynamic) bundled once, so that the per-call
 * binding cost of a Python / JNI / cgo caller is 7 pointers instead of 12 mixed arguments — the reference issues one
 * fake_quant_act_dynamic per linear per forward (module_utils.py:799-813: 224 calls per Llama-3-8B forward), where the
 * host cost per call bounds the achieved bandwidth. lcb_fakequant_plan_run(plan, ...) is exactly
 * lcb_fakequant(in, out, scales, zeros, <plan>, workspace, stream): same checks, same kernels, same status codes.
 * plan_create returns NULL on invalid arguments; plans are immutable and may be shared between threads and streams.
 */
void* lcb_fakequant_plan_create(int64_t rows, int64_t cols, int64_t unit_rows, int64_t unit_cols,
                                const lcb_quant_spec* spec, int dynamic);
int lcb_fakequant_plan_run(const void* plan, const void* in, void* out, void* scales, void* zeros, void* workspace,
                           void* stream);
void lcb_fakequant_plan_destroy(void* plan);

/*
 * Range of every quantization unit (same unit grid as lcb_fakequant): mins / maxs receive amin / amax over the unit in
 * `range_dtype` (the tensor dtype; LCB_F32 with abs_range != 0 = per_block, which ranges |x|.float() and counts the zero
 * padding of ragged edge blocks, quant.py:137-138, 637-638). Replaces get_minmax_range (quant.py:132-143).
 * workspace: NULL or lcb_fakequant_workspace_bytes() bytes (lets a whole-tensor range use every SM).
 */
int lcb_minmax(const void* in, void* mins, void* maxs, int64_t rows, int64_t cols, int64_t unit_rows, int64_t unit_cols,
               int32_t dtype, int32_t range_dtype, int32_t abs_range, void* workspace, void* stream);

/*
 * scales / zeros of n units from given (min, max) ranges: get_qparams (quant.py:545-559) for ranges that were modified
 * after the search (learnable bounds `sigmoid(f) * min / max`, quant.py:205-219; AutoClip v2, static calibration).
 * mins / maxs in spec->math_dtype, scales / zeros in spec->qparam_dtype; zeros may be NULL when spec->sym.
 */
int lcb_qparams_from_range(const void* mins, const void* maxs, void* scales, void* zeros, int64_t n,
                           const lcb_quant_spec* spec, void* stream);

/*
 * calib_algo = 'mse': the shrink-grid range search of BaseQuantizer.get_mse_range (quant.py:145-203) on a contiguous
 * [n_units, unit_len] tensor of spec->dtype (the reshape_tensor() layout of per_group / per_channel / per_token /
 * per_head; the reference itself only handles 2-D row units here). For i in [0, n_steps) — n_steps =
 * int(maxshrink * mse_grid) — the candidate range is (1 - i / mse_grid) * (min, max) of the range kept so far (the
 * reference writes the kept range through the tensors the next candidate is derived from, quant.py:164, 193-195), its
 * error sum |quant_dequant(x) - x| ** norm in fp32, kept when strictly smaller. mins / maxs: fp32 [n_units]
 * (the search runs on tensor.float(), quant.py:151). Uses spec->kind / sym / round_zp / qmin / qmax / dtype only.
 * unit_len * 4 bytes must fit the shared memory of one CTA (unit_len <= 51200). LCB_KIND_INT only: on a shrunk range the
 * fp8 grid overflows and the result is defined by qtorch.float_quantize, which cannot be pinned offline.
 */
int lcb_mse_range(const void* in, void* mins, void* maxs, int64_t n_units, int64_t unit_len, const lcb_quant_spec* spec,
                  int32_t n_steps, double mse_grid, float norm, void* stream);

/*
 * Dynamic fake-quant (LCB_OUT_FAKE, IntegerQuantizer, units of 1 x unit_cols with unit_cols * sizeof(dtype) <= 1 KiB)
 * of up to 16 tensors in ONE launch — e.g. the 7 linear weights of a transformer block at deploy time
 * (base_bq.py:899-1006 calls w_qdq once per linear). ins / outs / rows / cols are HOST arrays of n entries; every
 * tensor is a contiguous [rows[i], cols[i]] matrix of spec->dtype with cols[i] % unit_cols == 0. Same arithmetic as
 * lcb_fakequant. Returns LCB_E_UNSUPPORTED for specs outside that envelope (callers then loop over lcb_fakequant).
 */
int lcb_fakequant_multi(const void* const* ins, void* const* outs, const int64_t* rows, const int64_t* cols, int32_t n,
                        int64_t unit_cols, const lcb_quant_spec* spec, void* stream);

/*
 * Dynamic FP8 per_block (128 x 128, fp32 math and scales: weight_cast_to_fp8 / FloatQuantizer per_block, quant.py:34-43,
 * 1043-1221) of up to 64 bf16 / f16 weights in ONE launch — the expert matrices of a MoE block. spec->out_mode is
 * LCB_OUT_FAKE (outs[i]: spec->dtype) or LCB_OUT_FP8 (outs[i]: one byte per element); scales may be NULL, or an array whose
 * entries are NULL or fp32 [ceil(rows/128) * ceil(cols/128)]. cols[i] % 8 == 0 (16-byte row pitch for TMA). Same
 * arithmetic as lcb_fakequant. Returns LCB_E_UNSUPPORTED outside that envelope (callers then loop over lcb_fakequant).
 */
int lcb_fakequant_block_fp8_multi(const void* const* ins, void* const* outs, void* const* scales, const int64_t* rows,
                                  const int64_t* cols, int32_t n, const lcb_quant_spec* spec, void* stream);

/*
 * FloatQuantizer WITHOUT qtorch (`use_qtorch` unset: configs/quantization/methods/FP_Quant/rtn_we4m3ae4m3.yml and
 * friends): per-ELEMENT scales. Replaces FloatQuantizer.get_float_qparams (quant.py:1005-1027) + quant / dequant
 * (quant.py:1061-1076) for the 1-D granularities (per_channel / per_token / per_group / per_head):
 *   maxval = max(max, -min) of the unit;  bias = 2^e - log2(maxval) + log2(2 - 2^-m) - 1;
 *   xc = clamp(x, -maxval, maxval);  s = 2 ** (clamp(floor(log2|xc| + bias), min = 1) - m - bias);  s[s == 0] = 1;
 *   y = (round(xc / s + 0) - 0) * s
 * with one rounding to the tensor dtype after every op (e_bits < 5) or an fp32 chain and one final cast (e_bits >= 5,
 * quant.py:1016-1017; f32 tensors). in: [n_units, unit_size] contiguous in `dtype`. out (fake-quantized, `dtype`),
 * xc_out and scales_out (what get_tensor_qparams returns: `dtype` when e_bits < 5, f32 otherwise) may each be NULL.
 * An all-zero unit yields NaN like the reference (log2(0) + inf).
 * lcb_floatquant_elem_static: y = (round(x / s + 0) - 0) * s with caller-provided per-element scales
 * (fake_quant_*_static with the scales of get_tensor_qparams, quant.py:1083-1140); scales_dtype = dtype or LCB_F32.
 */
int lcb_floatquant_elem(const void* in, void* out, void* xc_out, void* scales_out, int64_t n_units, int64_t unit_size,
                        int32_t e_bits, int32_t m_bits, int32_t dtype, void* stream);
int lcb_floatquant_elem_static(const void* in, const void* scales, void* out, int64_t n, int32_t dtype, int32_t scales_dtype,
                               void* stream);

/*
 * Block-wise FP8 dequantisation y = x_fp8 * s[block] -> bf16/f16/f32
 * Replaces weight_cast_to_bf16 (quant.py:18-31, kernel.py:86-138).
 */
int lcb_fp8_block_dequant(const void* codes, const float* scales, void* out, int64_t rows,
                          int64_t cols, int64_t block, int32_t kind, int32_t out_dtype,
                          void* stream);

/*
 * act_quant (kernel.py:7-54, the Triton kernel of LlmcFp8Linear's activation side, module_utils.py:41-43): x is a
 * contiguous tensor of n_groups * group elements (bf16 / f16 / f32); for every group of `group` consecutive elements
 * s = max|x| / 448 (fp32) and y = e4m3(x / s) (fp32 division, RNE cast). An all-zero group yields s = 0 and NaN codes
 * (0 / 0), as the reference does. group in {8, 16, 32, 64, 128, 256}. y: uint8 [n_groups * group], scales fp32 [n_groups].
 */
int lcb_act_quant_fp8(const void* x, void* y, float* scales, int64_t n_groups, int64_t group, int32_t dtype,
                      void* stream);

/*
 * fp8_gemm (kernel.py:140-242) with the bias add of block_wise_fp8_forward_func (module_utils.py:41-46):
 *   c[m, n] = out_dtype( sum_kb  dot(a[m, kb], b[n, kb]) * a_s[m, kb] * b_s[n / 128, kb] )  (+ bias[n], rounded again)
 * a: e4m3 [M, K] row-major, a_s fp32 [M, ceil(K/128)]; b: e4m3 [N, K] row-major (a Linear weight), b_s fp32
 * [ceil(N/128), ceil(K/128)]; c: [M, N] of out_dtype; bias NULL or [N] of bias_dtype. The K block is 128 as in the
 * reference kernel (BLOCK_SIZE_K = 128, b_s indexed by offs_n // 128). fp32 accumulation: each K block is contracted by
 * the tensor pipe into TMEM and promoted with one FMA by (a_s * b_s). K % 16 == 0 and 16-byte aligned a / b (TMA).
 */
int lcb_fp8_gemm(const void* a, const float* a_s, const void* b, const float* b_s, void* c, const void* bias, int64_t M,
                 int64_t N, int64_t K, int32_t out_dtype, int32_t bias_dtype, void* stream);
/*
 * The same GEMM with a caller workspace of lcb_fp8_gemm_workspace_bytes(M, N, K) bytes (0: no workspace needed, NULL is
 * fine), 16-byte aligned, which MUST BE ZERO-FILLED before its first use and is LEFT ZERO-FILLED by every call (so one
 * zeroed buffer serves any number of calls that are ordered on a stream; calls that overlap on the GPU need their own).
 * With few output tiles (decode shapes: M <= 128 rows leave N / 256 CTAs for 148 SMs, and the call is bound by streaming
 * the weight) the K blocks of a tile are split over the idle SMs; every slice adds its fp32 partial tile into the
 * workspace (16-byte L2 reductions), the slice that arrives last reads the sums back, rounds to out_dtype, adds the bias,
 * stores the tile and clears sums and counter: ONE launch. Same contraction and promotion per K block; the fp32 sum over
 * K blocks runs in slice order instead of block order (differences at fp32 rounding level, not deterministic between
 * runs). No counterpart in the reference (its Triton kernel has no split-K).
 */
int64_t lcb_fp8_gemm_workspace_bytes(int64_t M, int64_t N, int64_t K);
int lcb_fp8_gemm_ws(const void* a, const float* a_s, const void* b, const float* b_s, void* c, const void* bias, int64_t M,
                    int64_t N, int64_t K, int32_t out_dtype, int32_t bias_dtype, void* workspace, void* stream);

/*
 * GPTQ Hessian accumulation  S += X^T X   (upper 128x128 blocks of S only; SYRK)
 * X is the layer input flattened to [tokens, C] row-major (the reference transposes a view of it, gptq.py:265-267),
 * f32 / f16 / bf16. S is fp32 [C, C] row-major, zero-initialised by the caller, and holds the UN-NORMALISED running
 * sum: the reference's H after k batches equals (2/k) * sum_{j<=k} X_j^T X_j (gptq.py:283-290: H *= n/(n+b);
 * H += (2/n) X^T X), so the scale is applied once by lcb_hessian_finalize instead of once per batch. Any number
 * of tokens may be passed per call (several calibration batches at once). Only 128x128 blocks (i, j >= i) of S are
 * written; lcb_hessian_finalize mirrors them.
 * The contraction runs on tcgen05 tensor cores, fp32 accumulation in TMEM: 16-bit inputs multiply exactly; f32 input
 * is split into bf16 hi + lo (|x - hi - lo| <= 2^-17 |x|) and accumulated as hi*hi + hi*lo + lo*hi.
 * workspace: lcb_hessian_workspace_bytes(tokens, C, x_dtype) bytes of device scratch (0 for 16-bit inputs).
 * Multi-GPU: every rank accumulates its own token shard into its own S; ONE all-reduce(SUM) of S per linear, then
 * finalize with scale = 2 / (total batches) — replaces the per-batch all-reduce of gptq.py:292-295.
 * Replaces GPTQ.add_batch (gptq.py:254-295).
 */
int lcb_hessian_accum(float* S, const void* X, int64_t tokens, int64_t C, int32_t x_dtype, void* workspace,
                      void* stream);
int64_t lcb_hessian_workspace_bytes(int64_t tokens, int64_t C, int32_t x_dtype);
/*
 * Allocation-free variant of lcb_hessian_accum for f32 X. No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid
 * it). Same arithmetic, kernels and S contract as lcb_hessian_accum(S, X, tokens, C, LCB_F32, workspace, stream); the
 * bf16 hi / lo split buffers live in a 16 MiB static __device__ array inside the library (held by every GPU that loads
 * it until exit) instead of a caller workspace. The call allocates nothing (no cudaMalloc / cudaFree, no allocating
 * library routine). Limit: tokens * C <= LCB_HESSIAN_NOALLOC_MAX_ELEMS (e.g. 1024 tokens x 4096 channels); beyond it
 * the call returns LCB_E_SHAPE with the reason in lcb_last_error() and launches nothing - split the tokens over several
 * calls (S is a running sum). Every call on one GPU shares the array: calls that overlap on the GPU (different streams
 * or threads) race; calls ordered on one stream do not. Nothing in the array has to start at zero.
 */
#define LCB_HESSIAN_NOALLOC_MAX_ELEMS (4ll << 20)
int lcb_hessian_accum_f32_noalloc(float* S, const float* X, int64_t tokens, int64_t C, void* stream);
/* H = scale * sym(S) in place: scales the upper blocks and mirrors them into the lower ones. scale = 2 / nsamples: the
 * normalisation gptq.py:283-290 applies per batch (H *= n / (n + b); H += (2 / n) X^T X), applied once. */
int lcb_hessian_finalize(float* S, int64_t C, double scale, void* stream);
/*
 * Token-sharded Hessians (replaces the all-reduce per batch of gptq.py:292-295 by ONE per linear):
 * lcb_hessian_pack copies the maintained part of S (block row i keeps columns i*128 .., i.e. half of S) into a contiguous
 * buffer of lcb_hessian_packed_elems(C) floats whose LAST element is this rank's batch count; the caller all-reduces
 * (SUM) that buffer; lcb_hessian_unpack_finalize writes the sums back and applies H = (2 / total count) * sym(S) with the
 * count read on the device — no second collective, no host synchronisation.
 */
int64_t lcb_hessian_packed_elems(int64_t C);
int lcb_hessian_pack(const float* S, float* packed, int64_t C, double count, void* stream);
int lcb_hessian_unpack_finalize(float* S, const float* packed, int64_t C, void* stream);

/*
 * GPTQ blocked column update (gptq.py:199-244).
 * W [R, C] fp32 row-major is updated in place (the reference mutates W too), Hinv [C, C] fp32 is the
 * upper Cholesky factor of the inverse damped Hessian (gptq.py:172-174).
 * Out (`tmp`, gptq.py:237) receives the error-compensated float columns, Losses the per-element
 * loss (gptq.py:238). For per_group dynamic (static_groups == 0) group qparams are searched on the
 * outer W at every group boundary (gptq.py:215-223) and written to gscales / gzeros, laid out
 * [C/group, R] fp32. For static_groups / per_channel the same arrays are inputs
 * (group index per column from `col_group`, int32 [C], may be NULL = col / group).
 * workspace: lcb_gptq_workspace_bytes(R, C, blocksize).
 * Replaces GPTQ.weight_transform + search_column_qparams (gptq.py:199-244, 359-366).
 */
int lcb_gptq_weight_transform(float* W, const float* Hinv, float* Out, float* Losses, float* gscales,
                              float* gzeros, const int32_t* col_group, int64_t R, int64_t C,
                              int64_t n_nonout, int32_t blocksize, int32_t group_size,
                              int32_t static_groups, const lcb_quant_spec* spec, void* workspace,
                              void* stream);
int64_t lcb_gptq_workspace_bytes(int64_t R, int64_t C, int32_t blocksize);

/*
 * In place: H <- U, the UPPER Cholesky factor of inverse(H + percdamp * mean(diag H) * I), i.e. U^T U = (H + damp I)^-1.
 * Replaces the damping and the three factorisations of GPTQ.process_hessian_and_weights (gptq.py:165-174:
 * torch.linalg.cholesky -> torch.cholesky_inverse -> torch.linalg.cholesky(upper=True)) by one blocked reverse
 * Cholesky H = R R^T (R upper) and one blocked triangular inverse U = R^-1, whose O(n^3) parts run as tcgen05 GEMMs
 * over bf16 pieces of the fp32 operands (splits = 2: hi + lo, 3 products; splits = 3: hi + mid + lo, 6 products);
 * diagonal blocks, substitutions and accumulators are fp32.
 * H: fp32 [n, n] symmetric positive definite, leading dimension ld; only its upper triangle is read; on return the
 * strict lower triangle is zero. n % 128 == 0 requires ld == n and a 16-byte aligned H (factorised in place);
 * any other n / ld goes through a padded copy inside the workspace.
 * info: device int32, 0 on success, else 1 + index of the first non-positive pivot (the reference raises
 * torch.linalg.LinAlgError there; the host wrapper does the same after reading info).
 * workspace: lcb_cholinv_workspace_bytes(n, splits) bytes.
 */
int lcb_cholinv_upper(float* H, int64_t n, int64_t ld, float percdamp, int32_t splits, void* workspace, int32_t* info,
                      void* stream);
int64_t lcb_cholinv_workspace_bytes(int64_t n, int32_t splits);

/*
 * out[r, c] = in[r, c] * scales[c] (op 0) or in[r, c] / scales[c] (op 1, IEEE division), computed in fp32 and rounded
 * to `dtype` — the tensor-dtype arithmetic of Awq.scaling_weight (awq.py:41-47: W.mul_(scales)) and
 * scaling_input (base_bq.py:877-889: x / scales). in / out / scales share `dtype`; out may alias in.
 */
int lcb_scale_cols(const void* in, const void* scales, void* out, int64_t rows, int64_t cols, int32_t dtype, int32_t op,
                   void* stream);

/*
 * loss[0] += sum_{t, r} rd(org_out[t, r] - rd(sum_c x[t, c] * w[r, c]))^2     (rd = round to `dtype`, bf16 / f16)
 * = F.linear(x, w) followed by the squared-error sum of Awq.calculate_loss (awq.py:134-145: (org_out - out) is taken
 * in the tensor dtype, then .float().pow(2)); the caller divides by T * R for the mean. One tcgen05 GEMM (fp32
 * accumulation in TMEM) whose epilogue does the rounding, the subtraction and the reduction: `out` is never stored.
 * x [T, C], w [R, C], org_out [T, R] row-major contiguous; loss is a device double (accumulated with atomics).
 */
int lcb_linear_mse(const void* x, const void* w, const void* org_out, double* loss, int64_t T, int64_t R, int64_t C,
                   int32_t dtype, void* stream);

/*
 * out[t, r] = rd(sum_c x[t, c] * w[r, c] (+ bias[r]))  — torch.nn.functional.linear on 16-bit tensors, fp32 accumulation
 * on tcgen05, one rounding to the tensor dtype. The projections INSIDE the inspect modules of Awq.search_scale_subset
 * (awq.py:110-126: the whole attention / MLP module is run for every candidate) go through this entry point, so that the
 * candidate loop holds no library GEMM. x [T, C], w [R, C], out [T, R] row-major contiguous; bias NULL or [R].
 * C % 8 == 0, R % 8 == 0, 16-byte aligned bases.
 */
int lcb_linear(const void* x, const void* w, const void* bias, void* out, int64_t T, int64_t R, int64_t C, int32_t dtype,
               void* stream);

/*
 * AWQ scale search over n_cand candidates for a subset whose inspect module is the linear itself (awq.py:202-248):
 *   Wq = fake_quant(W * s_n[None, :]);  x_n = x / s_n;  loss[n] += sum (org_out - x_n Wq^T)^2      n = 0 .. n_cand-1
 * via lcb_scale_cols + lcb_fakequant + lcb_linear_mse. cand_scales [n_cand, C] in the tensor dtype (awq.py:100-107
 * computes them in the activation dtype). spec: the weight quantizer (dynamic; unit = group_size columns, or the whole
 * row when group_size <= 0), dtype = math_dtype = the tensor dtype, out_mode = LCB_OUT_FAKE. loss: device doubles,
 * zeroed by the caller. workspace: lcb_awq_workspace_bytes(T, R, C, n_cand).
 */
int lcb_awq_candidates(const void* W, const void* x, const void* org_out, const void* cand_scales, double* loss,
                       int64_t T, int64_t R, int64_t C, int32_t n_cand, int32_t group_size,
                       const lcb_quant_spec* spec, void* workspace, void* stream);
int64_t lcb_awq_workspace_bytes(int64_t T, int64_t R, int64_t C, int32_t n_cand);

/*
 * AutoClip v1 search (auto_clip.py:84-191, clip_version 'v1', w_only): for every (output channel r, group g)
 *   for i_s in 0 .. n_steps-1:  max_val = org_max * (1 - i_s / n_grid);  min_val = org_min * (1 - i_s / n_grid)  (clip_sym:
 *       org_max = |w|.amax, min_val = -max_val);  q_w = fake_quant(clamp(w, min_val, max_val));
 *       err = mean_t ((x_g q_w_g^T) - (x_g w_g^T))^2;  keep the first step with the smallest err
 * W [R, C], x [T_rows, C] (the sampled tokens, auto_clip.py:131-133) row-major in spec->dtype (bf16 / f16); rows
 * tokens .. T_rows-1 of x must be zero (padding to a multiple of 8); `tokens` is the divisor of the mean.
 * best_max / best_min: [R, C / group] in the tensor dtype (group_size <= 0: one group per row). spec: the weight
 * quantizer (dynamic, LCB_OUT_FAKE, math_dtype == dtype). The per-step output is never materialised: products rounded to
 * the tensor dtype as the reference rounds them (HMUL2), summed in fp32 on the tensor pipe (mma.sync with a ones operand),
 * the squared error reduced over tokens in registers.
 * workspace: lcb_autoclip_workspace_bytes(R, C, T_rows, group_size).
 */
int lcb_autoclip_search(const void* W, const void* x, void* best_max, void* best_min, int64_t R, int64_t C, int64_t T_rows,
                        int64_t tokens, int32_t group_size, int32_t n_grid, int32_t n_steps, int32_t clip_sym,
                        const lcb_quant_spec* spec, void* workspace, void* stream);
int64_t lcb_autoclip_workspace_bytes(int64_t R, int64_t C, int64_t T_rows, int64_t group_size);

/*
 * The same search taken apart, for the cases where the per-step weights are not "clamp then dynamic fake-quant":
 * clip_version 'v2' (auto_clip.py:258-274: logit / sigmoid bounds -> get_qparams -> fake_quant_weight_static) and w+a
 * configs (auto_clip.py:164, 276-281: the candidate outputs use the fake-quantized inputs, org_out the plain ones).
 * The host owns the buffers and the loop; the per-step weights come from lcb_fakequant (static qparams).
 *   lcb_autoclip_init: org_max / org_min fp32 [R, C / group] (clip_sym: org_max = |w|.amax), best_max / best_min (tensor
 *       dtype) = org, min_err fp32 = 1e9 rounded to the tensor dtype (auto_clip.py:116-125).
 *   lcb_autoclip_step: is_org != 0: org[g, r, t] = (x * W).sum(-1) (tensor dtype, [C / group, R, T_rows]; the bounds and
 *       best_* pointers may be NULL). is_org == 0: W = this step's fake-quantized weights, x = the (fake-quantized) inputs:
 *       err = mean_t ((x * W).sum(-1) - org)^2 with the reference's roundings; where err < min_err: min_err = err,
 *       best_max = org_max * ratio, best_min = clip_sym ? -best_max : org_min * ratio (auto_clip.py:154-179).
 * Shapes / padding rules as lcb_autoclip_search.
 */
int lcb_autoclip_init(const void* W, void* best_max, void* best_min, float* org_max, float* org_min, float* min_err,
                      int64_t R, int64_t C, int32_t group_size, int32_t clip_sym, int32_t dtype, void* stream);
int lcb_autoclip_step(const void* Wq, const void* x, void* org, const float* org_max, const float* org_min, float* min_err,
                      void* best_max, void* best_min, int64_t R, int64_t C, int64_t T_rows, int64_t tokens,
                      int32_t group_size, float ratio, int32_t clip_sym, int32_t is_org, int32_t dtype, void* stream);

/*
 * Real-quant export packers (the reference packs with numpy on the host / a Python loop per column).
 * lcb_pack_rows: VllmRealQuantLinear.pack (module_utils.py:931-955): codes [rows, cols] (int8 or int32, code_bytes = 1 | 4)
 *   -> out int32 [rows, ceil(cols / (32 / bit))], code + 2^(bit-1) as unsigned in bits [bit i, bit i + bit) of word col / pf.
 * lcb_awq_gemm_pack: AutoawqRealQuantLinear.gemm_pack (module_utils.py:1099-1158), 4 bit: W [R, C] (dtype), scales fp16
 *   [C / group, R], zeros int32 [C / group, R] -> qweight int32 [C, R / 8], qzeros int32 [C / group, R / 8], nibble
 *   order 0 2 4 6 1 3 5 7; intweight = round((w + zeros * scales) / scales) in torch's promoted dtype.
 */
int lcb_pack_rows(const void* codes, int32_t* out, int64_t rows, int64_t cols, int32_t code_bytes, int32_t bit, void* stream);
int lcb_awq_gemm_pack(const void* W, const void* scales, const int32_t* zeros, int32_t* qweight, int32_t* qzeros, int64_t R,
                      int64_t C, int32_t group_size, int32_t dtype, void* stream);

/*
 * out[0] += sum_i (T(a[i] - b[i]))^2: the difference rounded to the tensor dtype, the square and the sum in fp32 / fp64 —
 * the arithmetic of `(org_out - out).float().pow(2).mean()` (calculate_loss, awq.py:134-145; the caller divides by n) in
 * one pass over the two tensors instead of four torch kernels. a, b: n contiguous elements of `dtype`; out: one double,
 * accumulated into (zero it first).
 */
int lcb_sqdiff_sum(const void* a, const void* b, double* out, int64_t n, int32_t dtype, void* stream);

/* column absmax of a [rows, cols] matrix -> fp32 [cols] (running max: out = max(out, colmax)).
 * Replaces SmoothQuant.get_act_scale / get_weight_scale (smoothquant.py:27-51) and the |x| reductions
 * of Awq.get_act_scale (awq.py:74-85) when mean != 0 (then it is a column |x| sum). */
int lcb_col_absstat(const void* x, float* out, int64_t rows, int64_t cols, int32_t dtype, int32_t mean,
                    void* stream);

/* version / capability probe; returns LCB_ABI_VERSION (no counterpart in the reference: it has no FFI) */
int lcb_abi_version(void);
/* name of the last failing CUDA call of this thread ("" if none); no counterpart in the reference: the text the Python
 * mirror puts into the RuntimeError it raises for LCB_E_CUDA */
const char* lcb_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* LCB200_H_ */
