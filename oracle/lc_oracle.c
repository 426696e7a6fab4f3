/*
 * lc_oracle.c — CPU restatement of the reference's grouped quant / dequant arithmetic.
 *
 * TEST INFRASTRUCTURE ONLY. Nothing under lightcompress_b200/ may import, link or execute this file;
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use it, and
 * only as the checker / the timed CPU baseline.
 *
 * Parity pinning: checked bit-for-bit against outputs of the reference itself
 * (/root/reference/llmc/compression/quantization/quant.py, run on CPU in the build container by
 * tests/golden/make_golden.py) — see tests/test_oracle_golden.py. The FP8 rounding is pinned against
 * torch's Tensor.to(float8_*) (the cast that ends FloatQuantizer.real_quant_*, quant.py:1183,1211);
 * qtorch.float_quantize itself is not available offline: "parity unpinned" for qtorch's tie rule.
 *
 * Model: every torch op of the reference = fp32 compute, result rounded to the math
 * dtype M. One C statement per reference op, in the reference's order:
 *   get_minmax_range   quant.py:132-143      get_qparams  quant.py:545-559
 *   IntegerQuantizer.quant / dequant  quant.py:699-717
 *   FloatQuantizer.quant / dequant    quant.py:1061-1081
 *   real_quant_* casts  quant.py:929-940, 1183, 1211
 * Build: gcc -O2 -fopenmp -ffp-contract=off -fno-fast-math -shared -fPIC (oracle/build.py)
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

#define LCO_F32 0
#define LCO_F16 1
#define LCO_BF16 2
#define KIND_INT 0
#define KIND_E4M3 1
#define KIND_E5M2 2
#define OUT_FAKE 0
#define OUT_I8 1
#define OUT_U8 2
#define OUT_I32 3
#define OUT_FP8 4
#define OUT_NONE 5
#define OUT_CODE_T 6

typedef struct {
  int32_t dtype, math_dtype, qparam_dtype, kind, sym, round_zp, out_mode, qp_math_plus1;
  float qmin, qmax, output_scale_factor, clip_ratio;
} lco_spec;

static inline float bits2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static inline uint32_t f2bits(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

static inline uint16_t f32_to_bf16(float f) {
  uint32_t u = f2bits(f);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);  /* NaN */
  u += 0x7fffu + ((u >> 16) & 1u);                                            /* RNE */
  return (uint16_t)(u >> 16);
}
static inline float bf16_to_f32(uint16_t h) { return bits2f((uint32_t)h << 16); }
static inline uint16_t f32_to_f16(float f) { _Float16 h = (_Float16)f; uint16_t u; memcpy(&u, &h, 2); return u; }
static inline float f16_to_f32(uint16_t u) { _Float16 h; memcpy(&h, &u, 2); return (float)h; }

static inline float rd(float x, int dt) {
  if (dt == LCO_BF16) return bf16_to_f32(f32_to_bf16(x));
  if (dt == LCO_F16) return f16_to_f32(f32_to_f16(x));
  return x;
}
static inline float load(const void* p, int64_t i, int dt) {
  if (dt == LCO_BF16) return bf16_to_f32(((const uint16_t*)p)[i]);
  if (dt == LCO_F16) return f16_to_f32(((const uint16_t*)p)[i]);
  return ((const float*)p)[i];
}
static inline void store(void* p, int64_t i, float v, int dt) {
  if (dt == LCO_BF16) ((uint16_t*)p)[i] = f32_to_bf16(v);
  else if (dt == LCO_F16) ((uint16_t*)p)[i] = f32_to_f16(v);
  else ((float*)p)[i] = v;
}
/* torch.max / clamp propagate NaN */
static inline float max_nan(float a, float b) { return (a != a || b != b) ? NAN : (a > b ? a : b); }
static inline float min_nan(float a, float b) { return (a != a || b != b) ? NAN : (a < b ? a : b); }
static inline float clamp_nan(float v, float lo, float hi) { return min_nan(max_nan(v, lo), hi); }

/* fp32 -> fp8 value, round-to-nearest-even, saturate to the largest finite (e4m3fn: 448, e5m2: 57344) */
static float round_fp8(float x, int kind) {
  const int mb = kind == KIND_E4M3 ? 3 : 2;
  const int bias = kind == KIND_E4M3 ? 7 : 15;
  const float maxv = kind == KIND_E4M3 ? 448.0f : 57344.0f;
  if (x != x) return x;
  float a = fabsf(x);
  if (a == 0.0f) return x;
  int e;
  if (isinf(a)) return copysignf(maxv, x);
  frexpf(a, &e);            /* a = m * 2^e, m in [0.5,1) -> floor(log2 a) = e-1 */
  int ex = e - 1;
  if (ex < 1 - bias) ex = 1 - bias;  /* subnormal range shares the min-normal quantum */
  float quantum = ldexpf(1.0f, ex - mb);
  float q = rintf(a / quantum) * quantum;  /* exact: power-of-two scaling; rintf = half-to-even */
  if (q > maxv) q = maxv;
  return copysignf(q, x);
}
static uint8_t encode_fp8(float q, int kind) {
  const int mb = kind == KIND_E4M3 ? 3 : 2;
  const int bias = kind == KIND_E4M3 ? 7 : 15;
  uint8_t sign = signbit(q) ? 0x80 : 0;
  float a = fabsf(q);
  if (q != q) return kind == KIND_E4M3 ? (uint8_t)(sign | 0x7f) : (uint8_t)(sign | 0x7e);
  if (a == 0.0f) return sign;
  int e;
  float m = frexpf(a, &e);  /* a = m 2^e */
  int ex = e - 1;
  if (ex < 1 - bias) {
    int mant = (int)(a / ldexpf(1.0f, 1 - bias - mb));
    return (uint8_t)(sign | mant);
  }
  int mant = (int)((m * 2.0f - 1.0f) * (float)(1 << mb));
  return (uint8_t)(sign | ((ex + bias) << mb) | mant);
}

typedef struct { float s, z, s_out, clo, chi; int clipped; } unitq;

/* quant.py:545-559 */
static void qparams(float mn, float mx, const lco_spec* sp, unitq* u) {
  const int md = sp->math_dtype;
  /* per_tensor: 0-dim range / 0-dim fp32 divisor promotes to fp32 (torch 0-dim promotion) */
  const int mq = sp->qp_math_plus1 > 0 ? sp->qp_math_plus1 - 1 : md;
  u->clipped = 0;
  if (sp->clip_ratio != 1.0f && sp->clip_ratio != -1.0f) {
    /* AutoClip v1 (auto_clip.py:112-117, 139-146, 258-260): max_val = org_max * ratio, min_val = org_min * ratio
       (clip_ratio < 0 = clip_sym: org_max = |w|.amax, min_val = -max_val); cur_w = clamp(w, min_val, max_val) and the
       quantizer then sees the clamped tensor, whose range is the clamped range (clamp is monotonic). */
    const float r = fabsf(sp->clip_ratio);
    float hi, lo;
    if (sp->clip_ratio < 0.0f) { hi = rd(max_nan(fabsf(mx), fabsf(mn)) * r, md); lo = -hi; }
    else { hi = rd(mx * r, md); lo = rd(mn * r, md); }
    mx = clamp_nan(mx, lo, hi);
    mn = clamp_nan(mn, lo, hi);
    u->clo = lo; u->chi = hi; u->clipped = 1;
  }
  float s, z = 0.0f;
  if (sp->sym || sp->kind != KIND_INT) {
    float a = max_nan(fabsf(mx), fabsf(mn));          /* torch.max(max_val.abs(), min_val.abs()) */
    a = rd(max_nan(a, 1e-5f), md);                    /* .clamp(min=1e-5) */
    s = rd(a / sp->qmax, mq);                         /* abs_max / qmax */
  } else {
    float d = rd(mx - mn, md);
    d = rd(max_nan(d, 1e-5f), md);
    s = rd(d / (sp->qmax - sp->qmin), mq);
    float t = rd(mn / s, mq);
    if (sp->round_zp) z = clamp_nan(rd(sp->qmin - rintf(t), mq), sp->qmin, sp->qmax);
    else z = rd(sp->qmin - t, mq);
  }
  /* FloatQuantizer.quant: scales[scales == 0] = 1 in place (quant.py:1062); get_tensor_qparams alone keeps the 0 */
  if (sp->kind != KIND_INT && s == 0.0f && sp->out_mode != OUT_NONE) s = 1.0f;
  u->s = s; u->z = z;
  u->s_out = (sp->output_scale_factor != 1.0f) ? rd(s * sp->output_scale_factor, md) : s;
}

/* quant.py:699-717 / 1061-1081; returns fake value (math dtype), *code = code value */
static float elem(float x, const unitq* u, const lco_spec* sp, float* code) {
  const int md = sp->math_dtype;
  if (u->clipped) x = clamp_nan(x, u->clo, u->chi);
  if (sp->kind == KIND_INT) {
    float q;
    if (sp->round_zp) {
      float t1 = rd(x / u->s, md);
      float t2 = rintf(t1);
      float t3 = rd(t2 + u->z, md);
      q = clamp_nan(t3, sp->qmin, sp->qmax);
    } else {
      float sc = rd(max_nan(u->s, 1e-9f), md);
      float t1 = rd(x / sc, md);
      float t1b = rd(t1 + u->z, md);
      q = clamp_nan(rintf(t1b), sp->qmin, sp->qmax);
    }
    *code = q;
    float t4 = rd(q - u->z, md);
    return rd(t4 * u->s_out, md);
  }
  float t1 = rd(x / u->s, md);
  { volatile float zp = 0.0f; t1 = rd(t1 + zp, md); }  /* "+ zeros", zeros = tensor(0.0): -0 becomes +0 */
  float q = round_fp8(t1, sp->kind);
  *code = q;
  return q * u->s_out;
}

static void store_out(void* out, int64_t e, float val, float code, const lco_spec* sp) {
  switch (sp->out_mode) {
    case OUT_FAKE: store(out, e, val, sp->dtype); break;
    case OUT_CODE_T: store(out, e, code, sp->dtype); break;
    case OUT_I8: ((int8_t*)out)[e] = (int8_t)(int)code; break;
    case OUT_U8: ((uint8_t*)out)[e] = (uint8_t)(int)code; break;
    case OUT_I32: ((int32_t*)out)[e] = (int32_t)code; break;
    case OUT_FP8: ((uint8_t*)out)[e] = encode_fp8(code, sp->kind); break;
    default: break;
  }
}

/* same contract as lcb_fakequant (include/lcb200.h) on host memory */
int lco_fakequant(const void* in, void* out, void* scales, void* zeros, int64_t rows, int64_t cols, int64_t ur,
                  int64_t uc, const lco_spec* sp, int dynamic) {
  if (rows == 0 || cols == 0) return 0;
  if (ur > rows) ur = rows;
  if (uc > cols) uc = cols;
  if (ur == 1 && uc < cols && cols % uc != 0) return -2;
  const int64_t nru = (rows + ur - 1) / ur, ncu = (cols + uc - 1) / uc;
  const int whole = (ur == rows && uc == cols);
  const int abs_range = (ur > 1 && !whole);  /* per_block: range of |x|.float(), quant.py:137-138 */
#pragma omp parallel for schedule(static)
  for (int64_t unit = 0; unit < nru * ncu; ++unit) {
    const int64_t r0 = (unit / ncu) * ur, c0 = (unit % ncu) * uc;
    const int64_t nr = (ur < rows - r0) ? ur : rows - r0, nc = (uc < cols - c0) ? uc : cols - c0;
    unitq u;
    if (dynamic) {
      float mx = -INFINITY, mn = INFINITY;
      for (int64_t r = 0; r < nr; ++r)
        for (int64_t c = 0; c < nc; ++c) {
          float x = load(in, (r0 + r) * cols + c0 + c, sp->dtype);
          if (abs_range) x = fabsf(x);
          mx = max_nan(mx, x);
          mn = min_nan(mn, x);
        }
      if (abs_range && (nr < ur || nc < uc)) mn = min_nan(mn, 0.0f);  /* zero padding, quant.py:637-638 */
      qparams(mn, mx, sp, &u);
      if (scales) store(scales, unit, u.s_out, sp->qparam_dtype);
      if (zeros && !sp->sym && sp->kind == KIND_INT) store(zeros, unit, u.z, sp->qparam_dtype);
    } else {
      u.clipped = 0;
      u.s = load(scales, unit, sp->qparam_dtype);
      u.z = (zeros && !sp->sym && sp->kind == KIND_INT) ? load(zeros, unit, sp->qparam_dtype) : 0.0f;
      if (sp->kind != KIND_INT && u.s == 0.0f) u.s = 1.0f;
      u.s_out = (sp->output_scale_factor != 1.0f) ? rd(u.s * sp->output_scale_factor, sp->math_dtype) : u.s;
    }
    if (sp->out_mode == OUT_NONE) continue;
    for (int64_t r = 0; r < nr; ++r)
      for (int64_t c = 0; c < nc; ++c) {
        const int64_t e = (r0 + r) * cols + c0 + c;
        float code;
        float val = elem(load(in, e, sp->dtype), &u, sp, &code);
        store_out(out, e, val, code, sp);
      }
  }
  return 0;
}

/* ---- identities the CUDA fast chain relies on (tests/test_fq_math_identity.py) --------------------- */
/* Markstein division with a correctly rounded reciprocal: returns number of mismatches vs x / s */
int64_t lco_check_markstein(const float* xs, int64_t nx, const float* ss, int64_t ns) {
  int64_t bad = 0;
#pragma omp parallel for reduction(+ : bad)
  for (int64_t j = 0; j < ns; ++j) {
    const float s = ss[j];
    const float r = 1.0f / s; /* IEEE: correctly rounded reciprocal == __frcp_rn */
    for (int64_t i = 0; i < nx; ++i) {
      const float x = xs[i];
      float q0 = x * r;
      float e = fmaf(-s, q0, x);
      float q = fmaf(e, r, q0);
      float ref = x / s;
      /* quotients below 2^-40 round to zero in every output format (int codes, fp8): a last-bit difference caused by
         an underflowing residual there cannot reach the output, so it is not counted */
      if (f2bits(q) != f2bits(ref) && !(q != q && ref != ref) && (fabsf(q) >= 0x1p-40f || fabsf(ref) >= 0x1p-40f)) ++bad;
    }
  }
  return bad;
}
/* magic-number rounding == rintf on |q| < 2^22 */
int64_t lco_check_magic_round(const float* qs, int64_t n) {
  int64_t bad = 0;
  for (int64_t i = 0; i < n; ++i) {
    volatile float t = qs[i] + 12582912.0f;
    float r = t - 12582912.0f;
    if (r != rintf(qs[i])) ++bad;
  }
  return bad;
}
float lco_round_fp8(float x, int kind) { return round_fp8(x, kind); }
int lco_encode_fp8(float x, int kind) { return encode_fp8(x, kind); }
