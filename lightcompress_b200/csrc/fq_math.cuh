// fq_math.cuh — the arithmetic of the grouped quant-dequant path.
//
// Two implementations of the same function:
//   * the LITERAL chain: one statement per torch op of the reference
//     (quant.py:545-559 get_qparams, :699-717 quant/dequant, :1061-1081 float quant), every
//     intermediate rounded to the math dtype — this is the specification;
//   * the FAST chain used by the vector kernels: algebraically identical for bit<=8, round_zp=True,
//     finite inputs. It replaces the IEEE division by a Markstein step on a per-unit correctly
//     rounded reciprocal (q0 = x*r; e = fma(-s,q0,x); q = fma(e,r,q0) == RN(x/s)), torch.round by the
//     1.5*2^23 magic add, and folds "+z, clamp, -z" into clamp(t, qmin-z, qmax-z) (all three are
//     exact on integers <= 256). tests/test_fq_math_identity.py proves the identities on the CPU.
#pragma once
#include "common.cuh"

namespace lcb {

struct UnitQ {
  float s;      // scale, rounded to math dtype
  float r;      // RN(1/s)
  float z;      // zero point
  float lo, hi; // qmin - z, qmax - z
  float s_out;  // s * output_scale_factor (rounded)
  int exact;    // 1 -> this unit must take the literal chain
  float clo = -INFINITY, chi = INFINITY;  // AutoClip v1 clamp bounds of this unit (fq_elem_literal clamps x when clipped)
  int clipped = 0;
};

// AutoClip v1 (auto_clip.py:112-117, 139-146, 258-260): max_val = org_max * ratio, min_val = org_min * ratio, or with
// clip_sym (clip < 0) org_max = |w|.amax and min_val = -max_val; cur_w = clamp(w, min_val, max_val) is what the quantizer
// sees, so its range is the clamped range (clamp is monotonic). Returns false when the spec does not clip.
__device__ __forceinline__ bool clip_bounds(float& mn, float& mx, float clip, int md, float& lo, float& hi);

struct SpecD {  // device copy of lcb_quant_spec + derived flags
  int dtype, math_dtype, qparam_dtype, kind, sym, round_zp, out_mode, fast_ok, qp_math;
  float qmin, qmax, osf, clip;
};

__device__ __forceinline__ float clamp_nan(float v, float lo, float hi) {
  return min_nan(max_nan(v, lo), hi);
}

__device__ __forceinline__ bool finite_f(float v) { return fabsf(v) <= 3.402823466e38f; }

__device__ __forceinline__ bool clip_bounds(float& mn, float& mx, float clip, int md, float& lo, float& hi) {
  if (clip == 1.0f || clip == -1.0f) return false;
  const float r = fabsf(clip);
  if (clip < 0.f) {
    hi = rd_rt(__fmul_rn(max_nan(fabsf(mx), fabsf(mn)), r), md);
    lo = -hi;
  } else {
    hi = rd_rt(__fmul_rn(mx, r), md);
    lo = rd_rt(__fmul_rn(mn, r), md);
  }
  mx = clamp_nan(mx, lo, hi);
  mn = clamp_nan(mn, lo, hi);
  return true;
}

// ---- qparams from a (min, max) range: quant.py:545-559 -------------------------------------------
__device__ __forceinline__ void qparams_from_range(float mn, float mx, const SpecD& sp, UnitQ& u) {
  const int md = sp.math_dtype;
  const int mq = sp.qp_math;  // dtype of the qparam divisions (== md except per_tensor 0-dim promotion)
  u.clipped = clip_bounds(mn, mx, sp.clip, md, u.clo, u.chi) ? 1 : 0;
  float s, z = 0.f;
  if (sp.sym || sp.kind != LCB_KIND_INT) {
    float a = max_nan(fabsf(mx), fabsf(mn));  // both already in math dtype: abs/max exact
    a = rd_rt(max_nan(a, 1e-5f), md);
    s = rd_rt(__fdiv_rn(a, sp.qmax), mq);
  } else {
    float d = rd_rt(__fsub_rn(mx, mn), md);
    d = rd_rt(max_nan(d, 1e-5f), md);
    s = rd_rt(__fdiv_rn(d, __fsub_rn(sp.qmax, sp.qmin)), mq);
    float t = rd_rt(__fdiv_rn(mn, s), mq);
    if (sp.round_zp) {
      z = clamp_nan(rd_rt(__fsub_rn(sp.qmin, rintf(t)), mq), sp.qmin, sp.qmax);
    } else {
      z = rd_rt(__fsub_rn(sp.qmin, t), mq);
    }
  }
  u.s = s;
  u.z = z;
}

// derive the per-unit fast-path constants once s and z are known
__device__ __forceinline__ void finish_unit(const SpecD& sp, UnitQ& u) {
  // FloatQuantizer.quant mutates scales in place: scales[scales == 0] = 1 (quant.py:1062); get_tensor_qparams alone
  // (LCB_OUT_NONE) returns the untouched zero
  if (sp.kind != LCB_KIND_INT && u.s == 0.f && sp.out_mode != LCB_OUT_NONE) u.s = 1.f;
  u.s_out = (sp.osf != 1.0f) ? rd_rt(__fmul_rn(u.s, sp.osf), sp.math_dtype) : u.s;
  u.r = __frcp_rn(u.s);
  u.lo = __fsub_rn(sp.qmin, u.z);
  u.hi = __fsub_rn(sp.qmax, u.z);
  // fast chain needs: spec eligible, s a positive normal number with a finite reciprocal, integral z
  // 1e-20 <= s <= 1e20: for |x| >= 2^-100 the Markstein residual e = x - s*q0 is exactly representable; for smaller
  // |x| the quotient is < 1e-10 and rounds to zero on every path (tests/test_fq_math_identity.py)
  bool ok = sp.fast_ok && (u.s >= 1e-20f) && (u.s <= 1e20f) && finite_f(u.z) && (u.z == rintf(u.z));
  u.exact = ok ? 0 : 1;
}

// ---- LITERAL element chain -----------------------------------------------------------------------
// returns (.x) the fake-quant value, already rounded to the math dtype, and (.y) the integer / fp8 code value
static __device__ __noinline__ float2 fq_elem_literal(float x, const UnitQ& u, const SpecD& sp) {
  const int md = sp.math_dtype;
  if (u.clipped) x = clamp_nan(x, u.clo, u.chi);
  if (sp.out_mode == LCB_OUT_DEQUANT) {   // x is a code: (q - z) * s   (quant.py:710-712 / 1076-1078)
    const float t4 = sp.kind == LCB_KIND_INT ? rd_rt(__fsub_rn(x, u.z), md) : x;
    return make_float2(rd_rt(__fmul_rn(t4, u.s_out), md), x);
  }
  if (sp.kind == LCB_KIND_INT) {
    float q;
    if (sp.round_zp) {
      float t1 = rd_rt(__fdiv_rn(x, u.s), md);
      float t2 = rintf(t1);
      float t3 = rd_rt(__fadd_rn(t2, u.z), md);
      q = clamp_nan(t3, sp.qmin, sp.qmax);
    } else {
      float sc = rd_rt(max_nan(u.s, 1e-9f), md);
      float t1 = rd_rt(__fdiv_rn(x, sc), md);
      float t1b = rd_rt(__fadd_rn(t1, u.z), md);
      q = clamp_nan(rintf(t1b), sp.qmin, sp.qmax);
    }
    float t4 = rd_rt(__fsub_rn(q, u.z), md);
    return make_float2(rd_rt(__fmul_rn(t4, u.s_out), md), q);
  } else {
    float t1 = rd_rt(__fadd_rn(rd_rt(__fdiv_rn(x, u.s), md), 0.0f), md);  // "+ zeros" with zeros = tensor(0.0): -0 -> +0
    uint8_t b = f32_to_fp8_byte(t1, sp.kind);
    float q = fp8_byte_to_f32(b, sp.kind);
    if (t1 != t1) q = t1;
    return make_float2(__fmul_rn(q, u.s_out), q);  // fp32 product, rounded to the storage dtype by the store
  }
}

// ---- FAST element chain --------------------------------------------------------------------------
template <typename M>  // M: math dtype as a type (TI<M>::rd)
__device__ __forceinline__ float fq_elem_fast_int(float x, const UnitQ& u, float* tq) {
  float q0 = __fmul_rn(x, u.r);
  float e = __fmaf_rn(-u.s, q0, x);
  float q = __fmaf_rn(e, u.r, q0);  // == RN(x / s)
  q = TI<M>::rd(q);
  const float MAGIC = 12582912.0f;  // 1.5 * 2^23: RNE to integer for |q| < 2^22
  float t = __fsub_rn(__fadd_rn(q, MAGIC), MAGIC);
  t = fminf(fmaxf(t, u.lo), u.hi);
  *tq = t;  // code = t + z
  return __fmul_rn(t, u.s_out);
}

template <typename M>
__device__ __forceinline__ float fq_elem_fast_fp8(float x, const UnitQ& u, int kind, float* code) {
  float q0 = __fmul_rn(x, u.r);
  float e = __fmaf_rn(-u.s, q0, x);
  float q = __fmaf_rn(e, u.r, q0);
  q = __fadd_rn(TI<M>::rd(q), 0.0f);  // "+ zeros": turns -0 into +0 like the reference
  float v = fp8_byte_to_f32(f32_to_fp8_byte(q, kind), kind);
  *code = v;
  return __fmul_rn(v, u.s_out);
}

}  // namespace lcb
