// Kernel (3), difference-propagating form ("jets"): the FD-SSM objective of the energy MLPs evaluated on the three
// streams (h0, a, b) with  h(x+v) = h0 + b + a,  h(x-v) = h0 + b - a  instead of on the stacked batch [x | x+v | x-v].
//
//   layer       z0 = W h0 + bias        P = W a          Q = W b            (ONE GEMM over the rows [h0 | a | b])
//   activation  h0' = s(z0)     a' = (s(z0+Q+P) - s(z0+Q-P))/2     b' = (s(z0+Q+P) + s(z0+Q-P))/2 - s(z0)
//   start       h0 = x          a  = v = eps u/||u||               b  = 0
//   head        dE = E(x+v) - E(x-v),  sE = E(x+v) + E(x-v) - 2 E(x)      loss_b = dE^2/8 - sE + (c1 - 2) E(x)
//
// Mathematically identical to the reference's stacked evaluation (DeepEBM/ESM/loss_train.py:212-234,
// SSM/losses/efficient_sm.py:47-73); numerically, rounding is relative to the DIFFERENCES (a ~ eps, b ~ eps^2) instead of
// to |E|, so TF32 tensor-core operands perturb loss and gradient at the 1e-4 level where the reference's own fp32 stacked
// arithmetic is off by 1e-2 .. 1 (hard parts 1-2; oracle/jet.py is the CPU restatement and
// tests/tools/jet_accuracy_study.py the measurement).  a', b' are evaluated in cancellation-free closed forms (below).
//
// Arithmetic: activations and gradients travel as TF32-rounded (round-to-nearest) fp32; weights are split W = hi + lo
// (both TF32) once per step and every forward / dgrad k-step issues A*hi + A*lo (PASSES = 2; PASSES = 1 drops the lo term:
// 2e-4 instead of 6e-5 gradient error at config 2); the weight-gradient GEMMs contract two rounded operands (1 pass).
#pragma once
#include "tc_gemm_host.cuh"

namespace fdsm {
namespace jet {

constexpr float LOG2E = 1.4426950408889634f, LN2 = 0.69314718055994531f;
#ifndef FDSM_JET_PACKED
#define FDSM_JET_PACKED 1      // 0: scalar closed forms everywhere (A/B builds: make EXTRA=-DFDSM_JET_PACKED=0)
#endif

__device__ __forceinline__ float ex2a(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float lg2a(float x) {
  float y;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// 1/d: the MUFU reciprocal (relative error <= 2^-23)
__device__ __forceinline__ float rcp_a(float d) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(d));
  return r;
}
// e^x - 1 for |x| <= SMALL: degree-7 Taylor (relative error < 1e-8 on that range); no MUFU
constexpr float SMALL = 0.25f;
__device__ __forceinline__ float expm1_s(float x) {
  float p = 1.984127e-4f;
  p = fmaf(p, x, 1.3888889e-3f);
  p = fmaf(p, x, 8.3333333e-3f);
  p = fmaf(p, x, 4.1666667e-2f);
  p = fmaf(p, x, 1.6666667e-1f);
  p = fmaf(p, x, 0.5f);
  return fmaf(p * x, x, x);
}
// e^x - 1, any x: the polynomial where it holds, MUFU exponential minus one beyond (< 1e-6 there)
__device__ __forceinline__ float expm1_j(float x) {
  return fabsf(x) <= 0.35f ? expm1_s(x) : ex2a(x * LOG2E) - 1.f;
}
// log(1 + n/d) = 2 atanh(n / (2d + n)) as an odd polynomial: relative error < 1e-7 for n/d in [-0.55, 1.1] (|s| <= 0.36)
__device__ __forceinline__ float log1p_ratio(float n, float d) {
  const float s = n * rcp_a(fmaf(2.f, d, n));
  const float u = s * s;
  float q = 0.076923077f;
  q = fmaf(q, u, 0.090909091f);
  q = fmaf(q, u, 0.11111111f);
  q = fmaf(q, u, 0.14285715f);
  q = fmaf(q, u, 0.2f);
  q = fmaf(q, u, 0.33333334f);
  const float s2 = s + s;
  return fmaf(s2 * u, q, s2);
}
// log(1 + t), any t > -1
__device__ __forceinline__ float log1p_j(float t) {
  return (t >= -0.5f && t <= 1.f) ? log1p_ratio(t, 1.f) : lg2a(1.f + t) * LN2;
}
__device__ __forceinline__ float tf32r(float x) { return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xFFFFE000u); }

struct Jet { float h0, a, b; };

// ---- two elements per instruction: Blackwell's packed fp32 pipe (FFMA2 / FADD2 / FMUL2: fma.rn.f32x2 ...) -----------------------
// The epilogues are bound by instruction issue and already work on two samples per lane; the polynomial parts of the closed
// forms below (about two thirds of the arithmetic) run on both samples with one instruction each.  Every lane of a packed
// instruction is the IEEE operation of its scalar form.
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) {
  float2 d;
  asm("{\n\t.reg .b64 ra, rb, rc, rd;\n\t"
      "mov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %5};\n\tmov.b64 rc, {%6, %7};\n\t"
      "fma.rn.f32x2 rd, ra, rb, rc;\n\t"
      "mov.b64 {%0, %1}, rd;\n\t}"
      : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y), "f"(c.x), "f"(c.y));
  return d;
}
__device__ __forceinline__ float2 mul2(float2 a, float2 b) {
  float2 d;
  asm("{\n\t.reg .b64 ra, rb, rd;\n\t"
      "mov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %5};\n\t"
      "mul.rn.f32x2 rd, ra, rb;\n\t"
      "mov.b64 {%0, %1}, rd;\n\t}"
      : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
  return d;
}
__device__ __forceinline__ float2 add2(float2 a, float2 b) {
  float2 d;
  asm("{\n\t.reg .b64 ra, rb, rd;\n\t"
      "mov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %5};\n\t"
      "add.rn.f32x2 rd, ra, rb;\n\t"
      "mov.b64 {%0, %1}, rd;\n\t}"
      : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
  return d;
}
__device__ __forceinline__ float2 f2(float c) { return make_float2(c, c); }
__device__ __forceinline__ float2 neg2(float2 a) { return make_float2(-a.x, -a.y); }
__device__ __forceinline__ float2 sub2(float2 a, float2 b) { return add2(a, neg2(b)); }
__device__ __forceinline__ float2 rcp2(float2 d) { return make_float2(rcp_a(d.x), rcp_a(d.y)); }
// the polynomials of expm1_s / log1p_ratio, two arguments at once (same coefficients, same evaluation order)
__device__ __forceinline__ float2 expm1_s2(float2 x) {
  float2 p = f2(1.984127e-4f);
  p = fma2(p, x, f2(1.3888889e-3f));
  p = fma2(p, x, f2(8.3333333e-3f));
  p = fma2(p, x, f2(4.1666667e-2f));
  p = fma2(p, x, f2(1.6666667e-1f));
  p = fma2(p, x, f2(0.5f));
  return fma2(mul2(p, x), x, x);
}
__device__ __forceinline__ float2 log1p_ratio2(float2 n, float2 d) {
  const float2 s = mul2(n, rcp2(fma2(f2(2.f), d, n)));
  const float2 u = mul2(s, s);
  float2 q = f2(0.076923077f);
  q = fma2(q, u, f2(0.090909091f));
  q = fma2(q, u, f2(0.11111111f));
  q = fma2(q, u, f2(0.14285715f));
  q = fma2(q, u, f2(0.2f));
  q = fma2(q, u, f2(0.33333334f));
  const float2 s2 = add2(s, s);
  return fma2(mul2(s2, u), q, s2);
}
struct Hyp2 { float2 em, emn, sinh, coshm1; };
__device__ __forceinline__ Hyp2 hyp_small2(float2 d) {
  Hyp2 r;
  r.em = expm1_s2(d);
  const float2 inv = rcp2(add2(f2(1.f), r.em));     // e^-d
  r.emn = neg2(mul2(r.em, inv));
  r.sinh = mul2(f2(0.5f), sub2(r.em, r.emn));
  r.coshm1 = mul2(mul2(mul2(f2(0.5f), r.em), r.em), inv);
  return r;
}

// quantities of one small argument d shared by the closed forms: expm1(d), expm1(-d), sinh d, cosh d - 1
struct Hyp { float em, emn, sinh, coshm1; };
template <bool SMALL_ARG>
__device__ __forceinline__ Hyp hyp(float d) {
  Hyp r;
  r.em = SMALL_ARG ? expm1_s(d) : expm1_j(d);
  const float inv = rcp_a(1.f + r.em);     // e^-d
  r.emn = -r.em * inv;
  r.sinh = 0.5f * (r.em - r.emn);          // opposite signs: no cancellation
  r.coshm1 = 0.5f * r.em * r.em * inv;     // (e^d - 1)^2 / (2 e^d)
  return r;
}

enum : int { JACT_SOFTPLUS = 0, JACT_ELU = 1 };

// ---- softplus (offset 0) / normalised softplus (offset ln 2: DeepEBM/ESM/models.py:19-28) ------------------------------
// forward, s0 = sigmoid(z0), e0 = 1 - s0:
//   (1 + e^{z0+d}) / (1 + e^{z0}) = 1 + s0 expm1(d)
//   a = 1/2 log1p( 2 s0 e^Q sinh P / (1 + s0 expm1(Q - P)) )
//   b = 1/2 log1p( s0 e0 S + s0^2 expm1(2Q) ),   S = expm1(Q+P) + expm1(Q-P) = 2 (expm1(Q) cosh P + (cosh P - 1))
__device__ __forceinline__ float softplus_j(float z) {
  return fmaxf(z, 0.f) + log1p_j(ex2a(-fabsf(z) * LOG2E));
}
// A perturbation response beyond SMALL (eps far above the reference's 0.01 .. 0.1, or exploding weights) leaves the
// regime the closed forms are conditioned for: the three softplus values are O(1) apart, combine them directly.
// Out of line: rare, and the epilogue loops must stay small enough for the instruction cache.
__device__ __noinline__ void softplus_jet_fwd_large(float z0, float P, float Q, float offset, Jet* out) {
  const float s0 = softplus_j(z0), sp = softplus_j(z0 + Q + P), sm = softplus_j(z0 + Q - P);
  out->h0 = s0 - offset;
  out->a = 0.5f * (sp - sm);
  out->b = 0.5f * ((sp - s0) + (sm - s0));
}
// main path: |P|, |Q| <= SMALL, every expm1 / log1p below is a short polynomial (6 MUFU operations in all)
__device__ __forceinline__ Jet softplus_jet_fwd_small(float z0, float P, float Q, float offset) {
  const float t = ex2a(-fabsf(z0) * LOG2E);          // e^-|z0|
  const float r = rcp_a(1.f + t);
  const float s0 = z0 >= 0.f ? r : t * r, e0 = z0 >= 0.f ? t * r : r;
  Jet o;
  o.h0 = fmaf(lg2a(1.f + t), LN2, fmaxf(z0, 0.f)) - offset;
  const Hyp hp = hyp<true>(P);
  const float emq = expm1_s(Q), eq = 1.f + emq;
  const float em = hp.emn + emq + hp.emn * emq;      // expm1(Q - P)
  o.a = 0.5f * log1p_ratio(2.f * s0 * eq * hp.sinh, fmaf(s0, em, 1.f));
  const float S = 2.f * fmaf(emq, 1.f + hp.coshm1, hp.coshm1);
  o.b = 0.5f * log1p_ratio(s0 * fmaf(e0, S, s0 * emq * (2.f + emq)), 1.f);
  return o;
}
__device__ __forceinline__ Jet softplus_jet_fwd(float z0, float P, float Q, float offset) {
  if (fmaxf(fabsf(P), fabsf(Q)) > SMALL) {
    Jet o;
    softplus_jet_fwd_large(z0, P, Q, offset, &o);
    return o;
  }
  return softplus_jet_fwd_small(z0, P, Q, offset);
}
// backward from the OUTPUTS: softplus(z+-) = sp0 + b +- a and sigmoid = 1 - exp(-softplus), e0 = exp(-sp0):
//   S1 = (s'(z+) - s'(z-))/2 = e0 e^-b sinh a     S2 = (s'(z+) + s'(z-))/2 - s'(z0) = -e0 (expm1(-b) cosh a + cosh a - 1)
//   dz0 = s0 g0 + S2 gb + S1 ga      dP = Sm ga + S1 gb      dQ = Sm gb + S1 ga       (Sm = S2 + s0)
__device__ __forceinline__ void jet_bwd_apply(float d0, float S1, float S2, float g0, float ga, float gb, float& dz0, float& dP,
                                              float& dQ) {
  const float Sm = S2 + d0;
  dz0 = fmaf(d0, g0, fmaf(S2, gb, S1 * ga));
  dP = fmaf(Sm, ga, S1 * gb);
  dQ = fmaf(Sm, gb, S1 * ga);
}
__device__ __noinline__ void softplus_jet_bwd_large(float sp0, float e0, float a, float b, float* s0, float* S1, float* S2) {
  *s0 = -expm1_j(-sp0);
  const Hyp ha = hyp<false>(a);
  const float emb = expm1_j(-b);
  *S1 = e0 * (1.f + emb) * ha.sinh;
  *S2 = -e0 * fmaf(emb, 1.f + ha.coshm1, ha.coshm1);
}
__device__ __forceinline__ void softplus_jet_bwd_small(const Jet& o, float offset, float g0, float ga, float gb, float& dz0,
                                                       float& dP, float& dQ) {
  const float e0 = ex2a(-(o.h0 + offset) * LOG2E);
  const Hyp ha = hyp<true>(o.a);
  const float emb = expm1_s(-o.b);
  jet_bwd_apply(1.f - e0, e0 * (1.f + emb) * ha.sinh, -e0 * fmaf(emb, 1.f + ha.coshm1, ha.coshm1), g0, ga, gb, dz0, dP, dQ);
}
__device__ __forceinline__ void softplus_jet_bwd(const Jet& o, float offset, float g0, float ga, float gb, float& dz0, float& dP,
                                                 float& dQ) {
  if (fmaxf(fabsf(o.a), fabsf(o.b)) <= SMALL) {       // always, when the forward took its main path
    softplus_jet_bwd_small(o, offset, g0, ga, gb, dz0, dP, dQ);
  } else {
    const float sp0 = o.h0 + offset;
    float s0, S1, S2;
    softplus_jet_bwd_large(sp0, ex2a(-sp0 * LOG2E), o.a, o.b, &s0, &S1, &S2);
    jet_bwd_apply(s0, S1, S2, g0, ga, gb, dz0, dP, dQ);
  }
}

// the two main paths above on TWO elements (sample pair of a lane) with the packed pipe
__device__ __forceinline__ void softplus_jet_fwd_small2(float2 z0, float2 P, float2 Q, float offset, Jet (&o)[2]) {
  const float2 t = make_float2(ex2a(-fabsf(z0.x) * LOG2E), ex2a(-fabsf(z0.y) * LOG2E));      // e^-|z0|
  const float2 t1 = add2(f2(1.f), t);
  const float2 r = rcp2(t1), tr = mul2(t, r);
  const float2 s0 = make_float2(z0.x >= 0.f ? r.x : tr.x, z0.y >= 0.f ? r.y : tr.y);
  const float2 e0 = make_float2(z0.x >= 0.f ? tr.x : r.x, z0.y >= 0.f ? tr.y : r.y);
  const float2 h0 = add2(fma2(make_float2(lg2a(t1.x), lg2a(t1.y)), f2(LN2), make_float2(fmaxf(z0.x, 0.f), fmaxf(z0.y, 0.f))), f2(-offset));
  const Hyp2 hp = hyp_small2(P);
  const float2 emq = expm1_s2(Q), eq = add2(f2(1.f), emq);
  const float2 em = add2(add2(hp.emn, emq), mul2(hp.emn, emq));      // expm1(Q - P)
  const float2 a = mul2(f2(0.5f), log1p_ratio2(mul2(mul2(mul2(f2(2.f), s0), eq), hp.sinh), fma2(s0, em, f2(1.f))));
  const float2 S = mul2(f2(2.f), fma2(emq, add2(f2(1.f), hp.coshm1), hp.coshm1));
  const float2 b = mul2(f2(0.5f), log1p_ratio2(mul2(s0, fma2(e0, S, mul2(mul2(s0, emq), add2(f2(2.f), emq)))), f2(1.f)));
  o[0] = Jet{h0.x, a.x, b.x};
  o[1] = Jet{h0.y, a.y, b.y};
}
__device__ __forceinline__ void softplus_jet_bwd_small2(const Jet (&o)[2], float offset, float2 g0, float2 ga, float2 gb, float2& dz0,
                                                        float2& dP, float2& dQ) {
  const float2 e0 = make_float2(ex2a(-(o[0].h0 + offset) * LOG2E), ex2a(-(o[1].h0 + offset) * LOG2E));
  const Hyp2 ha = hyp_small2(make_float2(o[0].a, o[1].a));
  const float2 emb = expm1_s2(make_float2(-o[0].b, -o[1].b));
  const float2 d0 = sub2(f2(1.f), e0);
  const float2 S1 = mul2(mul2(e0, add2(f2(1.f), emb)), ha.sinh);
  const float2 S2 = neg2(mul2(e0, fma2(emb, add2(f2(1.f), ha.coshm1), ha.coshm1)));
  const float2 Sm = add2(S2, d0);
  dz0 = fma2(d0, g0, fma2(S2, gb, mul2(S1, ga)));
  dP = fma2(Sm, ga, mul2(S1, gb));
  dQ = fma2(Sm, gb, mul2(S1, ga));
}

// ---- ELU (alpha = 1) --------------------------------------------------------------------------------------------------------
// all three arguments in the linear region: a = P, b = Q exactly; all three in the exponential region:
// a = e^{z0} e^Q sinh P, b = e^{z0} (expm1(Q) cosh P + cosh P - 1); otherwise the one-sided differences
// elu(z0 + d) - elu(z0) (each cancellation-free) are combined directly.
__device__ __noinline__ void elu_jet_fwd_large(float z0, float zp, float zm, Jet* out) {
  const float e0 = z0 > 0.f ? z0 : expm1_j(z0), ep = zp > 0.f ? zp : expm1_j(zp), em = zm > 0.f ? zm : expm1_j(zm);
  out->h0 = e0;
  out->a = 0.5f * (ep - em);
  out->b = 0.5f * ((ep - e0) + (em - e0));
}
__device__ __forceinline__ Jet elu_jet_fwd_small(float z0, float P, float Q) {
  const float zp = z0 + Q + P, zm = z0 + Q - P;
  const float em0 = expm1_j(fminf(z0, 0.f));          // elu(z0) where z0 <= 0
  const float e = 1.f + em0;                          // e^{min(z0, 0)}
  Jet o;
  o.h0 = z0 > 0.f ? z0 : em0;
  const Hyp hp = hyp<true>(P);
  const float emq = expm1_s(Q), eq = 1.f + emq;
  const float ep = hp.em + emq + hp.em * emq, em = hp.emn + emq + hp.emn * emq;   // expm1(Q +- P)
  const bool pos = z0 > 0.f && zp > 0.f && zm > 0.f, neg = z0 <= 0.f && zp <= 0.f && zm <= 0.f;
  // one-sided differences for the mixed case
  auto delta = [&](float d, float t, float emd) {
    // elu(z0 + d) - elu(z0), t = z0 + d, emd = expm1(d)
    if (z0 > 0.f) return t > 0.f ? d : expm1_j(fminf(t, 0.f)) - z0;
    return t > 0.f ? t - em0 : e * emd;
  };
  const float dp = delta(Q + P, zp, ep), dm = delta(Q - P, zm, em);
  o.a = pos ? P : (neg ? e * eq * hp.sinh : 0.5f * (dp - dm));
  o.b = pos ? Q : (neg ? e * fmaf(emq, 1.f + hp.coshm1, hp.coshm1) : 0.5f * (dp + dm));
  return o;
}
__device__ __forceinline__ Jet elu_jet_fwd(float z0, float P, float Q) {
  if (fmaxf(fabsf(P), fabsf(Q)) > SMALL) {            // see softplus_jet_fwd_large
    Jet o;
    elu_jet_fwd_large(z0, z0 + Q + P, z0 + Q - P, &o);
    return o;
  }
  return elu_jet_fwd_small(z0, P, Q);
}
__device__ __forceinline__ void elu_jet_bwd(const Jet& o, float g0, float ga, float gb, float& dz0, float& dP, float& dQ) {
  const float hp = o.h0 + o.b + o.a, hm = o.h0 + o.b - o.a;
  const float d0 = o.h0 > 0.f ? 1.f : o.h0 + 1.f;
  const float dp = hp > 0.f ? 1.f : hp + 1.f, dm = hm > 0.f ? 1.f : hm + 1.f;
  const bool neg = o.h0 <= 0.f && hp <= 0.f && hm <= 0.f;
  const float S1 = neg ? o.a : 0.5f * (dp - dm);
  const float S2 = neg ? o.b : 0.5f * (dp + dm) - d0;
  jet_bwd_apply(d0, S1, S2, g0, ga, gb, dz0, dP, dQ);
}

template <int ACT>
__device__ __forceinline__ Jet jet_fwd(float z0, float P, float Q, float offset) {
  if constexpr (ACT == JACT_ELU) return elu_jet_fwd(z0, P, Q);
  else return softplus_jet_fwd(z0, P, Q, offset);
}
template <int ACT>
__device__ __forceinline__ void jet_bwd(const Jet& o, float offset, float g0, float ga, float gb, float& dz0, float& dP, float& dQ) {
  if constexpr (ACT == JACT_ELU) elu_jet_bwd(o, g0, ga, gb, dz0, dP, dQ);
  else softplus_jet_bwd(o, offset, g0, ga, gb, dz0, dP, dQ);
}

// out-of-line single-element versions for the rare mixed case below (keeps the epilogue loops small)
template <int ACT>
__device__ __noinline__ void jet_fwd_any(float z0, float P, float Q, float offset, Jet* o) { *o = jet_fwd<ACT>(z0, P, Q, offset); }
__device__ __noinline__ void softplus_jet_bwd_any(const Jet* o, float offset, float g0, float ga, float gb, float* d) {
  softplus_jet_bwd(*o, offset, g0, ga, gb, d[0], d[1], d[2]);
}

// ---- exact sliced score matching: the same three streams as an order-2 Taylor jet of tau -> h(x + tau u) --------------------
// (h0, a, b) = (h, h', h''/2):  z0 = W h0 + bias, P = W a, Q = W b as before; the activation rule becomes
//     a' = s1 P          b' = 1/2 s2 P^2 + s1 Q          (s1, s2, s3: first three derivatives of the activation at z0)
// (the eps -> 0 limit of the finite-difference rule above), and the backward from the OUTPUTS needs no division:
//     s2 P = (1 - s0) a' ,   1/2 s3 P^2 + s2 Q = (1 - s0)(b' - a'^2 / 2)         (softplus: s1 = s0 = sigmoid(z0))
//     dz0 = s1 g0 + (s2 P) ga + (1/2 s3 P^2 + s2 Q) gb      dP = s1 ga + (s2 P) gb      dQ = s1 gb
// Reference: SSM/losses/sliced_sm.py:7-34, DeepEBM/ESM/loss_train.py:173-189 (double backward); oracle/forward_mode.py.
enum : int { JMODE_FD = 0, JMODE_TAYLOR = 1 };
template <int ACT>
__device__ __forceinline__ Jet taylor_fwd(float z0, float P, float Q, float offset) {
  Jet o;
  if constexpr (ACT == JACT_ELU) {
    const float e = ex2a(fminf(z0, 0.f) * LOG2E);
    const bool pos = z0 > 0.f;
    o.h0 = pos ? z0 : expm1_j(z0);
    o.a = pos ? P : e * P;
    o.b = pos ? Q : e * fmaf(0.5f * P, P, Q);
  } else {
    const float t = ex2a(-fabsf(z0) * LOG2E);
    const float r = rcp_a(1.f + t);
    const float s0 = z0 >= 0.f ? r : t * r, e0 = z0 >= 0.f ? t * r : r;
    o.h0 = fmaf(lg2a(1.f + t), LN2, fmaxf(z0, 0.f)) - offset;
    o.a = s0 * P;
    o.b = s0 * fmaf(0.5f * e0 * P, P, Q);
  }
  return o;
}
template <int ACT>
__device__ __forceinline__ void taylor_bwd(const Jet& o, float offset, float g0, float ga, float gb, float& dz0, float& dP, float& dQ) {
  float d1, S1, S2;
  if constexpr (ACT == JACT_ELU) {
    const bool pos = o.h0 > 0.f;
    d1 = pos ? 1.f : o.h0 + 1.f;
    S1 = pos ? 0.f : o.a;
    S2 = pos ? 0.f : o.b;
  } else {
    const float e0 = ex2a(-(o.h0 + offset) * LOG2E);
    d1 = 1.f - e0;
    S1 = e0 * o.a;
    S2 = e0 * fmaf(-0.5f * o.a, o.a, o.b);
  }
  dz0 = fmaf(d1, g0, fmaf(S1, ga, S2 * gb));
  dP = fmaf(d1, ga, S1 * gb);
  dQ = d1 * gb;
}

// N independent elements at once: one test for the main path, then N straight-line evaluations the compiler can interleave
// (the epilogues are latency-bound on a single element's dependency chain)
template <int ACT, int MODE, int N>
__device__ __forceinline__ void jet_fwd_n(const float (&z0)[N], const float (&P)[N], const float (&Q)[N], float offset, Jet (&o)[N]) {
  if constexpr (MODE == JMODE_TAYLOR) {
#pragma unroll
    for (int n = 0; n < N; ++n) o[n] = taylor_fwd<ACT>(z0[n], P[n], Q[n], offset);
    return;
  }
  float m = 0.f;
#pragma unroll
  for (int n = 0; n < N; ++n) m = fmaxf(m, fmaxf(fabsf(P[n]), fabsf(Q[n])));
  if (m <= SMALL) {
    if constexpr (ACT != JACT_ELU && N == 2 && FDSM_JET_PACKED) {
      softplus_jet_fwd_small2(make_float2(z0[0], z0[1]), make_float2(P[0], P[1]), make_float2(Q[0], Q[1]), offset, o);
    } else {
#pragma unroll
      for (int n = 0; n < N; ++n) {
        if constexpr (ACT == JACT_ELU) o[n] = elu_jet_fwd_small(z0[n], P[n], Q[n]);
        else o[n] = softplus_jet_fwd_small(z0[n], P[n], Q[n], offset);
      }
    }
  } else {      // (fully unrolled as well: a runtime index would push the arrays into local memory on the main path too)
#pragma unroll
    for (int n = 0; n < N; ++n) jet_fwd_any<ACT>(z0[n], P[n], Q[n], offset, &o[n]);
  }
}
template <int ACT, int MODE, int N>
__device__ __forceinline__ void jet_bwd_n(const Jet (&o)[N], float offset, const float (&g0)[N], const float (&ga)[N],
                                          const float (&gb)[N], float (&dz0)[N], float (&dP)[N], float (&dQ)[N]) {
  if constexpr (MODE == JMODE_TAYLOR) {
#pragma unroll
    for (int n = 0; n < N; ++n) taylor_bwd<ACT>(o[n], offset, g0[n], ga[n], gb[n], dz0[n], dP[n], dQ[n]);
  } else if constexpr (ACT == JACT_ELU) {
#pragma unroll
    for (int n = 0; n < N; ++n) elu_jet_bwd(o[n], g0[n], ga[n], gb[n], dz0[n], dP[n], dQ[n]);
  } else {
    float m = 0.f;
#pragma unroll
    for (int n = 0; n < N; ++n) m = fmaxf(m, fmaxf(fabsf(o[n].a), fabsf(o[n].b)));
    if (m <= SMALL) {
      if constexpr (N == 2 && FDSM_JET_PACKED) {
        float2 z2, p2, q2;
        softplus_jet_bwd_small2(o, offset, make_float2(g0[0], g0[1]), make_float2(ga[0], ga[1]), make_float2(gb[0], gb[1]), z2, p2, q2);
        dz0[0] = z2.x; dz0[1] = z2.y; dP[0] = p2.x; dP[1] = p2.y; dQ[0] = q2.x; dQ[1] = q2.y;
      } else {
#pragma unroll
        for (int n = 0; n < N; ++n) softplus_jet_bwd_small(o[n], offset, g0[n], ga[n], gb[n], dz0[n], dP[n], dQ[n]);
      }
    } else {
#pragma unroll
      for (int n = 0; n < N; ++n) {
        float d[3];
        softplus_jet_bwd_any(&o[n], offset, g0[n], ga[n], gb[n], d);
        dz0[n] = d[0]; dP[n] = d[1]; dQ[n] = d[2];
      }
    }
  }
}

}  // namespace jet
}  // namespace fdsm
