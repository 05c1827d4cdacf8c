// fp32 fast path of the rational-quadratic spline element math (K compile-time), used by the sm_100a
// layer kernels for the default configuration of every spline wrapper (float32, 8 bins).
//
// Same definition as rqs_math.cuh (normflow/utils/splines.py:20-193) and the same
// sequence of roundings on the knot path (softmax -> min + span*p -> running sum -> affine map), but
// written for the issue slots of one SM sub-partition instead of for generality:
//   * the width axis and the height axis are evaluated as ONE stream of packed f32x2 instructions
//     (FMUL2 / FADD2 / FFMA2: one issue slot for both axes);
//   * exp / log / reciprocal / sqrt use the MUFU approximations directly (ex2, lg2, rcp, sqrt: <= 2 ulp);
//     1/sum of the softmax gets one Newton step, log1p of the softplus an atanh series -- no libm
//     slow paths, no branches;
//   * only the two derivative parameters of the bin that was hit go through the softplus;
//   * batch-shared knot tables are stored as {X, Y, D} records (one 16-byte load per knot) and searched
//     by bisection when K == 8.
// Everything is __host__ __device__ (the host versions use libm for the approximations) so that
// tests/hostcheck can check the formulas against the oracle in the GPU-less container.
#pragma once
#include "rqs_math.cuh"

namespace vcnf {
namespace fast {

struct P2 { float x, y; };                       // (width axis, height axis) or any pair evaluated together
struct alignas(16) Knot4 { float x, y, d, pad; };   // table record: X knot, Y knot, knot derivative

#if defined(__CUDA_ARCH__)
VCNF_HD P2 p2fma(P2 a, P2 b, P2 c) {
  const float2 r = __ffma2_rn(make_float2(a.x, a.y), make_float2(b.x, b.y), make_float2(c.x, c.y));
  return P2{r.x, r.y};
}
VCNF_HD P2 p2add(P2 a, P2 b) {
  const float2 r = __fadd2_rn(make_float2(a.x, a.y), make_float2(b.x, b.y));
  return P2{r.x, r.y};
}
VCNF_HD P2 p2mul(P2 a, P2 b) {
  const float2 r = __fmul2_rn(make_float2(a.x, a.y), make_float2(b.x, b.y));
  return P2{r.x, r.y};
}
VCNF_HD float f_ex2(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
VCNF_HD float f_lg2(float x) { float y; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
VCNF_HD float f_rcp(float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
VCNF_HD float f_sqrt(float x) { float y; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
#else
VCNF_HD P2 p2fma(P2 a, P2 b, P2 c) { return P2{fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y)}; }
VCNF_HD P2 p2add(P2 a, P2 b) { return P2{a.x + b.x, a.y + b.y}; }
VCNF_HD P2 p2mul(P2 a, P2 b) { return P2{a.x * b.x, a.y * b.y}; }
VCNF_HD float f_ex2(float x) { return exp2f(x); }
VCNF_HD float f_lg2(float x) { return log2f(x); }
VCNF_HD float f_rcp(float x) { return 1.0f / x; }
VCNF_HD float f_sqrt(float x) { return sqrtf(x); }
#endif

constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

// softmax -> bin sizes -> knots of BOTH axes.  p[k] = (softmax(scale*uw)[k], softmax(scale*uh)[k]),
// kn[k] = (X_k, Y_k), k = 0..K.
template <int K>
VCNF_HD void axes(const float* prm, float scale, float bound, float lo, P2* p, P2* kn) {
  P2 v[K];
  const P2 sc{scale, scale};
#pragma unroll
  for (int k = 0; k < K; ++k) v[k] = p2mul(P2{prm[k], prm[K + k]}, sc);
  float mx = v[0].x, my = v[0].y;
#pragma unroll
  for (int k = 1; k < K; ++k) { mx = fmaxf(mx, v[k].x); my = fmaxf(my, v[k].y); }
  const P2 nm{-mx, -my}, l2e{kLog2e, kLog2e};
  P2 sum{0.f, 0.f};
#pragma unroll
  for (int k = 0; k < K; ++k) {
    const P2 d = p2mul(p2add(v[k], nm), l2e);
    p[k] = P2{f_ex2(d.x), f_ex2(d.y)};
    sum = p2add(sum, p[k]);
  }
  // 1/sum, sum in [1, K]: reciprocal approximation + one Newton step
  P2 r{f_rcp(sum.x), f_rcp(sum.y)};
  const P2 e = p2fma(P2{-sum.x, -sum.y}, r, P2{1.f, 1.f});
  r = p2fma(r, e, r);
  const float mn = RqsConst<float>::min_size();
  const float span = 1.0f - mn * float(K);
  const float two_b = bound - (-bound);
  const P2 mn2{mn, mn}, span2{span, span}, tb2{two_b, two_b}, nb2{lo, lo};
  const float hi = lo + two_b;
  P2 run{0.f, 0.f};
  kn[0] = nb2;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    p[k] = p2mul(p[k], r);
    run = p2add(run, p2fma(span2, p[k], mn2));
    kn[k + 1] = (k + 1 == K) ? P2{hi, hi} : p2fma(tb2, run, nb2);         // end knot pinned (splines.py:118)
  }
}

// Storage indices of the raw derivative parameters of knots i and i+1 (-1: the boundary constant);
// same table as rqs_raw_deriv.
VCNF_HD void deriv_slots(int i, int K, int mode, int* s0, int* s1) {
  if (mode == DERIV_FREE) { *s0 = i; *s1 = i + 1; return; }
  const bool lin = (mode == DERIV_LINEAR) || (mode == DERIV_LIST_LINEAR);
  const int off = (mode == DERIV_LINEAR) ? -1 : 0;
  *s0 = (lin && i == 0) ? -1 : i + off;
  *s1 = (i + 1 == K) ? (lin ? -1 : 0) : i + 1 + off;
}

// softplus (beta 1; the threshold-20 branch of torch is the identity to fp32 precision here) of a
// pair, and exp(-|r|) for the gradient:  softplus(r) = max(r, 0) + log1p(exp(-|r|)),
// log1p(e) = 2 atanh(e / (2 + e)), e in (0, 1].
VCNF_HD P2 softplus2(P2 r, P2* e_out) {
  const P2 na{-fabsf(r.x), -fabsf(r.y)};
  const P2 t = p2mul(na, P2{kLog2e, kLog2e});
  const P2 e{f_ex2(t.x), f_ex2(t.y)};
  *e_out = e;
  const P2 den = p2add(e, P2{2.f, 2.f});
  const P2 s = p2mul(e, P2{f_rcp(den.x), f_rcp(den.y)});
  const P2 s2 = p2mul(s, s);
  P2 q{1.f / 15.f, 1.f / 15.f};
  q = p2fma(q, s2, P2{1.f / 13.f, 1.f / 13.f});
  q = p2fma(q, s2, P2{1.f / 11.f, 1.f / 11.f});
  q = p2fma(q, s2, P2{1.f / 9.f, 1.f / 9.f});
  q = p2fma(q, s2, P2{1.f / 7.f, 1.f / 7.f});
  q = p2fma(q, s2, P2{1.f / 5.f, 1.f / 5.f});
  q = p2fma(q, s2, P2{1.f / 3.f, 1.f / 3.f});
  q = p2fma(q, s2, P2{1.f, 1.f});
  const P2 two_s = p2add(s, s);
  return p2fma(two_s, q, P2{fmaxf(r.x, 0.f), fmaxf(r.y, 0.f)});
}
// d softplus / dr = sigmoid(r) from e = exp(-|r|)
VCNF_HD float sigmoid_from_e(float r, float e) {
  const float inv = f_rcp(1.f + e);
  return r >= 0.f ? inv : e * inv;
}

// Bin record of input v from the raw parameter vector [uw(K) | uh(K) | ud(ND)] (both axes, search on
// the input axis of the direction, the two knot derivatives).  Always computes (v outside the interval
// gives a meaningless but harmless record; callers select).  e01 receives exp(-|r0|), exp(-|r1|).
template <int K, bool INV>
VCNF_HD void locate(const float* prm, float v, float bound, float scale, int mode, P2* p, RqsBin<float>* b,
                    P2* e01) {
  P2 kn[K + 1];
  axes<K>(prm, scale, bound, rqs_lo(bound, mode), p, kn);
  P2 lo = kn[0], hi = kn[1];
  int idx = 0;
#pragma unroll
  for (int k = 1; k < K; ++k) {          // knots increase strictly: the last k with v >= knot_k wins
    const bool ge = v >= (INV ? kn[k].y : kn[k].x);
    lo.x = ge ? kn[k].x : lo.x;  lo.y = ge ? kn[k].y : lo.y;
    hi.x = ge ? kn[k + 1].x : hi.x;  hi.y = ge ? kn[k + 1].y : hi.y;
    idx = ge ? k : idx;
  }
  b->idx = idx;
  const P2 wh = p2add(hi, P2{-lo.x, -lo.y});
  b->x0 = lo.x; b->w = wh.x; b->y0 = lo.y; b->h = wh.y;
  deriv_slots(idx, K, mode, &b->s0, &b->s1);
  const float* ud = prm + 2 * K;
  const float edge = RqsConst<float>::edge();
  b->r0 = b->s0 < 0 ? edge : ud[b->s0 < 0 ? 0 : b->s0];
  b->r1 = b->s1 < 0 ? edge : ud[b->s1 < 0 ? 0 : b->s1];
  const P2 sp = softplus2(P2{b->r0, b->r1}, e01);
  b->d0 = RqsConst<float>::min_deriv() + sp.x;
  b->d1 = RqsConst<float>::min_deriv() + sp.y;
}

// Forward closed form (splines.py:178-193).
VCNF_HD void rq_forward(float x, const RqsBin<float>& b, float* y, float* lad) {
  const float rw = f_rcp(b.w);
  const float s = b.h * rw;
  const float th = (x - b.x0) * rw;
  const float om = 1.f - th;
  const float t1 = th * om;
  const float tt = th * th;
  const float num = b.h * fmaf(s, tt, b.d0 * t1);
  const float q = b.d0 + b.d1 - 2.f * s;
  const float den = fmaf(q, t1, s);
  const float rden = f_rcp(den);
  *y = fmaf(num, rden, b.y0);
  const float E = fmaf(b.d1, tt, fmaf(2.f * s, t1, b.d0 * om * om));
  *lad = kLn2 * f_lg2(s * s * E * rden * rden);
}

// Inverse closed form (splines.py:152-177); *ok = 0 on a negative discriminant (the reference asserts).
VCNF_HD void rq_inverse(float v, const RqsBin<float>& b, float* x, float* lad, int* ok) {
  const float rw = f_rcp(b.w);
  const float s = b.h * rw;
  const float dv = v - b.y0;
  const float q = b.d0 + b.d1 - 2.f * s;
  const float qa = fmaf(dv, q, b.h * (s - b.d0));
  const float qb = fmaf(-dv, q, b.h * b.d0);
  const float qc = -s * dv;
  const float disc = fmaf(qb, qb, -4.f * qa * qc);
  *ok = disc >= 0.f ? 1 : 0;
  const float root = (2.f * qc) * f_rcp(-qb - f_sqrt(disc));
  *x = fmaf(root, b.w, b.x0);
  const float om = 1.f - root;
  const float t1 = root * om;
  const float den = fmaf(q, t1, s);
  const float E = fmaf(b.d1, root * root, fmaf(2.f * s, t1, b.d0 * om * om));
  *lad = kLn2 * f_lg2(den * den * f_rcp(s * s * E));
}

// One element with its own parameter vector (conditional parameters).
template <int K, bool INV>
VCNF_HD void element(const float* prm, float v, float bound, float scale, int mode, int outside_zero,
                     float* out, float* lad, int* bin, int* flag) {
  P2 p[K], e01;
  RqsBin<float> b;
  locate<K, INV>(prm, v, bound, scale, mode, p, &b, &e01);
  float o, l;
  int ok = 1;
  if (INV) rq_inverse(v, b, &o, &l, &ok); else rq_forward(v, b, &o, &l);
  const float left = rqs_lo(bound, mode);
  const bool inside = (v >= left) && (v <= left + (bound - (-bound)));
  *out = inside ? o : (outside_zero ? 0.f : v);
  *lad = inside ? l : 0.f;
  *bin = inside ? b.idx : -1;
  if (inside && !ok) *flag |= 1;
}

// ---- batch-shared parameters: knot records built once per CTA -------------------------------------
// rec[k] = {X_k, Y_k, D_k}, k = 0..K (same arithmetic as the per-element path).
template <int K>
VCNF_HD void build_records(const float* prm, int mode, float bound, Knot4* rec) {
  P2 p[K], kn[K + 1];
  axes<K>(prm, 1.f, bound, rqs_lo(bound, mode), p, kn);
  const float edge = RqsConst<float>::edge();
  for (int k = 0; k <= K; k += 2) {
    int s;
    const float ra = rqs_raw_deriv(prm + 2 * K, k, K, mode, edge, &s);
    const float rb = (k + 1 <= K) ? rqs_raw_deriv(prm + 2 * K, k + 1, K, mode, edge, &s) : edge;
    P2 e;
    const P2 sp = softplus2(P2{ra, rb}, &e);
    rec[k] = Knot4{kn[k].x, kn[k].y, RqsConst<float>::min_deriv() + sp.x, 0.f};
    if (k + 1 <= K) rec[k + 1] = Knot4{kn[k + 1].x, kn[k + 1].y, RqsConst<float>::min_deriv() + sp.y, 0.f};
  }
}

template <int K, bool INV>
VCNF_HD void table_locate(const Knot4* rec, float v, RqsBin<float>* b) {
  int i = 0;
  if (K == 8) {                                   // bisection: 3 dependent loads instead of 7
    i = (v >= (INV ? rec[4].y : rec[4].x)) ? 4 : 0;
    i += (v >= (INV ? rec[i + 2].y : rec[i + 2].x)) ? 2 : 0;
    i += (v >= (INV ? rec[i + 1].y : rec[i + 1].x)) ? 1 : 0;
  } else {
#pragma unroll
    for (int k = 1; k < K; ++k) i += (v >= (INV ? rec[k].y : rec[k].x)) ? 1 : 0;
  }
  const Knot4 lo = rec[i], hi = rec[i + 1];
  b->idx = i;
  b->x0 = lo.x; b->w = hi.x - lo.x;
  b->y0 = lo.y; b->h = hi.y - lo.y;
  b->d0 = lo.d; b->d1 = hi.d;
  b->r0 = b->r1 = 0.f;
  b->s0 = b->s1 = -1;
}

template <int K, bool INV>
VCNF_HD void table_element(const Knot4* rec, float v, float bound, int outside_zero, float* out, float* lad,
                           int* bin, int* flag) {
  RqsBin<float> b;
  table_locate<K, INV>(rec, v, &b);
  float o, l;
  int ok = 1;
  if (INV) rq_inverse(v, b, &o, &l, &ok); else rq_forward(v, b, &o, &l);
  // the end records ARE the interval (-bound / bound for the tail modes, 0 / 2 bound for the bounded spline)
  const bool inside = (v >= (INV ? rec[0].y : rec[0].x)) && (v <= (INV ? rec[K].y : rec[K].x));
  *out = inside ? o : (outside_zero ? 0.f : v);
  *lad = inside ? l : 0.f;
  *bin = inside ? b.idx : -1;
  if (inside && !ok) *flag |= 1;
}

// ---- backward ---------------------------------------------------------------------------------------
// Adjoint of rq_forward (same derivation as rqs_math.cuh: rq_forward_adjoint; divisions -> reciprocals).
VCNF_HD void rq_forward_adjoint(float x, const RqsBin<float>& b, float gy, float gl, float* gx,
                                RqsBinGrad<float>* g) {
  const float H = b.h;
  const float rW = f_rcp(b.w);
  const float s = H * rW;
  const float th = (x - b.x0) * rW;
  const float om = 1.f - th;
  const float t1 = th * om;
  const float tt = th * th;
  const float q = b.d0 + b.d1 - 2.f * s;
  const float A = fmaf(s, tt, b.d0 * t1);
  const float N = H * A;
  const float D = fmaf(q, t1, s);
  const float E = fmaf(b.d1, tt, fmaf(2.f * s, t1, b.d0 * om * om));
  const float M = s * s * E;
  const float rD = f_rcp(D), rM = f_rcp(M);

  const float Nb = gy * rD;
  const float Db = -(fmaf(Nb, N, 2.f * gl)) * rD;
  const float Mb = gl * rM;
  const float Eb = Mb * s * s;
  float sb = fmaf(Mb * 2.f * s, E, fmaf(Eb * 2.f, t1, Db));
  float d1b = Eb * tt;
  float thb = Eb * 2.f * b.d1 * th;
  float t1b = fmaf(Eb * 2.f, s, Db * q);
  float d0b = Eb * om * om;
  float omb = Eb * 2.f * b.d0 * om;
  const float qb = Db * t1;
  float Hb = Nb * A;
  const float Ab = Nb * H;
  sb = fmaf(Ab, tt, sb);
  thb = fmaf(Ab * 2.f * s, th, thb);
  d0b = fmaf(Ab, t1, d0b) + qb;
  t1b = fmaf(Ab, b.d0, t1b);
  d1b += qb;
  sb = fmaf(-2.f, qb, sb);
  Hb = fmaf(sb, rW, Hb);
  float Wb = -sb * s * rW;
  thb = fmaf(t1b, om, thb);
  omb = fmaf(t1b, th, omb);
  thb -= omb;
  const float dxb = thb * rW;
  Wb = fmaf(-dxb, th, Wb);
  *gx = dxb;
  g->x0 = -dxb; g->w = Wb; g->y0 = gy; g->h = Hb; g->d0 = d0b; g->d1 = d1b;
}

// Adjoint of one located element wrt its input and the six bin quantities (either direction); the
// inverse direction by implicit differentiation (rqs_math.cuh: rqs_bin_adjoint).
template <bool INV>
VCNF_HD void bin_adjoint(float v, const RqsBin<float>& b, float gout, float glad, float* gv, RqsBinGrad<float>* g) {
  if (!INV) {
    rq_forward_adjoint(v, b, gout, glad, gv, g);
    return;
  }
  float x, li;
  int ok;
  rq_inverse(v, b, &x, &li, &ok);
  const float rfp = f_ex2(li * kLog2e);              // 1 / f'(x) = exp(lad_inv)
  float gx1;
  RqsBinGrad<float> g1;
  rq_forward_adjoint(x, b, 0.f, -glad, &gx1, &g1);
  const float xb = gout + gx1;
  const float lam = -xb * rfp;
  float gx2;
  RqsBinGrad<float> g2;
  rq_forward_adjoint(x, b, lam, 0.f, &gx2, &g2);
  *gv = xb * rfp;
  g->x0 = g1.x0 + g2.x0; g->w = g1.w + g2.w; g->y0 = g1.y0 + g2.y0;
  g->h = g1.h + g2.h; g->d0 = g1.d0 + g2.d0; g->d1 = g1.d1 + g2.d1;
}

// Backward of `element`: writes gv and the full parameter-vector gradient gprm[0 .. 2K+nd).
template <int K, bool INV>
VCNF_HD void element_bwd(const float* prm, float v, float bound, float scale, int mode, int outside_zero,
                         float gout, float glad, float* gv, float* gprm) {
  const int nd = rqs_num_derivs(K, mode);
  const float left = rqs_lo(bound, mode);
  if (!((v >= left) && (v <= left + (bound - (-bound))))) {
    *gv = outside_zero ? 0.f : gout;                 // identity tail: d out / d v = 1
    for (int j = 0; j < 2 * K + nd; ++j) gprm[j] = 0.f;
    return;
  }
  P2 p[K], e01;
  RqsBin<float> b;
  locate<K, INV>(prm, v, bound, scale, mode, p, &b, &e01);
  RqsBinGrad<float> g;
  bin_adjoint<INV>(v, b, gout, glad, gv, &g);
  // knots: x0 = X[i], w = X[i+1] - X[i] -> adj X[i] = g.x0 - g.w, adj X[i+1] = g.w (same for Y); knot_k,
  // 1 <= k <= K-1, depends on p_j for j < k through the running sum; the softmax Jacobian follows.
  const int i = b.idx;
  const float kk = (bound - (-bound)) * (1.0f - RqsConst<float>::min_size() * float(K));
  const P2 k2{kk, kk};
  const P2 zero{0.f, 0.f};
  const P2 glo = p2mul(P2{g.x0 - g.w, g.y0 - g.h}, k2), ghi = p2mul(P2{g.w, g.h}, k2);
  const P2 bq = (i + 1 <= K - 1) ? ghi : zero;       // applies to p_j, j <= i
  const P2 ab = p2add((i >= 1) ? glo : zero, bq);    // applies to p_j, j <  i
  P2 pb[K];
  P2 dot = zero;
#pragma unroll
  for (int j = 0; j < K; ++j) {
    pb[j] = (j < i) ? ab : ((j == i) ? bq : zero);
    dot = p2fma(p[j], pb[j], dot);
  }
  const P2 ndot{-dot.x, -dot.y}, sc{scale, scale};
#pragma unroll
  for (int j = 0; j < K; ++j) {
    const P2 gu = p2mul(p2mul(p[j], p2add(pb[j], ndot)), sc);
    gprm[j] = gu.x;
    gprm[K + j] = gu.y;
  }
  for (int j = 0; j < nd; ++j) gprm[2 * K + j] = 0.f;
  if (b.s0 >= 0) gprm[2 * K + b.s0] = g.d0 * sigmoid_from_e(b.r0, e01.x);
  if (b.s1 >= 0) gprm[2 * K + b.s1] = g.d1 * sigmoid_from_e(b.r1, e01.y);
}

// Backward of `table_element`: accumulates the adjoints of the touched knots into gtab (the generic table
// layout of rqs_math.cuh: X | Y | D with stride K+1; the caller owns it exclusively).
template <int K, bool INV>
VCNF_HD void table_element_bwd(const Knot4* rec, float v, float bound, int outside_zero, float gout, float glad,
                               float* gv, float* gtab) {
  if (!((v >= (INV ? rec[0].y : rec[0].x)) && (v <= (INV ? rec[K].y : rec[K].x)))) {
    *gv = outside_zero ? 0.f : gout;
    return;
  }
  RqsBin<float> b;
  table_locate<K, INV>(rec, v, &b);
  RqsBinGrad<float> g;
  bin_adjoint<INV>(v, b, gout, glad, gv, &g);
  const int KT = K + 1, i = b.idx;
  gtab[i] += g.x0 - g.w;
  gtab[i + 1] += g.w;
  gtab[KT + i] += g.y0 - g.h;
  gtab[KT + i + 1] += g.h;
  gtab[2 * KT + i] += g.d0;
  gtab[2 * KT + i + 1] += g.d1;
}

}  // namespace fast
}  // namespace vcnf
