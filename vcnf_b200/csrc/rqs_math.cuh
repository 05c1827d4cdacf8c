// Rational-quadratic spline element math (forward, inverse, backward), shared by every
// spline kernel.  One call handles ONE scalar input with its own K-bin parameter vector.
//
// Follows the reference definition in normflow/utils/splines.py:20-193:
//   knots   : softmax -> min + (1 - min*K)*p -> running sum -> affine map to [-B, B] with the
//             two end knots pinned (splines.py:109-119, 123-133)
//   search  : bin = #{k : x >= knot_k} - 1, last knot raised by 1e-6 (splines.py:12-17)
//   forward : splines.py:178-193      inverse : splines.py:152-177 (stable quadratic root)
//   tails   : identity outside [-B, B] (string tails) or zero (list tails, hazard A.5)
//
// The functions are __host__ __device__ so tests/hostcheck can run the very same code on the
// CPU against the oracle; the product only ever calls them from kernels.
#pragma once
#include <math.h>

#if defined(__CUDACC__)
#define VCNF_HD __host__ __device__ __forceinline__
#else
#define VCNF_HD inline
#endif

namespace vcnf {

// How the K+1 knot derivatives are read from the per-element parameter vector.
enum DerivMode : int {
  DERIV_LINEAR = 0,         // 'linear'  : K-1 stored, both ends = const       (splines.py:36-40)
  DERIV_CIRCULAR = 1,       // 'circular': K   stored, last = first           (splines.py:44-46)
  DERIV_LIST_LINEAR = 2,    // tails list, this feature linear  : K+1 stored, ends overwritten
  DERIV_LIST_CIRCULAR = 3,  // tails list, this feature circular: K+1 stored, last = first
  DERIV_FREE = 4            // tails=None (splines.py:88-193 called directly): K+1 stored and used as they are; the
                            // spline lives on [0, 2 bound] (bound = 0.5: the reference's default box [0, 1])
};

// Left end of the spline's interval: -bound for every tail mode, 0 for the bounded spline (tails=None).  The right end is
// lo + 2 bound in both cases, so everything downstream works on knot DIFFERENCES and is unchanged.
template <typename T>
VCNF_HD T rqs_lo(T bound, int mode) { return mode == DERIV_FREE ? T(0) : -bound; }

VCNF_HD int rqs_num_derivs(int K, int mode) {
  return mode == DERIV_LINEAR ? K - 1 : (mode == DERIV_CIRCULAR ? K : K + 1);
}

template <typename T> struct RqsConst;
template <> struct RqsConst<float> {
  static VCNF_HD float min_size() { return 1e-3f; }
  static VCNF_HD float min_deriv() { return 1e-3f; }
  // float32(np.log(np.exp(1 - 1e-3) - 1))
  static VCNF_HD float edge() { return 0.5397424101829529f; }
  static VCNF_HD float search_eps() { return 1e-6f; }
};
template <> struct RqsConst<double> {
  static VCNF_HD double min_size() { return 1e-3; }
  static VCNF_HD double min_deriv() { return 1e-3; }
  static VCNF_HD double edge() { return 0.5397424172369522; }
  static VCNF_HD double search_eps() { return 1e-6; }
};

VCNF_HD float vexp(float x) { return expf(x); }
VCNF_HD double vexp(double x) { return exp(x); }
VCNF_HD float vlog(float x) { return logf(x); }
VCNF_HD double vlog(double x) { return log(x); }
VCNF_HD float vlog1p(float x) { return log1pf(x); }
VCNF_HD double vlog1p(double x) { return log1p(x); }
// Division inside the closed forms (NOT on the knot / bin-search path, which stays IEEE so that bin
// indices keep agreeing with the reference): on the device in fp32 this is the 2-ulp fast divide.
VCNF_HD float vdiv(float a, float b) {
#if defined(__CUDA_ARCH__)
  return __fdividef(a, b);
#else
  return a / b;
#endif
}
VCNF_HD double vdiv(double a, double b) { return a / b; }
VCNF_HD float vsqrt(float x) { return sqrtf(x); }
VCNF_HD double vsqrt(double x) { return sqrt(x); }

// torch softplus, beta=1, threshold=20 (hazard A.9)
template <typename T> VCNF_HD T softplus20(T x) { return x > T(20) ? x : vlog1p(vexp(x)); }
template <typename T> VCNF_HD T softplus20_grad(T x) { return x > T(20) ? T(1) : T(1) / (T(1) + vexp(-x)); }

// Raw (pre-softplus) derivative parameter of knot j in [0, K]; *src receives the index into the
// stored derivative vector, or -1 when the value is the boundary constant.
template <typename T>
VCNF_HD T rqs_raw_deriv(const T* ud, int j, int K, int mode, T edge, int* src) {
  int s;
  if (mode == DERIV_LINEAR) s = (j == 0 || j == K) ? -1 : j - 1;
  else if (mode == DERIV_CIRCULAR) s = (j == K) ? 0 : j;
  else if (mode == DERIV_LIST_LINEAR) s = (j == 0 || j == K) ? -1 : j;
  else if (mode == DERIV_FREE) s = j;
  else s = (j == K) ? 0 : j;
  *src = s;
  return s < 0 ? edge : ud[s];
}

// Everything the closed forms need about the bin the input fell into.
template <typename T> struct RqsBin {
  int idx;          // bin index in [0, K-1]
  T x0, w;          // left knot and width on the search axis' partner ... (input axis: X, W)
  T y0, h;          // output-axis knot and height
  T d0, d1;         // knot derivatives at both ends
  T r0, r1;         // raw derivative parameters (for the softplus gradient)
  int s0, s1;       // their storage indices (-1 = constant)
};

// Normalised bin sizes of one axis: p[k] = softmax(scale*u)[k]; returns knots[0..K].
template <typename T, int KMAX>
VCNF_HD void rqs_axis(const T* u, T scale, int K, T bound, T lo, T* p, T* knots) {
  T m = u[0] * scale;
#pragma unroll
  for (int k = 1; k < KMAX; ++k) if (k < K) { T v = u[k] * scale; m = v > m ? v : m; }
  T sum = T(0);
#pragma unroll
  for (int k = 0; k < KMAX; ++k) if (k < K) { p[k] = vexp(u[k] * scale - m); sum += p[k]; }
  const T inv = T(1) / sum;
  const T mn = RqsConst<T>::min_size();
  const T span = T(1) - mn * T(K);
  const T two_b = bound - (-bound);
  T run = T(0);
  knots[0] = lo;
#pragma unroll
  for (int k = 0; k < KMAX; ++k) if (k < K) {
    p[k] *= inv;
    run += mn + span * p[k];
    knots[k + 1] = (k + 1 == K) ? lo + two_b : two_b * run + lo;    // end knot pinned (splines.py:118)
  }
}

template <typename T, int KMAX>
VCNF_HD int rqs_search(const T* knots, int K, T v) {
  int cnt = 0;
#pragma unroll
  for (int k = 0; k <= KMAX; ++k) if (k <= K) {
    T kn = knots[k];
    if (k == K) kn += RqsConst<T>::search_eps();
    cnt += (v >= kn) ? 1 : 0;
  }
  int b = cnt - 1;
  return b < 0 ? 0 : (b > K - 1 ? K - 1 : b);
}

template <typename T, int KMAX>
VCNF_HD T rqs_pick(const T* a, int n, int i) {
  T v = a[0];
#pragma unroll
  for (int k = 1; k <= KMAX; ++k) if (k < n && k == i) v = a[k];
  return v;
}

// Forward closed form (splines.py:178-193).
template <typename T>
VCNF_HD void rq_forward(T x, const RqsBin<T>& b, T* y, T* lad) {
  const T s = vdiv(b.h, b.w);
  const T th = vdiv(x - b.x0, b.w);
  const T t1 = th * (T(1) - th);
  const T num = b.h * (s * th * th + b.d0 * t1);
  const T den = s + (b.d0 + b.d1 - T(2) * s) * t1;
  *y = b.y0 + vdiv(num, den);
  const T dn = s * s * (b.d1 * th * th + T(2) * s * t1 + b.d0 * (T(1) - th) * (T(1) - th));
  *lad = vlog(dn) - T(2) * vlog(den);
}

// Inverse closed form (splines.py:152-177); v is a point on the output axis.  *ok = 0 when the
// discriminant is negative (the reference asserts, splines.py:164); the result is then NaN.
template <typename T>
VCNF_HD void rq_inverse(T v, const RqsBin<T>& b, T* x, T* lad, int* ok) {
  const T s = vdiv(b.h, b.w);
  const T dv = v - b.y0;
  const T q = b.d0 + b.d1 - T(2) * s;
  const T qa = dv * q + b.h * (s - b.d0);
  const T qb = b.h * b.d0 - dv * q;
  const T qc = -s * dv;
  const T disc = qb * qb - T(4) * qa * qc;
  *ok = disc >= T(0) ? 1 : 0;
  const T root = vdiv(T(2) * qc, -qb - vsqrt(disc));
  *x = root * b.w + b.x0;
  const T t1 = root * (T(1) - root);
  const T den = s + q * t1;
  const T dn = s * s * (b.d1 * root * root + T(2) * s * t1 + b.d0 * (T(1) - root) * (T(1) - root));
  *lad = -(vlog(dn) - T(2) * vlog(den));
}

// Adjoint of rq_forward: given dLoss/dy (gy) and dLoss/dlad (gl) accumulate adjoints of x and of
// the six bin quantities.
template <typename T> struct RqsBinGrad { T x0, w, y0, h, d0, d1; };

template <typename T>
VCNF_HD void rq_forward_adjoint(T x, const RqsBin<T>& b, T gy, T gl, T* gx, RqsBinGrad<T>* g) {
  const T W = b.w, H = b.h;
  const T s = H / W;
  const T th = (x - b.x0) / W;
  const T om = T(1) - th;
  const T t1 = th * om;
  const T q = b.d0 + b.d1 - T(2) * s;
  const T A = s * th * th + b.d0 * t1;
  const T N = H * A;
  const T D = s + q * t1;
  const T E = b.d1 * th * th + T(2) * s * t1 + b.d0 * om * om;
  const T M = s * s * E;

  const T Nb = gy / D;
  const T Db = -gy * N / (D * D) - T(2) * gl / D;
  const T Mb = gl / M;
  T sb = Mb * T(2) * s * E;
  const T Eb = Mb * s * s;
  T d1b = Eb * th * th;
  T thb = Eb * T(2) * b.d1 * th;
  sb += Eb * T(2) * t1;
  T t1b = Eb * T(2) * s;
  T d0b = Eb * om * om;
  T omb = Eb * T(2) * b.d0 * om;
  sb += Db;
  const T qb = Db * t1;
  t1b += Db * q;
  T Hb = Nb * A;
  const T Ab = Nb * H;
  sb += Ab * th * th;
  thb += Ab * T(2) * s * th;
  d0b += Ab * t1;
  t1b += Ab * b.d0;
  d0b += qb;
  d1b += qb;
  sb -= T(2) * qb;
  Hb += sb / W;
  T Wb = -sb * s / W;
  thb += t1b * om;
  omb += t1b * th;
  thb -= omb;
  const T dxb = thb / W;
  Wb -= thb * th / W;
  *gx = dxb;
  g->x0 = -dxb; g->w = Wb; g->y0 = gy; g->h = Hb; g->d0 = d0b; g->d1 = d1b;
}


// ---------------------------------------------------------------------------------------------
// Element drivers: raw parameter vector -> output (+ adjoints).
// Parameter vector layout (neural_spline/coupling.py:310-312): [uw(K) | uh(K) | ud(ND)].
// `wh_scale` multiplies uw and uh before the softmax (1/sqrt(hidden) on the coupling path,
// coupling.py:314-316; 1 on the autoregressive path, hazard A.8).
// ---------------------------------------------------------------------------------------------
struct RqsElemCfg {
  int K;             // bins
  int mode;          // DerivMode
  int inverse;       // 0: x -> y (splines.py:178), 1: y -> x (splines.py:152)
  int outside_zero;  // 1: list tails write 0 outside the interval (hazard A.5)
};

// Builds the bin record for input v.  Returns false when v is outside [-bound, bound] (or NaN).
template <typename T, int KMAX>
VCNF_HD bool rqs_locate(const T* prm, T v, T bound, T wh_scale, const RqsElemCfg& c,
                        T* pw, T* ph, RqsBin<T>* b) {
  const T lo = rqs_lo(bound, c.mode);
  if (!(v >= lo && v <= lo + (bound - (-bound)))) return false;
  const int K = c.K;
  T X[KMAX + 1], Y[KMAX + 1];
  rqs_axis<T, KMAX>(prm, wh_scale, K, bound, lo, pw, X);
  rqs_axis<T, KMAX>(prm + K, wh_scale, K, bound, lo, ph, Y);
  const int i = c.inverse ? rqs_search<T, KMAX>(Y, K, v) : rqs_search<T, KMAX>(X, K, v);
  b->idx = i;
  b->x0 = rqs_pick<T, KMAX>(X, K + 1, i);
  b->w = rqs_pick<T, KMAX>(X, K + 1, i + 1) - b->x0;
  b->y0 = rqs_pick<T, KMAX>(Y, K + 1, i);
  b->h = rqs_pick<T, KMAX>(Y, K + 1, i + 1) - b->y0;
  const T edge = RqsConst<T>::edge();
  b->r0 = rqs_raw_deriv(prm + 2 * K, i, K, c.mode, edge, &b->s0);
  b->r1 = rqs_raw_deriv(prm + 2 * K, i + 1, K, c.mode, edge, &b->s1);
  b->d0 = RqsConst<T>::min_deriv() + softplus20(b->r0);
  b->d1 = RqsConst<T>::min_deriv() + softplus20(b->r1);
  return true;
}

// out: transformed value; lad: log|d out / d v|; bin: -1 outside; flag |= 1 on negative discriminant.
template <typename T, int KMAX>
VCNF_HD void rqs_element(const T* prm, T v, T bound, T wh_scale, const RqsElemCfg& c,
                         T* out, T* lad, int* bin, int* flag) {
  T pw[KMAX], ph[KMAX];
  RqsBin<T> b;
  if (!rqs_locate<T, KMAX>(prm, v, bound, wh_scale, c, pw, ph, &b)) {
    *out = c.outside_zero ? T(0) : v;
    *lad = T(0);
    *bin = -1;
    return;
  }
  *bin = b.idx;
  if (c.inverse) {
    int ok;
    rq_inverse(v, b, out, lad, &ok);
    if (!ok) *flag |= 1;
  } else {
    rq_forward(v, b, out, lad);
  }
}

// Scatter the adjoints of one axis' two touched knots back to the K raw logits.
//   knot_k = 2B*sum_{j<k} (min + span*p_j) - B  for 1 <= k <= K-1; knots 0 and K are constants.
template <typename T, int KMAX>
VCNF_HD void rqs_axis_adjoint(const T* p, int K, T bound, T scale, int i, T g_lo, T g_hi, T* gu) {
  // g_lo = adjoint of knot[i], g_hi = adjoint of knot[i+1]
  const T span = T(1) - RqsConst<T>::min_size() * T(K);
  const T k = (bound - (-bound)) * span;
  const T a = (i >= 1) ? g_lo * k : T(0);             // applies to p_j, j < i
  const T bq = (i + 1 <= K - 1) ? g_hi * k : T(0);    // applies to p_j, j < i+1
  T dot = T(0);
#pragma unroll
  for (int j = 0; j < KMAX; ++j) if (j < K) {
    const T pb = (j < i ? a : T(0)) + (j < i + 1 ? bq : T(0));
    dot += p[j] * pb;
  }
#pragma unroll
  for (int j = 0; j < KMAX; ++j) if (j < K) {
    const T pb = (j < i ? a : T(0)) + (j < i + 1 ? bq : T(0));
    gu[j] = p[j] * (pb - dot) * scale;
  }
}

// Adjoint of one located element wrt its input and the six bin quantities (either direction).
// gout = dLoss/d out, glad = dLoss/d lad.
template <typename T>
VCNF_HD void rqs_bin_adjoint(T v, const RqsBin<T>& b, int inverse, T gout, T glad, T* gv, RqsBinGrad<T>* g) {
  if (!inverse) {
    rq_forward_adjoint(v, b, gout, glad, gv, g);
    return;
  }
  // v = f(x; phi) solved for x; lad_inv = -lad_fwd(x, phi).  Implicit differentiation:
  //   pass 1: adjoint of (-glad * lad_fwd) wrt x and phi
  //   pass 2: dx/dv = 1/f', dx/dphi = -(df/dphi)/f'  weighted by the total adjoint of x
  T x, li; int ok;
  rq_inverse(v, b, &x, &li, &ok);
  const T fprime = vexp(-li);                        // f'(x) = exp(lad_fwd) = exp(-lad_inv)
  T gx1; RqsBinGrad<T> g1;
  rq_forward_adjoint(x, b, T(0), -glad, &gx1, &g1);
  const T xb = gout + gx1;
  const T lam = -xb / fprime;
  T gx2; RqsBinGrad<T> g2;
  rq_forward_adjoint(x, b, lam, T(0), &gx2, &g2);
  *gv = xb / fprime;
  g->x0 = g1.x0 + g2.x0; g->w = g1.w + g2.w; g->y0 = g1.y0 + g2.y0;
  g->h = g1.h + g2.h; g->d0 = g1.d0 + g2.d0; g->d1 = g1.d1 + g2.d1;
}

// Backward of rqs_element.  gout = dLoss/d out, glad = dLoss/d lad.  Writes gv (dLoss/d v) and
// the full parameter-vector gradient gprm[0 .. 2K+ND) (zeros where nothing flows).
template <typename T, int KMAX>
VCNF_HD void rqs_element_bwd(const T* prm, T v, T bound, T wh_scale, const RqsElemCfg& c,
                             T gout, T glad, T* gv, T* gprm) {
  const int K = c.K;
  const int nd = rqs_num_derivs(K, c.mode);
  T pw[KMAX], ph[KMAX];
  RqsBin<T> b;
  if (!rqs_locate<T, KMAX>(prm, v, bound, wh_scale, c, pw, ph, &b)) {
    *gv = c.outside_zero ? T(0) : gout;               // identity tail: d out/d v = 1 (B.2 item 5)
    for (int j = 0; j < 2 * K + nd; ++j) gprm[j] = T(0);
    return;
  }
  RqsBinGrad<T> g;
  rqs_bin_adjoint(v, b, c.inverse, gout, glad, gv, &g);
  // knots: x0 = X[i], w = X[i+1]-X[i]  ->  adj X[i] = g.x0 - g.w, adj X[i+1] = g.w  (same for Y)
  rqs_axis_adjoint<T, KMAX>(pw, K, bound, wh_scale, b.idx, g.x0 - g.w, g.w, gprm);
  rqs_axis_adjoint<T, KMAX>(ph, K, bound, wh_scale, b.idx, g.y0 - g.h, g.h, gprm + K);
  for (int j = 0; j < nd; ++j) gprm[2 * K + j] = T(0);
  if (b.s0 >= 0) gprm[2 * K + b.s0] += g.d0 * softplus20_grad(b.r0);
  if (b.s1 >= 0) gprm[2 * K + b.s1] += g.d1 * softplus20_grad(b.r1);
}

// ---------------------------------------------------------------------------------------------
// Knot tables for BATCH-SHARED parameters (PiecewiseRationalQuadraticCDF,
// neural_spline/coupling.py:165-246): the knots of a feature are the same for every sample, so
// they are built once per CTA instead of once per element.  Table layout, stride KT = K+1:
//   tab[0 .. K]       X knots        tab[KT .. KT+K]   Y knots        tab[2KT .. 2KT+K]  derivatives
// The arithmetic is the very same rqs_axis / softplus20 sequence as the per-element path, so both
// give identical values for identical parameters.
// ---------------------------------------------------------------------------------------------
template <typename T, int KMAX>
VCNF_HD void rqs_build_table(const T* prm, int K, int mode, T bound, T* tab) {
  T p[KMAX], kn[KMAX + 1];
  const int KT = K + 1;
  const T lo = rqs_lo(bound, mode);
  rqs_axis<T, KMAX>(prm, T(1), K, bound, lo, p, kn);
  for (int k = 0; k <= K; ++k) tab[k] = kn[k];
  rqs_axis<T, KMAX>(prm + K, T(1), K, bound, lo, p, kn);
  for (int k = 0; k <= K; ++k) tab[KT + k] = kn[k];
  const T edge = RqsConst<T>::edge();
  for (int k = 0; k <= K; ++k) {
    int s;
    const T r = rqs_raw_deriv(prm + 2 * K, k, K, mode, edge, &s);
    tab[2 * KT + k] = RqsConst<T>::min_deriv() + softplus20(r);
  }
}

template <typename T, int KMAX>
VCNF_HD bool rqs_table_locate(const T* tab, int K, T v, T bound, int inverse, RqsBin<T>* b) {
  const int KT = K + 1;
  const T* X = tab;
  const T* Y = tab + KT;
  // (the end knots ARE the interval: -bound / bound for the tail modes, 0 / 2 bound for the bounded spline)
  if (!(v >= (inverse ? Y[0] : X[0]) && v <= (inverse ? Y[K] : X[K]))) return false;
  const T* Dv = tab + 2 * KT;
  const int i = rqs_search<T, KMAX>(inverse ? Y : X, K, v);
  b->idx = i;
  b->x0 = X[i];
  b->w = X[i + 1] - b->x0;
  b->y0 = Y[i];
  b->h = Y[i + 1] - b->y0;
  b->d0 = Dv[i];
  b->d1 = Dv[i + 1];
  b->r0 = b->r1 = T(0);
  b->s0 = b->s1 = -1;
  return true;
}

template <typename T, int KMAX>
VCNF_HD void rqs_table_element(const T* tab, T v, T bound, const RqsElemCfg& c, T* out, T* lad, int* bin,
                               int* flag) {
  RqsBin<T> b;
  if (!rqs_table_locate<T, KMAX>(tab, c.K, v, bound, c.inverse, &b)) {
    *out = c.outside_zero ? T(0) : v;
    *lad = T(0);
    *bin = -1;
    return;
  }
  *bin = b.idx;
  if (c.inverse) {
    int ok;
    rq_inverse(v, b, out, lad, &ok);
    if (!ok) *flag |= 1;
  } else {
    rq_forward(v, b, out, lad);
  }
}

// Backward of rqs_table_element: accumulates the adjoints of the touched table entries into gtab
// (same layout as tab; the caller owns it exclusively, plain +=).
template <typename T, int KMAX>
VCNF_HD void rqs_table_element_bwd(const T* tab, T v, T bound, const RqsElemCfg& c, T gout, T glad, T* gv,
                                   T* gtab) {
  RqsBin<T> b;
  if (!rqs_table_locate<T, KMAX>(tab, c.K, v, bound, c.inverse, &b)) {
    *gv = c.outside_zero ? T(0) : gout;
    return;
  }
  RqsBinGrad<T> g;
  rqs_bin_adjoint(v, b, c.inverse, gout, glad, gv, &g);
  const int KT = c.K + 1, i = b.idx;
  gtab[i] += g.x0 - g.w;
  gtab[i + 1] += g.w;
  gtab[KT + i] += g.y0 - g.h;
  gtab[KT + i + 1] += g.h;
  gtab[2 * KT + i] += g.d0;
  gtab[2 * KT + i + 1] += g.d1;
}

// Chain rule from table adjoints back to the raw parameter vector of one feature
// (softmax / running-sum / softplus; the pinned end knots carry no gradient, B.2 item 1).
template <typename T, int KMAX>
VCNF_HD void rqs_table_to_raw_grad(const T* prm, int K, int mode, T bound, const T* gtab, T* gprm) {
  const int KT = K + 1;
  const int nd = rqs_num_derivs(K, mode);
  const T span = T(1) - RqsConst<T>::min_size() * T(K);
  const T kk = (bound - (-bound)) * span;
  T p[KMAX], kn[KMAX + 1];
  for (int axis = 0; axis < 2; ++axis) {
    rqs_axis<T, KMAX>(prm + axis * K, T(1), K, bound, rqs_lo(bound, mode), p, kn);
    const T* gk = gtab + axis * KT;
    // knot_k (1 <= k <= K-1) depends on p_j for j < k
    T suffix = T(0), dot = T(0);
    T a[KMAX];
    for (int j = K - 1; j >= 0; --j) {
      if (j + 1 <= K - 1) suffix += gk[j + 1];
      a[j] = suffix * kk;
    }
    for (int j = 0; j < K; ++j) dot += p[j] * a[j];
    for (int j = 0; j < K; ++j) gprm[axis * K + j] = p[j] * (a[j] - dot);
  }
  for (int j = 0; j < nd; ++j) gprm[2 * K + j] = T(0);
  const T edge = RqsConst<T>::edge();
  for (int k = 0; k <= K; ++k) {
    int s;
    const T r = rqs_raw_deriv(prm + 2 * K, k, K, mode, edge, &s);
    if (s >= 0) gprm[2 * K + s] += gtab[2 * KT + k] * softplus20_grad(r);
  }
}

}  // namespace vcnf
