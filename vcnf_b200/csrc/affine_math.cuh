// Element math of the affine bijector family (forward, inverse, adjoints).  __host__ __device__
// so tests/hostcheck can run it on the CPU.
//
//   coupling      : normflow/flows/affine/coupling.py:113-168 (exp / sigmoid / sigmoid_inv / no scale)
//   masked affine : normflow/flows/affine/coupling.py:202-222
//   const affine  : normflow/flows/affine/coupling.py:37-53 (ActNorm: flows/normalization.py:20-38)
//   MAF           : normflow/flows/affine/autoregressive.py:76-90 (scale = sigmoid(u+2)+1e-3)
#pragma once
#include "rqs_math.cuh"

namespace vcnf {

enum ScaleMap : int { SCALE_EXP = 0, SCALE_SIGMOID = 1, SCALE_SIGMOID_INV = 2, SCALE_NONE = 3, SCALE_MAF = 4 };

template <typename T> VCNF_HD T sigmoid_(T x) { return T(1) / (T(1) + vexp(-x)); }

// y = transform(z; shift t, raw scale s).  ld = this element's log|det| contribution.
//   inverse == 0 is the layer's `forward`, inverse == 1 its `inverse`.
template <typename T>
VCNF_HD void affine_apply(int map, int inverse, T z, T t, T s, T* y, T* ld) {
  if (map == SCALE_EXP) {
    if (!inverse) { *y = z * vexp(s) + t; *ld = s; }
    else { *y = (z - t) * vexp(-s); *ld = -s; }
  } else if (map == SCALE_SIGMOID) {
    const T sc = sigmoid_(s + T(2));
    if (!inverse) { *y = z / sc + t; *ld = -vlog(sc); }
    else { *y = (z - t) * sc; *ld = vlog(sc); }
  } else if (map == SCALE_SIGMOID_INV) {
    const T sc = sigmoid_(s + T(2));
    if (!inverse) { *y = z * sc + t; *ld = vlog(sc); }
    else { *y = (z - t) / sc; *ld = -vlog(sc); }
  } else if (map == SCALE_MAF) {      // params [.., 2]: (unconstrained scale, shift)
    const T sc = sigmoid_(s + T(2)) + T(1e-3);
    if (!inverse) { *y = sc * z + t; *ld = vlog(sc); }
    else { *y = (z - t) / sc; *ld = -vlog(sc); }
  } else {
    *y = inverse ? z - t : z + t;
    *ld = T(0);
  }
}

// Adjoint: g = dLoss/dy, gl = dLoss/d(log-det of this sample).  Outputs gz, gt, gs.
template <typename T>
VCNF_HD void affine_adjoint(int map, int inverse, T z, T t, T s, T g, T gl, T* gz, T* gt, T* gs) {
  if (map == SCALE_EXP) {
    if (!inverse) { const T e = vexp(s); *gz = g * e; *gt = g; *gs = g * z * e + gl; }
    else { const T e = vexp(-s); *gz = g * e; *gt = -g * e; *gs = -g * (z - t) * e - gl; }
  } else if (map == SCALE_SIGMOID || map == SCALE_SIGMOID_INV) {
    const T sc = sigmoid_(s + T(2));
    const T dsc = T(1) - sc;                     // d log(sc) / ds
    const bool mul = (map == SCALE_SIGMOID) == (inverse != 0);   // y multiplies by sc?
    if (mul) {
      if (!inverse) { *gz = g * sc; *gt = g; *gs = (g * z * sc + gl) * dsc; }
      else { *gz = g * sc; *gt = -g * sc; *gs = (g * (z - t) * sc + gl) * dsc; }
    } else {
      if (!inverse) { *gz = g / sc; *gt = g; *gs = -(g * z / sc + gl) * dsc; }
      else { *gz = g / sc; *gt = -g / sc; *gs = -(g * (z - t) / sc + gl) * dsc; }
    }
  } else if (map == SCALE_MAF) {
    const T sg = sigmoid_(s + T(2));
    const T sc = sg + T(1e-3);
    const T dsc = sg * (T(1) - sg);              // d sc / ds
    if (!inverse) { *gz = g * sc; *gt = g; *gs = (g * z + gl / sc) * dsc; }
    else { *gz = g / sc; *gt = -g / sc; *gs = -(g * (z - t) / (sc * sc) + gl / sc) * dsc; }
  } else {
    *gz = g; *gt = inverse ? -g : g; *gs = T(0);
  }
}

template <typename T> VCNF_HD bool finite_(T v) { return v - v == T(0); }

// MaskedAffineFlow element, literal to the reference incl. NaN-ing of non-finite s, t
// (coupling.py:205-208) and arbitrary (not only 0/1) mask values b.
template <typename T>
VCNF_HD void masked_affine_apply(int inverse, T b, T z, T t, T s, T* y, T* ld) {
  const T qnan = T(NAN);
  if (!finite_(s)) s = qnan;
  if (!finite_(t)) t = qnan;
  const T zm = b * z;
  if (!inverse) { *y = zm + (T(1) - b) * (z * vexp(s) + t); *ld = (T(1) - b) * s; }
  else { *y = zm + (T(1) - b) * (z - t) * vexp(-s); *ld = -((T(1) - b) * s); }
}

template <typename T>
VCNF_HD void masked_affine_adjoint(int inverse, T b, T z, T t, T s, T g, T gl, T* gz, T* gt, T* gs) {
  const T qnan = T(NAN);
  if (!finite_(s)) s = qnan;
  if (!finite_(t)) t = qnan;
  const T nb = T(1) - b;
  if (!inverse) {
    const T e = vexp(s);
    *gz = g * (b + nb * e); *gt = nb * g; *gs = nb * (g * z * e + gl);
  } else {
    const T e = vexp(-s);
    *gz = g * (b + nb * e); *gt = -nb * g * e; *gs = -nb * (g * (z - t) * e + gl);
  }
}

}  // namespace vcnf
