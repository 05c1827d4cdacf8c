// LU-parametrised weight of LULinearPermute (normflow/flows/mixing.py:352-396): from the parameter
// vectors to the dense D x D matrix W = L U and log|det W| in ONE launch, and back in one launch.
//   L: unit lower triangular, strictly-lower entries = lower_entries (np.tril_indices(D, -1) order:
//      index(i, j) = i (i - 1) / 2 + j);
//   U: strictly-upper entries = upper_entries (np.triu_indices(D, 1) order: index(i, j) =
//      i D - i (i + 1) / 2 + j - i - 1), diag(U) = softplus(unconstrained_upper_diag) + eps (:382-384);
//   log|det W| = sum_i log diag(U)_i (:394-396).
// The reference builds L and U with four index_put calls on zero matrices and multiplies activations by
// U and L separately; as ATen ops that is ~15 tiny kernels forward and as many backward PER LAYER of a
// model whose parameters are all D^2 = 841 numbers -- 1.2 ms of a 15 ms training step of cfg 2 went
// there.  D <= 64, one CTA; not a bandwidth problem, a launch-count problem.
#include "common.cuh"
#include "../../include/vcnf_b200.h"

namespace vcnf {

constexpr int kLuThreads = 256;
constexpr int kLuMaxD = 64;

template <typename T> __device__ __forceinline__ T lu_exp(T x);
template <> __device__ __forceinline__ float lu_exp<float>(float x) { return expf(x); }
template <> __device__ __forceinline__ double lu_exp<double>(double x) { return exp(x); }
template <typename T> __device__ __forceinline__ T lu_log(T x);
template <> __device__ __forceinline__ float lu_log<float>(float x) { return logf(x); }
template <> __device__ __forceinline__ double lu_log<double>(double x) { return log(x); }
template <typename T> __device__ __forceinline__ T lu_log1p(T x);
template <> __device__ __forceinline__ float lu_log1p<float>(float x) { return log1pf(x); }
template <> __device__ __forceinline__ double lu_log1p<double>(double x) { return log1p(x); }

// torch softplus (beta 1, threshold 20) and its derivative
template <typename T> __device__ __forceinline__ T lu_softplus(T x) { return x > T(20) ? x : lu_log1p(lu_exp(x)); }
template <typename T> __device__ __forceinline__ T lu_softplus_grad(T x) { return x > T(20) ? T(1) : T(1) / (T(1) + lu_exp(-x)); }

template <typename T>
__device__ __forceinline__ void lu_load(const T* lower, const T* upper, const T* diag_raw, T eps, int D, T* Ls, T* Us) {
  const int ld = D + 1;
  for (int e = threadIdx.x; e < D * D; e += kLuThreads) {
    const int i = e / D, j = e - i * D;
    Ls[i * ld + j] = j < i ? lower[i * (i - 1) / 2 + j] : (j == i ? T(1) : T(0));
    Us[i * ld + j] = j > i ? upper[i * D - i * (i + 1) / 2 + j - i - 1] : (j == i ? lu_softplus(diag_raw[i]) + eps : T(0));
  }
}

template <typename T>
__device__ __forceinline__ void lu_compose_fwd_body(const T* lower, const T* upper, const T* diag_raw, T eps, int D, T* W,
                                                    T* logabsdet) {
  extern __shared__ __align__(16) unsigned char lu_smem[];
  const int ld = D + 1;
  T* Ls = reinterpret_cast<T*>(lu_smem);
  T* Us = Ls + D * ld;
  lu_load(lower, upper, diag_raw, eps, D, Ls, Us);
  __syncthreads();
  for (int e = threadIdx.x; e < D * D; e += kLuThreads) {
    const int i = e / D, j = e - i * D;
    const int kmax = i < j ? i : j;                       // L[i][k] = 0 for k > i, U[k][j] = 0 for k > j
    T acc = T(0);
#pragma unroll 8
    for (int k = 0; k <= kmax; ++k) acc += Ls[i * ld + k] * Us[k * ld + j];
    W[e] = acc;
  }
  if (threadIdx.x < 32) {                                 // fixed-order sum of the log-diagonal
    T acc = T(0);
    for (int i = threadIdx.x; i < D; i += 32) acc += lu_log(Us[i * ld + i]);
    acc = warp_sum(acc);
    if (threadIdx.x == 0) *logabsdet = acc;
  }
}

template <typename T>
__global__ void __launch_bounds__(kLuThreads) lu_compose_fwd_kernel(const T* lower, const T* upper, const T* diag_raw, T eps,
                                                                    int D, T* W, T* logabsdet) {
  lu_compose_fwd_body(lower, upper, diag_raw, eps, D, W, logabsdet);
}

// All LU-parametrised layers of a model in ONE launch each way (blockIdx.x = layer): the weights do not depend on the
// activations, so the 16 single-CTA launches of a cfg-2 step (6.3 us forward, 10.4 us backward each, serial on the
// stream) run side by side before the first / after the last layer.  Pointers travel in the kernel parameters.
constexpr int kLuBatchMax = 32;
struct LuBatch {
  const void* lower[kLuBatchMax]; const void* upper[kLuBatchMax]; const void* diag[kLuBatchMax];
  void* out0[kLuBatchMax]; void* out1[kLuBatchMax]; void* out2[kLuBatchMax];      // fwd: W, logabsdet, - ; bwd: g_lower, g_upper, g_diag
  const void* in0[kLuBatchMax]; const void* in1[kLuBatchMax];                      // bwd: gW, glogabsdet
};

template <typename T>
__global__ void __launch_bounds__(kLuThreads) lu_compose_fwd_batched_kernel(const LuBatch b, T eps, int D) {
  const int l = blockIdx.x;
  lu_compose_fwd_body(static_cast<const T*>(b.lower[l]), static_cast<const T*>(b.upper[l]), static_cast<const T*>(b.diag[l]),
                      eps, D, static_cast<T*>(b.out0[l]), static_cast<T*>(b.out1[l]));
}

// Sampling direction (mixing.py:412-429: two triangular solves per call): W^-1 = U^-1 L^-1 and log|det W| from
// the parameter vectors in ONE launch.  The reference (and round 1's first version) ran index_put x 4, two
// batched fp64 trsm, casts and fills: ~25 tiny launches, 1.2 ms of a 5.3 ms sample(65536) call of cfg 2.
// One CTA (one solving warp) per column c of the identity, so the D independent solves run on D SMs.  Lane l
// keeps y[l] and y[l + 32] in registers (D <= 64).  The arithmetic
// is fp64 whatever the parameter dtype (like the fp64 solve it replaces), the result is rounded once.
constexpr int kLuInvThreads = 128;

__device__ __forceinline__ double warp_sum_f64(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

template <typename T>
__device__ __forceinline__ void lu_inverse_body(const T* lower, const T* upper, const T* diag_raw, T eps, int D, T* Winv,
                                                T* logabsdet) {
  extern __shared__ __align__(16) unsigned char lu_smem[];
  const int ld = D + 1;
  T* Ls = reinterpret_cast<T*>(lu_smem);
  T* Us = Ls + D * ld;
  for (int e = threadIdx.x; e < D * D; e += kLuInvThreads) {
    const int i = e / D, j = e - i * D;
    Ls[i * ld + j] = j < i ? lower[i * (i - 1) / 2 + j] : (j == i ? T(1) : T(0));
    Us[i * ld + j] = j > i ? upper[i * D - i * (i + 1) / 2 + j - i - 1] : (j == i ? lu_softplus(diag_raw[i]) + eps : T(0));
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) {
    const int c = blockIdx.x;
    // column-oriented substitution: once y_i is final it is broadcast (one shuffle) and every lane updates the
    // entries it owns -- no reduction on the dependent chain (the dot-product form cost ~700 clk per step)
    double y0 = lane == c ? 1.0 : 0.0, y1 = lane + 32 == c ? 1.0 : 0.0;      // y[lane], y[lane + 32] = e_c
    const double rd0 = lane < D ? 1.0 / static_cast<double>(Us[lane * ld + lane]) : 0.0;
    const double rd1 = lane + 32 < D ? 1.0 / static_cast<double>(Us[(lane + 32) * ld + lane + 32]) : 0.0;
    for (int i = c; i < D; ++i) {                         // L y = e_c (unit diagonal)
      const double yi = __shfl_sync(0xffffffffu, i < 32 ? y0 : y1, i & 31);
      if (lane > i && lane < D) y0 -= static_cast<double>(Ls[lane * ld + i]) * yi;
      if (lane + 32 > i && lane + 32 < D) y1 -= static_cast<double>(Ls[(lane + 32) * ld + i]) * yi;
    }
    for (int i = D - 1; i >= 0; --i) {                    // U z = y, in place
      if (lane == (i & 31)) { if (i < 32) y0 *= rd0; else y1 *= rd1; }
      const double zi = __shfl_sync(0xffffffffu, i < 32 ? y0 : y1, i & 31);
      if (lane < i) y0 -= static_cast<double>(Us[lane * ld + i]) * zi;
      if (lane + 32 < i) y1 -= static_cast<double>(Us[(lane + 32) * ld + i]) * zi;
    }
    if (lane < D) Winv[lane * D + c] = static_cast<T>(y0);              // column c of the inverse
    if (lane + 32 < D) Winv[(lane + 32) * D + c] = static_cast<T>(y1);
  }
  if (blockIdx.x == 0 && warp == 1) {
    T acc = T(0);
    for (int i = lane; i < D; i += 32) acc += lu_log(Us[i * ld + i]);
    acc = warp_sum(acc);
    if (lane == 0) *logabsdet = acc;
  }
}

template <typename T>
__global__ void __launch_bounds__(kLuInvThreads) lu_inverse_kernel(const T* lower, const T* upper, const T* diag_raw, T eps,
                                                                   int D, T* Winv, T* logabsdet) {
  lu_inverse_body(lower, upper, diag_raw, eps, D, Winv, logabsdet);
}

// n layers in one launch: grid (D, n), blockIdx.y = layer (a sampling pass inverts every LU layer of the model once).
template <typename T>
__global__ void __launch_bounds__(kLuInvThreads) lu_inverse_batched_kernel(const LuBatch b, T eps, int D) {
  const int l = blockIdx.y;
  lu_inverse_body(static_cast<const T*>(b.lower[l]), static_cast<const T*>(b.upper[l]), static_cast<const T*>(b.diag[l]), eps,
                  D, static_cast<T*>(b.out0[l]), static_cast<T*>(b.out1[l]));
}

// Given gW = dLoss/dW and gl = dLoss/d log|det|:  gL = gW U^T (strictly lower part), gU = L^T gW (upper part),
// g_diag_raw[i] = (gU[i][i] + gl / diag(U)_i) * softplus'(raw_i).
template <typename T>
__device__ __forceinline__ void lu_compose_bwd_body(const T* lower, const T* upper, const T* diag_raw, T eps,
                                                                    int D, const T* gW, const T* gl, T* g_lower, T* g_upper,
                                                                    T* g_diag) {
  extern __shared__ __align__(16) unsigned char lu_smem[];
  const int ld = D + 1;
  T* Ls = reinterpret_cast<T*>(lu_smem);
  T* Us = Ls + D * ld;
  T* Gs = Us + D * ld;
  lu_load(lower, upper, diag_raw, eps, D, Ls, Us);
  for (int e = threadIdx.x; e < D * D; e += kLuThreads) {
    const int i = e / D, j = e - i * D;
    Gs[i * ld + j] = gW ? gW[e] : T(0);
  }
  __syncthreads();
  const T glv = gl ? *gl : T(0);
  for (int e = threadIdx.x; e < D * D; e += kLuThreads) {
    const int i = e / D, j = e - i * D;
    if (j < i) {                                          // gL[i][j] = sum_c gW[i][c] U[j][c], c >= j
      T acc = T(0);
#pragma unroll 8
      for (int c = j; c < D; ++c) acc += Gs[i * ld + c] * Us[j * ld + c];
      g_lower[i * (i - 1) / 2 + j] = acc;
    } else {                                              // gU[i][j] = sum_r L[r][i] gW[r][j], r >= i
      T acc = T(0);
#pragma unroll 8
      for (int r = i; r < D; ++r) acc += Ls[r * ld + i] * Gs[r * ld + j];
      if (j > i) g_upper[i * D - i * (i + 1) / 2 + j - i - 1] = acc;
      else g_diag[i] = (acc + glv / Us[i * ld + i]) * lu_softplus_grad(diag_raw[i]);
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(kLuThreads) lu_compose_bwd_kernel(const T* lower, const T* upper, const T* diag_raw, T eps,
                                                                    int D, const T* gW, const T* gl, T* g_lower, T* g_upper,
                                                                    T* g_diag) {
  lu_compose_bwd_body(lower, upper, diag_raw, eps, D, gW, gl, g_lower, g_upper, g_diag);
}

template <typename T>
__global__ void __launch_bounds__(kLuThreads) lu_compose_bwd_batched_kernel(const LuBatch b, T eps, int D) {
  const int l = blockIdx.x;
  lu_compose_bwd_body(static_cast<const T*>(b.lower[l]), static_cast<const T*>(b.upper[l]), static_cast<const T*>(b.diag[l]),
                      eps, D, static_cast<const T*>(b.in0[l]), static_cast<const T*>(b.in1[l]), static_cast<T*>(b.out0[l]),
                      static_cast<T*>(b.out1[l]), static_cast<T*>(b.out2[l]));
}

// ---------------------------------------------------------------------------------------------------
// Glow's LU-parametrised 1x1 convolution weight (normflow/flows/mixing.py:83-102,117-118):
//   W = P (tril(L, -1) + I) (triu(U, 1) + diag(sign_S exp(log_S))),  log|det W| = sum log_S,
// with P, L, U full C x C matrices (the reference masks them with tril / triu on every call).  As library ops the
// assembly and its autograd are ~30 launches per GlowBlock and step (tril, triu, diag, exp, two matmuls, sum and their
// adjoints): 1 700 of the 5 100 launches of a cfg-5 step.  One CTA each way, C <= 64.
template <typename T>
__device__ __forceinline__ void glow_load(const T* L, const T* U, const T* sign_s, const T* log_s, int C, T* Ls, T* Us) {
  const int ld = C + 1;
#pragma unroll 4
  for (int e = threadIdx.x; e < C * C; e += kLuThreads) {
    const int i = e / C, j = e - i * C;
    Ls[i * ld + j] = j < i ? L[e] : (j == i ? T(1) : T(0));
    Us[i * ld + j] = j > i ? U[e] : (j == i ? sign_s[i] * lu_exp(log_s[i]) : T(0));
  }
}

// With an ActNorm in front (s, t != NULL; the density direction of a GlowBlock runs ActNorm.inverse then the 1x1
// convolution, flows/affine/glow.py:59-75) the two per-pixel affine maps are one:  W ((z - t) e^-s) = W'' (z - t) with
// W''[o][i] = W[o][i] e^{-s_i}; the kernel then writes W'' (and its transpose, for the input gradient) and
// logdet = ld_scale (sum log_S - sum s).
template <typename T>
__global__ void __launch_bounds__(kLuThreads) glow_w_fwd_kernel(const T* P, const T* L, const T* U, const T* sign_s,
                                                                const T* log_s, const T* an_s, int C, T ld_scale, T* W,
                                                                T* Wt, T* logdet) {
  extern __shared__ __align__(16) unsigned char lu_smem[];
  const int ld = C + 1;
  T* Ls = reinterpret_cast<T*>(lu_smem);
  T* Us = Ls + C * ld;
  T* Ms = Us + C * ld;
  T* Ps = Ms + C * ld;
  glow_load(L, U, sign_s, log_s, C, Ls, Us);
#pragma unroll 4
  for (int e = threadIdx.x; e < C * C; e += kLuThreads) Ps[(e / C) * ld + e % C] = P[e];   // coalesced, loads in flight
  __syncthreads();
  for (int e = threadIdx.x; e < C * C; e += kLuThreads) {
    const int i = e / C, j = e - i * C;
    const int kmax = i < j ? i : j;
    T acc = T(0);
#pragma unroll 8
    for (int k = 0; k <= kmax; ++k) acc += Ls[i * ld + k] * Us[k * ld + j];
    Ms[i * ld + j] = acc;
  }
  __syncthreads();
  for (int e = threadIdx.x; e < C * C; e += kLuThreads) {
    const int i = e / C, j = e - i * C;
    T acc = T(0);
#pragma unroll 8
    for (int k = 0; k < C; ++k) acc += Ps[i * ld + k] * Ms[k * ld + j];
    if (an_s) acc *= lu_exp(-an_s[j]);
    W[e] = acc;
    if (Wt) Wt[j * C + i] = acc;
  }
  if (threadIdx.x < 32) {
    T acc = T(0);
    for (int i = threadIdx.x; i < C; i += 32) acc += log_s[i] - (an_s ? an_s[i] : T(0));
    acc = warp_sum(acc);
    if (threadIdx.x == 0) *logdet = ld_scale * acc;
  }
}

// A = P^T gW;  gL = strict lower part of A Us^T;  gU = strict upper part of Ls^T A;
// g_log_s[k] = (Ls^T A)[k][k] sign_k exp(log_s_k) + ld_scale g_logdet.  gL / gU are full matrices (zeros elsewhere).
// With the ActNorm folded in (an_s, an_t, W2 = the forward's W'', gb != NULL) `gW` is the RAW product
// G[o][i] = sum gy[o] z[i] of the per-pixel map's gradient kernel and gb[o] = sum gy[o]; the map was W'' (z - t), so
//   gW''[o][i] = G[o][i] - t_i gb[o],   g_t[i] = - sum_o W''[o][i] gb[o],
//   g_s[i] = - sum_o gW''[o][i] W''[o][i] - ld_scale g_logdet,   gW[o][i] = gW''[o][i] e^{-s_i}.
template <typename T>
__global__ void __launch_bounds__(kLuThreads) glow_w_bwd_kernel(const T* P, const T* L, const T* U, const T* sign_s,
                                                                const T* log_s, const T* an_s, const T* an_t, const T* W2,
                                                                const T* gb, int C, T ld_scale, const T* gW,
                                                                const T* g_logdet, T* gL, T* gU, T* g_log_s, T* g_an_s,
                                                                T* g_an_t) {
  extern __shared__ __align__(16) unsigned char lu_smem[];
  const int ld = C + 1;
  T* Ls = reinterpret_cast<T*>(lu_smem);
  T* Us = Ls + C * ld;
  T* As = Us + C * ld;
  T* Ps = As + C * ld;
  T* Gs = Ps + C * ld;
  glow_load(L, U, sign_s, log_s, C, Ls, Us);
  const T glv = g_logdet ? *g_logdet * ld_scale : T(0);
#pragma unroll 4
  for (int e = threadIdx.x; e < C * C; e += kLuThreads) {            // staged first: a dot product over global memory
    Ps[(e / C) * ld + e % C] = P[e];                                 // is one L2 round trip per term (26 us at C = 48)
    T g = gW ? gW[e] : T(0);
    if (an_s) g -= an_t[e % C] * gb[e / C];                          // gW''
    Gs[(e / C) * ld + e % C] = g;
  }
  __syncthreads();
  if (an_s) {
    for (int i = threadIdx.x; i < C; i += kLuThreads) {               // column sums for the ActNorm parameters
      T gs = T(0), gt = T(0);
#pragma unroll 8
      for (int o = 0; o < C; ++o) {
        const T w2 = __ldg(W2 + o * C + i);
        gs -= Gs[o * ld + i] * w2;
        gt -= w2 * gb[o];
      }
      g_an_s[i] = gs - glv;
      g_an_t[i] = gt;
    }
    __syncthreads();
    for (int e = threadIdx.x; e < C * C; e += kLuThreads) Gs[(e / C) * ld + e % C] *= lu_exp(-an_s[e % C]);   // gW
    __syncthreads();
  }
  for (int e = threadIdx.x; e < C * C; e += kLuThreads) {
    const int k = e / C, j = e - k * C;
    T acc = T(0);
#pragma unroll 8
    for (int i = 0; i < C; ++i) acc += Ps[i * ld + k] * Gs[i * ld + j];
    As[k * ld + j] = acc;
  }
  __syncthreads();
  for (int e = threadIdx.x; e < C * C; e += kLuThreads) {
    const int i = e / C, j = e - i * C;
    T gl = T(0), gu = T(0);
    if (j < i) {
#pragma unroll 8
      for (int c = j; c < C; ++c) gl += As[i * ld + c] * Us[j * ld + c];
    } else {
#pragma unroll 8
      for (int r = i; r < C; ++r) gu += Ls[r * ld + i] * As[r * ld + j];
      if (j == i) {
        g_log_s[i] = gu * Us[i * ld + i] + glv;
        gu = T(0);
      }
    }
    gL[e] = gl;
    gU[e] = gu;
  }
}

}  // namespace vcnf

using namespace vcnf;

extern "C" int vcnf_lu_compose_fwd(const void* lower, const void* upper, const void* diag_raw, double eps, int D, void* W,
                                   void* logabsdet, int dtype, void* stream) {
  if (!diag_raw || !W || !logabsdet || D <= 0 || (D > 1 && (!lower || !upper)) || (dtype != VCNF_F32 && dtype != VCNF_F64))
    return VCNF_E_BADARG;
  if (D > kLuMaxD) return VCNF_E_UNSUPPORTED;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t smem = 2 * static_cast<size_t>(D) * (D + 1) * (dtype == VCNF_F32 ? 4 : 8);
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(lu_compose_fwd_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    cudaFuncSetAttribute(lu_compose_fwd_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    attr_set = true;
  }
  if (dtype == VCNF_F32)
    lu_compose_fwd_kernel<float><<<1, kLuThreads, smem, st>>>((const float*)lower, (const float*)upper, (const float*)diag_raw,
                                                           (float)eps, D, (float*)W, (float*)logabsdet);
  else
    lu_compose_fwd_kernel<double><<<1, kLuThreads, smem, st>>>((const double*)lower, (const double*)upper, (const double*)diag_raw,
                                                            eps, D, (double*)W, (double*)logabsdet);
  return launch_status();
}

extern "C" int vcnf_lu_compose_bwd(const void* lower, const void* upper, const void* diag_raw, double eps, int D,
                                   const void* gW, const void* glogabsdet, void* g_lower, void* g_upper, void* g_diag,
                                   int dtype, void* stream) {
  if (!diag_raw || !g_diag || D <= 0 || (D > 1 && (!lower || !upper || !g_lower || !g_upper)) ||
      (dtype != VCNF_F32 && dtype != VCNF_F64))
    return VCNF_E_BADARG;
  if (D > kLuMaxD) return VCNF_E_UNSUPPORTED;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t smem = 3 * static_cast<size_t>(D) * (D + 1) * (dtype == VCNF_F32 ? 4 : 8);
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(lu_compose_bwd_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    cudaFuncSetAttribute(lu_compose_bwd_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    attr_set = true;
  }
  if (dtype == VCNF_F32)
    lu_compose_bwd_kernel<float><<<1, kLuThreads, smem, st>>>((const float*)lower, (const float*)upper, (const float*)diag_raw,
                                                           (float)eps, D, (const float*)gW, (const float*)glogabsdet,
                                                           (float*)g_lower, (float*)g_upper, (float*)g_diag);
  else
    lu_compose_bwd_kernel<double><<<1, kLuThreads, smem, st>>>((const double*)lower, (const double*)upper, (const double*)diag_raw,
                                                            eps, D, (const double*)gW, (const double*)glogabsdet,
                                                            (double*)g_lower, (double*)g_upper, (double*)g_diag);
  return launch_status();
}

extern "C" int vcnf_lu_inverse(const void* lower, const void* upper, const void* diag_raw, double eps, int D, void* Winv,
                               void* logabsdet, int dtype, void* stream) {
  if (!diag_raw || !Winv || !logabsdet || D <= 0 || (D > 1 && (!lower || !upper)) || (dtype != VCNF_F32 && dtype != VCNF_F64))
    return VCNF_E_BADARG;
  if (D > kLuMaxD) return VCNF_E_UNSUPPORTED;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t es = dtype == VCNF_F32 ? 4 : 8;
  const size_t smem = 2 * static_cast<size_t>(D) * (D + 1) * es;
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(lu_inverse_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, 120 * 1024);
    cudaFuncSetAttribute(lu_inverse_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, 120 * 1024);
    attr_set = true;
  }
  if (dtype == VCNF_F32)
    lu_inverse_kernel<float><<<D, kLuInvThreads, smem, st>>>(static_cast<const float*>(lower), static_cast<const float*>(upper),
                                                          static_cast<const float*>(diag_raw), static_cast<float>(eps), D,
                                                          static_cast<float*>(Winv), static_cast<float*>(logabsdet));
  else
    lu_inverse_kernel<double><<<D, kLuInvThreads, smem, st>>>(static_cast<const double*>(lower), static_cast<const double*>(upper),
                                                           static_cast<const double*>(diag_raw), eps, D,
                                                           static_cast<double*>(Winv), static_cast<double*>(logabsdet));
  return launch_status();
}

extern "C" int vcnf_lu_inverse_batched(const void* const* lower, const void* const* upper, const void* const* diag_raw,
                                       double eps, int D, int n, void* const* Winv, void* const* logabsdet, int dtype,
                                       void* stream) {
  if (n == 0) return 0;
  if (!diag_raw || !Winv || !logabsdet || D <= 0 || n < 0 || (D > 1 && (!lower || !upper)) ||
      (dtype != VCNF_F32 && dtype != VCNF_F64))
    return VCNF_E_BADARG;
  if (D > kLuMaxD || n > kLuBatchMax) return VCNF_E_UNSUPPORTED;
  LuBatch b{};
  for (int l = 0; l < n; ++l) {
    b.lower[l] = lower ? lower[l] : nullptr; b.upper[l] = upper ? upper[l] : nullptr; b.diag[l] = diag_raw[l];
    b.out0[l] = Winv[l]; b.out1[l] = logabsdet[l];
    if (!b.diag[l] || !b.out0[l] || !b.out1[l] || (D > 1 && (!b.lower[l] || !b.upper[l]))) return VCNF_E_BADARG;
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t smem = 2 * static_cast<size_t>(D) * (D + 1) * (dtype == VCNF_F32 ? 4 : 8);
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(lu_inverse_batched_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, 120 * 1024);
    cudaFuncSetAttribute(lu_inverse_batched_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, 120 * 1024);
    attr_set = true;
  }
  const dim3 grid(static_cast<unsigned>(D), static_cast<unsigned>(n));
  if (dtype == VCNF_F32) lu_inverse_batched_kernel<float><<<grid, kLuInvThreads, smem, st>>>(b, static_cast<float>(eps), D);
  else lu_inverse_batched_kernel<double><<<grid, kLuInvThreads, smem, st>>>(b, eps, D);
  return launch_status();
}

extern "C" int vcnf_glow_w_fwd(const void* P, const void* L, const void* U, const void* sign_s, const void* log_s,
                               const void* an_s, int C, double ld_scale, void* W, void* Wt, void* logdet, int dtype,
                               void* stream) {
  if (!P || !L || !U || !sign_s || !log_s || !W || !logdet || C <= 0 || (dtype != VCNF_F32 && dtype != VCNF_F64))
    return VCNF_E_BADARG;
  if (C > kLuMaxD) return VCNF_E_UNSUPPORTED;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t smem = 4 * static_cast<size_t>(C) * (C + 1) * (dtype == VCNF_F32 ? 4 : 8);
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(glow_w_fwd_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, 140 * 1024);
    cudaFuncSetAttribute(glow_w_fwd_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, 140 * 1024);
    attr_set = true;
  }
  if (dtype == VCNF_F32)
    glow_w_fwd_kernel<float><<<1, kLuThreads, smem, st>>>((const float*)P, (const float*)L, (const float*)U,
                                                       (const float*)sign_s, (const float*)log_s, (const float*)an_s, C,
                                                       (float)ld_scale, (float*)W, (float*)Wt, (float*)logdet);
  else
    glow_w_fwd_kernel<double><<<1, kLuThreads, smem, st>>>((const double*)P, (const double*)L, (const double*)U,
                                                        (const double*)sign_s, (const double*)log_s, (const double*)an_s,
                                                        C, ld_scale, (double*)W, (double*)Wt, (double*)logdet);
  return launch_status();
}

extern "C" int vcnf_glow_w_bwd(const void* P, const void* L, const void* U, const void* sign_s, const void* log_s,
                               const void* an_s, const void* an_t, const void* W2, const void* gb, int C, double ld_scale,
                               const void* gW, const void* g_logdet, void* gL, void* gU, void* g_log_s, void* g_an_s,
                               void* g_an_t, int dtype, void* stream) {
  if (!P || !L || !U || !sign_s || !log_s || !gL || !gU || !g_log_s || C <= 0 || (dtype != VCNF_F32 && dtype != VCNF_F64) ||
      (an_s && (!an_t || !W2 || !gb || !g_an_s || !g_an_t)))
    return VCNF_E_BADARG;
  if (C > kLuMaxD) return VCNF_E_UNSUPPORTED;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t smem = 5 * static_cast<size_t>(C) * (C + 1) * (dtype == VCNF_F32 ? 4 : 8);
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(glow_w_bwd_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, 170 * 1024);
    cudaFuncSetAttribute(glow_w_bwd_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, 170 * 1024);
    attr_set = true;
  }
  if (dtype == VCNF_F32)
    glow_w_bwd_kernel<float><<<1, kLuThreads, smem, st>>>(
        (const float*)P, (const float*)L, (const float*)U, (const float*)sign_s, (const float*)log_s, (const float*)an_s,
        (const float*)an_t, (const float*)W2, (const float*)gb, C, (float)ld_scale, (const float*)gW,
        (const float*)g_logdet, (float*)gL, (float*)gU, (float*)g_log_s, (float*)g_an_s, (float*)g_an_t);
  else
    glow_w_bwd_kernel<double><<<1, kLuThreads, smem, st>>>(
        (const double*)P, (const double*)L, (const double*)U, (const double*)sign_s, (const double*)log_s,
        (const double*)an_s, (const double*)an_t, (const double*)W2, (const double*)gb, C, ld_scale, (const double*)gW,
        (const double*)g_logdet, (double*)gL, (double*)gU, (double*)g_log_s, (double*)g_an_s, (double*)g_an_t);
  return launch_status();
}

extern "C" int vcnf_lu_compose_fwd_batched(const void* const* lower, const void* const* upper, const void* const* diag_raw,
                                           double eps, int D, int n, void* const* W, void* const* logabsdet, int dtype,
                                           void* stream) {
  if (n == 0) return 0;
  if (!diag_raw || !W || !logabsdet || D <= 0 || n < 0 || (D > 1 && (!lower || !upper)) ||
      (dtype != VCNF_F32 && dtype != VCNF_F64))
    return VCNF_E_BADARG;
  if (D > kLuMaxD || n > kLuBatchMax) return VCNF_E_UNSUPPORTED;
  LuBatch b{};
  for (int l = 0; l < n; ++l) {
    b.lower[l] = lower ? lower[l] : nullptr; b.upper[l] = upper ? upper[l] : nullptr; b.diag[l] = diag_raw[l];
    b.out0[l] = W[l]; b.out1[l] = logabsdet[l];
    if (!b.diag[l] || !b.out0[l] || !b.out1[l] || (D > 1 && (!b.lower[l] || !b.upper[l]))) return VCNF_E_BADARG;
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t smem = 2 * static_cast<size_t>(D) * (D + 1) * (dtype == VCNF_F32 ? 4 : 8);
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(lu_compose_fwd_batched_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    cudaFuncSetAttribute(lu_compose_fwd_batched_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    attr_set = true;
  }
  if (dtype == VCNF_F32) lu_compose_fwd_batched_kernel<float><<<n, kLuThreads, smem, st>>>(b, static_cast<float>(eps), D);
  else lu_compose_fwd_batched_kernel<double><<<n, kLuThreads, smem, st>>>(b, eps, D);
  return launch_status();
}

extern "C" int vcnf_lu_compose_bwd_batched(const void* const* lower, const void* const* upper, const void* const* diag_raw,
                                           double eps, int D, int n, const void* const* gW, const void* const* glogabsdet,
                                           void* const* g_lower, void* const* g_upper, void* const* g_diag, int dtype,
                                           void* stream) {
  if (n == 0) return 0;
  if (!diag_raw || !g_diag || D <= 0 || n < 0 || (D > 1 && (!lower || !upper || !g_lower || !g_upper)) ||
      (dtype != VCNF_F32 && dtype != VCNF_F64))
    return VCNF_E_BADARG;
  if (D > kLuMaxD || n > kLuBatchMax) return VCNF_E_UNSUPPORTED;
  LuBatch b{};
  for (int l = 0; l < n; ++l) {
    b.lower[l] = lower ? lower[l] : nullptr; b.upper[l] = upper ? upper[l] : nullptr; b.diag[l] = diag_raw[l];
    b.in0[l] = gW ? gW[l] : nullptr; b.in1[l] = glogabsdet ? glogabsdet[l] : nullptr;
    b.out0[l] = g_lower ? g_lower[l] : nullptr; b.out1[l] = g_upper ? g_upper[l] : nullptr; b.out2[l] = g_diag[l];
    if (!b.diag[l] || !b.out2[l] || (D > 1 && (!b.lower[l] || !b.upper[l] || !b.out0[l] || !b.out1[l]))) return VCNF_E_BADARG;
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t smem = 3 * static_cast<size_t>(D) * (D + 1) * (dtype == VCNF_F32 ? 4 : 8);
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(lu_compose_bwd_batched_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    cudaFuncSetAttribute(lu_compose_bwd_batched_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    attr_set = true;
  }
  if (dtype == VCNF_F32) lu_compose_bwd_batched_kernel<float><<<n, kLuThreads, smem, st>>>(b, static_cast<float>(eps), D);
  else lu_compose_bwd_batched_kernel<double><<<n, kLuThreads, smem, st>>>(b, eps, D);
  return launch_status();
}
