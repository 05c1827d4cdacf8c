// Adam over ONE flat parameter / gradient / moment buffer (the optimiser step of the reference's training loops,
// runcred.py:192,246-258: torch.optim.Adam(lr, weight_decay) after loss.backward()).  The step of cfg 2 updates
// 1.8 M parameters spread over ~260 tensors; the library's multi-tensor kernels need 0.58 ms for it (chunk tables,
// one device step counter per tensor).  On flat buffers it is one HBM-bound pass: 28 B per parameter.
// Same arithmetic as torch.optim.Adam (amsgrad off, maximize off): g += wd * p; m = b1 m + (1 - b1) g;
// v = b2 v + (1 - b2) g^2; p -= lr / (1 - b1^t) * m / (sqrt(v) / sqrt(1 - b2^t) + eps).
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/vcnf_b200.h"
#include "common.cuh"

namespace vcnf {

constexpr int kAdamThreads = 256;

template <typename T>
struct AdamCoef {
  T b1, b2, omb1, omb2, wd, eps;
};

// ALGO 0: Adam.  ALGO 1: Adamax (the Glow notebook's optimiser, example/glow.ipynb:88: torch.optim.Adamax): the second
// moment is the running infinity norm u = max(b2 u, |g| + eps) and the step is lr / (1 - b1^t) * m / u.
template <typename T, int ALGO>
__device__ __forceinline__ void adam_one(T& p, T g, T& m, T& v, const AdamCoef<T>& c, T step_size, T inv_sqrt_bc2) {
  g = g + c.wd * p;                                   // wd == 0: exact
  m = c.b1 * m + c.omb1 * g;
  if (ALGO == 0) {
    v = c.b2 * v + c.omb2 * g * g;
    const T denom = sqrt(v) * inv_sqrt_bc2 + c.eps;
    p = p - step_size * (m / denom);
  } else {
    const T a = c.b2 * v, b = (g < T(0) ? -g : g) + c.eps;
    v = (a > b || b != b) ? (b != b ? b : a) : b;      // max that propagates NaN like torch.amax
    p = p - step_size * (m / v);
  }
}

// `step` (device, fp32 like torch's capturable step tensors) holds the number of steps taken so far; every CTA reads
// it before any CTA can advance it: the LAST CTA to finish (ticket counter) writes step + 1 and clears the ticket,
// so a captured launch replays correctly with no second kernel.
template <typename T, typename V, int W, int ALGO>
__global__ void __launch_bounds__(kAdamThreads) adam_flat_kernel(T* __restrict__ p, const T* __restrict__ g,
                                                                 T* __restrict__ m, T* __restrict__ v,
                                                                 float* step, unsigned* ticket, int64_t n, double lr,
                                                                 double beta1, double beta2, double eps, double wd,
                                                                 double grad_scale, const double* lr_dev,
                                                                 const void* guard, int guard_dtype) {
  __shared__ T s_step_size, s_inv_sqrt_bc2;
  __shared__ float s_t;
  __shared__ int s_skip;
  if (threadIdx.x == 0) {
    // guard: a device scalar (the step's loss); a NaN / Inf there skips the whole update -- parameters, both
    // moments and the step counter stay as they are (the reference's loops test the loss on the host before
    // optimizer.step(), runcred.py:249-250; a replayed graph has no host in the loop)
    int skip = 0;
    if (guard) {
      const double gv = guard_dtype == VCNF_F64 ? *static_cast<const double*>(guard)
                                                : static_cast<double>(*static_cast<const float*>(guard));
      skip = !(gv - gv == 0.0);
    }
    s_skip = skip;
    if (lr_dev) lr = *lr_dev;
    const float t = *step + 1.f;
    const double bc1 = 1.0 - pow(beta1, static_cast<double>(t));
    const double bc2 = 1.0 - pow(beta2, static_cast<double>(t));
    s_step_size = static_cast<T>(lr / bc1);
    s_inv_sqrt_bc2 = static_cast<T>(1.0 / sqrt(bc2));
    s_t = t;
  }
  __syncthreads();
  if (s_skip) return;
  AdamCoef<T> c;
  const T gsc = static_cast<T>(grad_scale);
  c.b1 = static_cast<T>(beta1);
  c.b2 = static_cast<T>(beta2);
  c.omb1 = static_cast<T>(1.0 - beta1);
  c.omb2 = static_cast<T>(1.0 - beta2);
  c.wd = static_cast<T>(wd);
  c.eps = static_cast<T>(eps);
  const T step_size = s_step_size, isb = s_inv_sqrt_bc2;
  const int64_t nv = n / W;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * kAdamThreads;
  V* pv = reinterpret_cast<V*>(p);
  const V* gv = reinterpret_cast<const V*>(g);
  V* mv = reinterpret_cast<V*>(m);
  V* vv = reinterpret_cast<V*>(v);
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * kAdamThreads + threadIdx.x; i < nv; i += stride) {
    V P = pv[i], G = __ldcs(gv + i), M = mv[i], Vv = vv[i];
    T* pp = reinterpret_cast<T*>(&P);
    T* gp = reinterpret_cast<T*>(&G);
    T* mp = reinterpret_cast<T*>(&M);
    T* vp = reinterpret_cast<T*>(&Vv);
#pragma unroll
    for (int k = 0; k < W; ++k) adam_one<T, ALGO>(pp[k], gp[k] * gsc, mp[k], vp[k], c, step_size, isb);
    pv[i] = P;
    mv[i] = M;
    vv[i] = Vv;
  }
  if (blockIdx.x == 0) {                                // ragged tail (n not a multiple of the vector width)
    for (int64_t i = nv * W + threadIdx.x; i < n; i += kAdamThreads)
      adam_one<T, ALGO>(p[i], g[i] * gsc, m[i], v[i], c, step_size, isb);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned prev = atomicAdd(ticket, 1u);
    if (prev == gridDim.x - 1) {
      *step = s_t;
      *ticket = 0u;
      __threadfence();
    }
  }
}

}  // namespace vcnf

using namespace vcnf;

static int adam_launch(void* p, const void* g, void* m, void* v, void* step, void* ticket, int64_t n, double lr,
                       double beta1, double beta2, double eps, double weight_decay, double grad_scale,
                       const void* lr_dev, const void* guard, int guard_dtype, int dtype, void* stream, int algo = 0) {
  if (n < 0 || (dtype != VCNF_F32 && dtype != VCNF_F64) || !step || !ticket ||
      (guard && guard_dtype != VCNF_F32 && guard_dtype != VCNF_F64))
    return VCNF_E_BADARG;
  if (n == 0) return 0;
  if (!p || !g || !m || !v) return VCNF_E_BADARG;
  const uintptr_t bits = reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) |
                         reinterpret_cast<uintptr_t>(m) | reinterpret_cast<uintptr_t>(v);
  if (bits & 15u) return VCNF_E_BADARG;               // flat buffers come from the allocator: 16 B aligned
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int w = dtype == VCNF_F32 ? 4 : 2;
  const int64_t nv = n / w;
  int64_t blocks = (nv + kAdamThreads - 1) / kAdamThreads;
  const int64_t cap = 8 * static_cast<int64_t>(sm_count());
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
#define VCNF_ADAM_LAUNCH(T, V, W, ALGO)                                                                              \
  adam_flat_kernel<T, V, W, ALGO><<<static_cast<unsigned>(blocks), kAdamThreads, 0, st>>>(                            \
      static_cast<T*>(p), static_cast<const T*>(g), static_cast<T*>(m), static_cast<T*>(v), static_cast<float*>(step), \
      static_cast<unsigned*>(ticket), n, lr, beta1, beta2, eps, weight_decay, grad_scale,                             \
      static_cast<const double*>(lr_dev), guard, guard_dtype)
  if (dtype == VCNF_F32) {
    if (algo == 0) VCNF_ADAM_LAUNCH(float, float4, 4, 0);
    else VCNF_ADAM_LAUNCH(float, float4, 4, 1);
  } else {
    if (algo == 0) VCNF_ADAM_LAUNCH(double, double2, 2, 0);
    else VCNF_ADAM_LAUNCH(double, double2, 2, 1);
  }
#undef VCNF_ADAM_LAUNCH
  return launch_status();
}

extern "C" int vcnf_adam_flat(void* p, const void* g, void* m, void* v, void* step, void* ticket, int64_t n, double lr,
                              double beta1, double beta2, double eps, double weight_decay, int dtype, void* stream) {
  return adam_launch(p, g, m, v, step, ticket, n, lr, beta1, beta2, eps, weight_decay, 1.0, nullptr, nullptr, 0, dtype,
                     stream);
}

extern "C" int vcnf_adam_flat_guarded(void* p, const void* g, void* m, void* v, void* step, void* ticket, int64_t n,
                                      const void* lr_dev, double beta1, double beta2, double eps, double weight_decay,
                                      double grad_scale, const void* guard, int guard_dtype, int dtype, void* stream) {
  if (!lr_dev) return VCNF_E_BADARG;
  return adam_launch(p, g, m, v, step, ticket, n, 0.0, beta1, beta2, eps, weight_decay, grad_scale, lr_dev, guard,
                     guard_dtype, dtype, stream);
}

extern "C" int vcnf_adamax_flat_guarded(void* p, const void* g, void* m, void* u, void* step, void* ticket, int64_t n,
                                        const void* lr_dev, double beta1, double beta2, double eps, double weight_decay,
                                        double grad_scale, const void* guard, int guard_dtype, int dtype, void* stream) {
  if (!lr_dev) return VCNF_E_BADARG;
  return adam_launch(p, g, m, u, step, ticket, n, 0.0, beta1, beta2, eps, weight_decay, grad_scale, lr_dev, guard,
                     guard_dtype, dtype, stream, 1);
}
