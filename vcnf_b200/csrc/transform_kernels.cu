// Preprocessing either side of the path and the checkerboard split (row f2):
//   * logit_kernel      : transforms.Logit as MultiscaleFlow.transform (normflow/transforms.py:7-39) -- elementwise
//                         logit / sigmoid with the per-sample log|det| reduced in the same pass (one CTA per sample);
//                         also the data-loader Logit of utils/preprocessing.py:5-22 (no log-det)
//   * axpb_kernel       : out = a x + b + c noise -- Scale, Jitter (utils/preprocessing.py:25-56), Shift
//                         (transforms.py:42-62), in place when out == x
//   * checkerboard      : Split / Merge 'checkerboard(_inv)' (normflow/flows/reshape.py:30-44,56-72) as index maps
// HBM-bound: s (2 n + 1) bytes per sample.
#include "common.cuh"
#include "../../include/vcnf_b200.h"

namespace vcnf {
namespace tf {

constexpr int kThreads = 256;

template <typename T> __device__ __forceinline__ T tlog(T x);
template <> __device__ __forceinline__ float tlog<float>(float x) { return logf(x); }
template <> __device__ __forceinline__ double tlog<double>(double x) { return log(x); }
template <typename T> __device__ __forceinline__ T texp(T x);
template <> __device__ __forceinline__ float texp<float>(float x) { return expf(x); }
template <> __device__ __forceinline__ double texp<double>(double x) { return exp(x); }
template <typename T> __device__ __forceinline__ T tlog1p(T x);
template <> __device__ __forceinline__ float tlog1p<float>(float x) { return log1pf(x); }
template <> __device__ __forceinline__ double tlog1p<double>(double x) { return log1p(x); }

// logsigmoid(x) = min(x, 0) - log1p(exp(-|x|))   (what ATen evaluates)
template <typename T>
__device__ __forceinline__ T logsigmoid(T x) { return (x < T(0) ? x : T(0)) - tlog1p(texp(-(x < T(0) ? -x : x))); }

// to_logit = 1: u = a0 + beta z, y = log u - log(1 - u), ld = n log beta - sum(log u + log(1 - u))   (Logit.inverse)
// to_logit = 0: y = (sigmoid(z) - a0) / beta, ld = -n log beta + sum(logsigmoid(z) + logsigmoid(-z))  (Logit.forward)
// BWD: gz = gy dy/dz + gld dld/dz.
template <typename T, bool BWD>
__global__ void __launch_bounds__(kThreads) logit_kernel(const T* __restrict__ z, const T* __restrict__ gy,
                                                         const T* __restrict__ gld, T* __restrict__ out,
                                                         T* __restrict__ logdet, int64_t B, int n, T a0, T beta,
                                                         int to_logit) {
  __shared__ T red[kThreads / 32];
  for (int64_t b = blockIdx.x; b < B; b += gridDim.x) {
    T acc = T(0);
    const T gl = (BWD && gld) ? gld[b] : T(0);
    for (int j = threadIdx.x; j < n; j += kThreads) {
      const T v = z[b * n + j];
      if (to_logit) {
        const T u = a0 + beta * v;
        if (!BWD) {
          const T lu = tlog(u), l1 = tlog(T(1) - u);
          out[b * n + j] = lu - l1;
          acc += lu + l1;
        } else {
          const T iu = T(1) / u, i1 = T(1) / (T(1) - u);
          out[b * n + j] = (gy ? gy[b * n + j] : T(0)) * beta * (iu + i1) + gl * beta * (i1 - iu);
        }
      } else {
        const T s = T(1) / (T(1) + texp(-v));
        if (!BWD) {
          out[b * n + j] = (s - a0) / beta;
          acc += logsigmoid(v) + logsigmoid(-v);
        } else {
          out[b * n + j] = (gy ? gy[b * n + j] : T(0)) * s * (T(1) - s) / beta + gl * (T(1) - T(2) * s);
        }
      }
    }
    if (!BWD && logdet) {
      acc = warp_sum(acc);
      __syncthreads();
      if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
      __syncthreads();
      if (threadIdx.x == 0) {
        T t = T(0);
        for (int w = 0; w < kThreads / 32; ++w) t += red[w];
        const T nl = static_cast<T>(n) * tlog(beta);
        logdet[b] = to_logit ? nl - t : t - nl;
      }
    }
  }
}

template <typename T>
__global__ void axpb_kernel(const T* __restrict__ x, const T* __restrict__ noise, T* __restrict__ out, int64_t total,
                            T a, T b, T c) {
  for (int64_t e = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; e < total;
       e += static_cast<int64_t>(gridDim.x) * blockDim.x)
    out[e] = a * x[e] + b + (noise ? c * noise[e] : T(0));
}

// z [B, M, W] (M = product of the middle dims, W even) <-> z1, z2 [B, M, W/2].  cb(c, h, w) = (sum of the non-batch
// indices) % 2 for 'checkerboard' (reshape.py:33-40: cb0), flipped for 'checkerboard_inv'; z1 takes cb == 1.
// mid_parity[m] = parity of the middle indices of row m (host table).  merge = 0: split, 1: merge.
template <typename T>
__global__ void checkerboard_kernel(T* __restrict__ z, T* __restrict__ z1, T* __restrict__ z2,
                                    const uint8_t* __restrict__ mid_parity, int64_t total_half, int M, int Wh, int inv,
                                    int merge) {
  for (int64_t e = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; e < total_half;
       e += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int k = static_cast<int>(e % Wh);
    const int64_t row = e / Wh;                      // b * M + m
    const int p = (mid_parity[row % M] ^ inv) & 1;   // parity the rest of the index must NOT cancel
    // cb = (p + w) % 2; z1 <- cb == 1: w = 2k + 1 - p; z2 <- cb == 0: w = 2k + p
    const int64_t base = row * (2 * Wh);
    if (!merge) {
      z1[e] = z[base + 2 * k + 1 - p];
      z2[e] = z[base + 2 * k + p];
    } else {
      z[base + 2 * k + 1 - p] = z1[e];
      z[base + 2 * k + p] = z2[e];
    }
  }
}

}  // namespace tf
}  // namespace vcnf

using namespace vcnf;

static unsigned tf_grid(int64_t total, int threads) {
  const int64_t cap = static_cast<int64_t>(sm_count() > 0 ? sm_count() : VCNF_SM_COUNT_DEFAULT) * 16;
  int64_t blocks = (total + threads - 1) / threads;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return static_cast<unsigned>(blocks);
}

extern "C" {

int vcnf_logit_fwd(const void* z, void* out, void* logdet, int64_t B, int n, double a0, double beta, int to_logit,
                   int dtype, void* stream) {
  if (B == 0) return 0;
  if (!z || !out || B < 0 || n <= 0 || beta == 0.0 || (dtype != VCNF_F32 && dtype != VCNF_F64)) return VCNF_E_BADARG;
  const unsigned grid = tf_grid(B * tf::kThreads, tf::kThreads);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == VCNF_F32)
    tf::logit_kernel<float, false><<<grid, tf::kThreads, 0, st>>>(static_cast<const float*>(z), nullptr, nullptr,
                                                                  static_cast<float*>(out), static_cast<float*>(logdet),
                                                                  B, n, static_cast<float>(a0), static_cast<float>(beta),
                                                                  to_logit);
  else
    tf::logit_kernel<double, false><<<grid, tf::kThreads, 0, st>>>(static_cast<const double*>(z), nullptr, nullptr,
                                                                   static_cast<double*>(out),
                                                                   static_cast<double*>(logdet), B, n, a0, beta, to_logit);
  return launch_status();
}

int vcnf_logit_bwd(const void* z, const void* gy, const void* gld, void* gz, int64_t B, int n, double a0, double beta,
                   int to_logit, int dtype, void* stream) {
  if (B == 0) return 0;
  if (!z || !gz || B < 0 || n <= 0 || beta == 0.0 || (dtype != VCNF_F32 && dtype != VCNF_F64)) return VCNF_E_BADARG;
  const unsigned grid = tf_grid(B * tf::kThreads, tf::kThreads);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == VCNF_F32)
    tf::logit_kernel<float, true><<<grid, tf::kThreads, 0, st>>>(static_cast<const float*>(z),
                                                                 static_cast<const float*>(gy),
                                                                 static_cast<const float*>(gld), static_cast<float*>(gz),
                                                                 nullptr, B, n, static_cast<float>(a0),
                                                                 static_cast<float>(beta), to_logit);
  else
    tf::logit_kernel<double, true><<<grid, tf::kThreads, 0, st>>>(static_cast<const double*>(z),
                                                                  static_cast<const double*>(gy),
                                                                  static_cast<const double*>(gld),
                                                                  static_cast<double*>(gz), nullptr, B, n, a0, beta,
                                                                  to_logit);
  return launch_status();
}

int vcnf_axpb(const void* x, const void* noise, void* out, int64_t total, double a, double b, double c, int dtype,
              void* stream) {
  if (total == 0) return 0;
  if (!x || !out || total < 0 || (dtype != VCNF_F32 && dtype != VCNF_F64)) return VCNF_E_BADARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == VCNF_F32)
    tf::axpb_kernel<float><<<tf_grid(total, 256), 256, 0, st>>>(static_cast<const float*>(x),
                                                                static_cast<const float*>(noise),
                                                                static_cast<float*>(out), total, static_cast<float>(a),
                                                                static_cast<float>(b), static_cast<float>(c));
  else
    tf::axpb_kernel<double><<<tf_grid(total, 256), 256, 0, st>>>(static_cast<const double*>(x),
                                                                 static_cast<const double*>(noise),
                                                                 static_cast<double*>(out), total, a, b, c);
  return launch_status();
}

int vcnf_checkerboard(void* z, void* z1, void* z2, const uint8_t* mid_parity, int64_t B, int M, int W, int inv, int merge,
                      int dtype, void* stream) {
  if (B == 0) return 0;
  if (!z || !z1 || !z2 || !mid_parity || B < 0 || M <= 0 || W <= 0 || (W & 1) || (dtype != VCNF_F32 && dtype != VCNF_F64))
    return VCNF_E_BADARG;
  const int64_t total_half = B * M * (W / 2);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == VCNF_F32)
    tf::checkerboard_kernel<float><<<tf_grid(total_half, 256), 256, 0, st>>>(
        static_cast<float*>(z), static_cast<float*>(z1), static_cast<float*>(z2), mid_parity, total_half, M, W / 2,
        inv ? 1 : 0, merge);
  else
    tf::checkerboard_kernel<double><<<tf_grid(total_half, 256), 256, 0, st>>>(
        static_cast<double*>(z), static_cast<double*>(z1), static_cast<double*>(z2), mid_parity, total_half, M, W / 2,
        inv ? 1 : 0, merge);
  return launch_status();
}

}  // extern "C"
