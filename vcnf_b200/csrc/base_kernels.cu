// Base-distribution ends of the flow (reference normflow/distributions/base.py:609-652 DiagGaussian,
// :715-775 ClassCondDiagGaussian, :957-1034 GaussianMixture): one kernel per end instead of ~10-20 ATen launches and
// several [B, d] / [B, K, d] temporaries.
//   gauss_logprob : log N(z; loc_y, exp(log_scale_y)) summed over the event dims, added to the running log-density
//   gauss_sample  : z = loc_y + exp(log_scale_y) eps and its log-density from the same pass (eps is handed in)
//   gmm_logprob   : log sum_k w_k N(z; mu_k, sigma_k) with softmax weights
// plus their backward kernels (input gradient + parameter gradients reduced over the batch in a fixed order).
// All are HBM-bound streaming kernels: s (D + 1) bytes per sample forward, s (2 D + 1) backward.
#include "common.cuh"
#include "../../include/vcnf_b200.h"

namespace vcnf {
namespace base {

constexpr int kThreads = 256;
constexpr int kChunks = 148;
constexpr double kHalfLog2Pi = 0.91893853320467274178;

template <typename T> __device__ __forceinline__ T texp(T x);
template <> __device__ __forceinline__ float texp<float>(float x) { return expf(x); }
template <> __device__ __forceinline__ double texp<double>(double x) { return exp(x); }

template <typename T>
__device__ __forceinline__ T block_sum(T v, T* red) {
  v = warp_sum(v);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  T t = T(0);
  for (int w = 0; w < kThreads / 32; ++w) t += red[w];
  return t;
}

// One CTA per row when n >= 256, else one warp per row.  mode 0: log-prob of z; mode 1: sample from eps (z out).
template <typename T>
__global__ void __launch_bounds__(kThreads) gauss_fwd_kernel(const T* __restrict__ zin, const T* __restrict__ loc,
                                                             const T* __restrict__ ls, int64_t js, int64_t cs,
                                                             const int64_t* __restrict__ y, T log_temp,
                                                             const T* __restrict__ accum, T* __restrict__ zout,
                                                             T* __restrict__ logp, int64_t B, int n, int mode,
                                                             int warp_rows) {
  __shared__ T red[kThreads / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t rows_per_cta = warp_rows ? kThreads / 32 : 1;
  for (int64_t r0 = blockIdx.x * rows_per_cta; r0 < B; r0 += static_cast<int64_t>(gridDim.x) * rows_per_cta) {
    const int64_t b = warp_rows ? r0 + warp : r0;
    const bool live = b < B;
    const int64_t c = (live && y) ? y[b] : 0;
    T acc = T(0);
    if (live) {
      const int start = warp_rows ? lane : threadIdx.x, step = warp_rows ? 32 : kThreads;
      for (int j = start; j < n; j += step) {
        const T l = __ldg(ls + j * js + c * cs) + log_temp;
        const T v = zin[b * n + j];
        if (mode == 0) {
          const T e = (v - __ldg(loc + j * js + c * cs)) * texp(-l);
          acc += l + T(0.5) * e * e;
        } else {
          zout[b * n + j] = __ldg(loc + j * js + c * cs) + texp(l) * v;
          acc += l + T(0.5) * v * v;
        }
      }
    }
    T tot;
    if (warp_rows) tot = warp_sum(acc);
    else tot = block_sum(acc, red);
    if (live && (warp_rows ? lane == 0 : threadIdx.x == 0))
      logp[b] = (accum ? accum[b] : T(0)) + static_cast<T>(-kHalfLog2Pi * n) - tot;
  }
}

// Backward, input gradient only (frozen base parameters: cfg 2 / 3): a flat elementwise pass.
template <typename T>
__global__ void __launch_bounds__(kThreads) gauss_gz_kernel(const T* __restrict__ zin, const T* __restrict__ loc,
                                                            const T* __restrict__ ls, int64_t js, int64_t cs,
                                                            const int64_t* __restrict__ y, T log_temp,
                                                            const T* __restrict__ g_logp, T* __restrict__ gz,
                                                            int64_t total, int n) {
  for (int64_t e = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; e < total;
       e += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t b = e / n;
    const int j = static_cast<int>(e - b * n);
    const int64_t c = y ? y[b] : 0;
    const T il = texp(-(__ldg(ls + j * js + c * cs) + log_temp));
    gz[e] = -g_logp[b] * (zin[e] - __ldg(loc + j * js + c * cs)) * il * il;
  }
}

// Backward with parameter gradients.  gz (mode 0: d/dz; mode 1: nothing, eps has no gradient) and per-chunk partial
// parameter gradients partial[chunk][2][classes][n].  A CTA owns a contiguous chunk of rows; thread t keeps column
// j = t % npad of the rows t / npad, t / npad + G, ... (G = 256 / npad row groups, npad = n rounded up so that whole
// groups fit), so all 256 threads work for narrow rows too; its sums go to a private slot [group][class][j] of the
// shared table (single writer, no atomics), the G slots are added in group order at the end (fixed order).
template <typename T>
__global__ void __launch_bounds__(kThreads) gauss_bwd_kernel(const T* __restrict__ zin, const T* __restrict__ loc,
                                                             const T* __restrict__ ls, int64_t js, int64_t cs,
                                                             const int64_t* __restrict__ y, T log_temp,
                                                             const T* __restrict__ g_logp, const T* __restrict__ g_zout,
                                                             T* __restrict__ gz, T* __restrict__ partial, int64_t B,
                                                             int n, int classes, int mode, int64_t rows_per_chunk,
                                                             int G) {
  extern __shared__ unsigned char smem_raw[];
  T* tab = reinterpret_cast<T*>(smem_raw);          // [G][2][classes][n]
  const int chunk = blockIdx.x;
  const int64_t b0 = chunk * rows_per_chunk, b1 = b0 + rows_per_chunk < B ? b0 + rows_per_chunk : B;
  const int64_t tsize = static_cast<int64_t>(classes) * n;
  for (int64_t i = threadIdx.x; i < 2 * tsize * G; i += kThreads) tab[i] = T(0);
  __syncthreads();
  const int per = kThreads / G;                     // threads per row group
  const int grp = threadIdx.x / per, j0 = threadIdx.x - grp * per;
  T* mine = tab + static_cast<int64_t>(grp) * 2 * tsize;
  if (grp < G) {
    for (int64_t b = b0 + grp; b < b1; b += G) {
      const int64_t c = y ? y[b] : 0;
      const T g = g_logp ? g_logp[b] : T(0);
      for (int j = j0; j < n; j += per) {
        const T l = __ldg(ls + j * js + c * cs) + log_temp;
        const T v = zin[b * n + j];
        T gl, gs;
        if (mode == 0) {
          const T il = texp(-l);
          const T e = (v - __ldg(loc + j * js + c * cs)) * il;
          const T gzv = -g * e * il;
          if (gz) gz[b * n + j] = gzv;
          gl = -gzv;
          gs = g * (e * e - T(1));
        } else {
          const T go = g_zout ? g_zout[b * n + j] : T(0);
          gl = go;
          gs = go * texp(l) * v - g;
        }
        mine[c * n + j] += gl;
        mine[tsize + c * n + j] += gs;
      }
    }
  }
  __syncthreads();
  for (int64_t i = threadIdx.x; i < 2 * tsize; i += kThreads) {
    T acc = T(0);
    for (int g2 = 0; g2 < G; ++g2) acc += tab[static_cast<int64_t>(g2) * 2 * tsize + i];
    partial[static_cast<int64_t>(chunk) * 2 * tsize + i] = acc;
  }
}

template <typename T>
__global__ void reduce_chunks_kernel(const T* __restrict__ partial, int chunks, int64_t stride, int64_t total,
                                     T* __restrict__ out) {
  const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= total) return;
  T acc = T(0);
  for (int c = 0; c < chunks; ++c) acc += partial[static_cast<int64_t>(c) * stride + i];
  out[i] = acc;
}

// ---- Gaussian mixture (base.py:1023-1034).  CTA = kGmmRows rows: T[b][k] in shared memory, thread per (b, k).
constexpr int kGmmRows = 32;
template <typename T>
__global__ void __launch_bounds__(kThreads) gmm_kernel(const T* __restrict__ z, const T* __restrict__ mu,
                                                       const T* __restrict__ ls, const T* __restrict__ ws,
                                                       const T* __restrict__ accum, T* __restrict__ logp,
                                                       const T* __restrict__ g_logp, T* __restrict__ gz,
                                                       T* __restrict__ partial, int64_t B, int K, int d,
                                                       int64_t rows_per_chunk, int bwd) {
  extern __shared__ unsigned char smem_raw[];
  T* s_lw = reinterpret_cast<T*>(smem_raw);         // [K] log softmax weights
  T* s_t = s_lw + K;                                // [kGmmRows][K]
  T* s_z = s_t + kGmmRows * K;                      // [kGmmRows][d]
  T* s_g = s_z + kGmmRows * d;                      // [kGmmRows]
  T* s_red = s_g + kGmmRows;                        // [8]
  T* s_il = s_red + 8;                              // [K*d] exp(-log_scale): one exp per parameter, not per sample
  T* s_sl = s_il + K * d;                           // [K]   sum_j log_scale
  T* s_acc = s_sl + K;                              // bwd: [K*d*2 + K] accumulators of this CTA's chunk
  // log-softmax of the weight scores
  T m = -INFINITY;
  for (int k = threadIdx.x; k < K; k += kThreads) m = fmax(m, ws[k]);
  for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = m;
  __syncthreads();
  m = s_red[0];
  for (int w = 1; w < kThreads / 32; ++w) m = fmax(m, s_red[w]);
  T se = T(0);
  for (int k = threadIdx.x; k < K; k += kThreads) se += texp(ws[k] - m);
  se = block_sum(se, s_red);
  const T lse_w = m + log(se);
  for (int k = threadIdx.x; k < K; k += kThreads) s_lw[k] = ws[k] - lse_w;
  for (int i = threadIdx.x; i < K * d; i += kThreads) s_il[i] = texp(-__ldg(ls + i));
  for (int k = threadIdx.x; k < K; k += kThreads) {
    T t = T(0);
    for (int j = 0; j < d; ++j) t += __ldg(ls + k * d + j);
    s_sl[k] = t;
  }
  const int nacc = K * d * 2 + K;
  if (bwd)
    for (int i = threadIdx.x; i < nacc; i += kThreads) s_acc[i] = T(0);
  __syncthreads();
  const int chunk = blockIdx.x;
  const int64_t c0 = chunk * rows_per_chunk, c1 = c0 + rows_per_chunk < B ? c0 + rows_per_chunk : B;
  for (int64_t r0 = c0; r0 < c1; r0 += kGmmRows) {
    const int rows = static_cast<int>(c1 - r0 < kGmmRows ? c1 - r0 : kGmmRows);
    for (int i = threadIdx.x; i < rows * d; i += kThreads) s_z[i] = z[r0 * d + i];
    if (bwd)
      for (int i = threadIdx.x; i < rows; i += kThreads) s_g[i] = g_logp ? g_logp[r0 + i] : T(0);
    __syncthreads();
    for (int i = threadIdx.x; i < rows * K; i += kThreads) {
      const int b = i / K, k = i - b * K;
      T q = T(0);
      for (int j = 0; j < d; ++j) {
        const T e = (s_z[b * d + j] - __ldg(mu + k * d + j)) * s_il[k * d + j];
        q += e * e;
      }
      s_t[b * K + k] = static_cast<T>(-kHalfLog2Pi * d) + s_lw[k] - T(0.5) * q - s_sl[k];
    }
    __syncthreads();
    // per-row logsumexp: one warp per row
    for (int b = threadIdx.x >> 5; b < rows; b += kThreads / 32) {
      const int lane = threadIdx.x & 31;
      T mx = -INFINITY;
      for (int k = lane; k < K; k += 32) mx = fmax(mx, s_t[b * K + k]);
      for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      T s = T(0);
      for (int k = lane; k < K; k += 32) s += texp(s_t[b * K + k] - mx);
      s = warp_sum(s);
      const T lse = mx + log(s);
      if (!bwd && lane == 0) logp[r0 + b] = (accum ? accum[r0 + b] : T(0)) + lse;
      if (bwd)
        for (int k = lane; k < K; k += 32) s_t[b * K + k] = texp(s_t[b * K + k] - lse);   // responsibilities
    }
    __syncthreads();
    if (bwd) {
      // parameter gradients: thread per (k, j), rows in order
      for (int i = threadIdx.x; i < K * d; i += kThreads) {
        const int k = i / d, j = i - k * d;
        const T il = s_il[i], mk = __ldg(mu + i);
        T am = T(0), al = T(0);
        for (int b = 0; b < rows; ++b) {
          const T w = s_g[b] * s_t[b * K + k];
          const T e = (s_z[b * d + j] - mk) * il;
          am += w * e * il;
          al += w * (e * e - T(1));
        }
        s_acc[i] += am;
        s_acc[K * d + i] += al;
      }
      for (int k = threadIdx.x; k < K; k += kThreads) {
        const T wk = texp(s_lw[k]);
        T aw = T(0);
        for (int b = 0; b < rows; ++b) aw += s_g[b] * (s_t[b * K + k] - wk);
        s_acc[2 * K * d + k] += aw;
      }
      if (gz) {
        for (int i = threadIdx.x; i < rows * d; i += kThreads) {
          const int b = i / d, j = i - b * d;
          T a = T(0);
          for (int k = 0; k < K; ++k) {
            const T il = s_il[k * d + j];
            a -= s_t[b * K + k] * (s_z[i] - __ldg(mu + k * d + j)) * il * il;
          }
          gz[r0 * d + i] = s_g[b] * a;
        }
      }
      __syncthreads();
    }
  }
  if (bwd) {
    __syncthreads();
    for (int i = threadIdx.x; i < nacc; i += kThreads) partial[static_cast<int64_t>(chunk) * nacc + i] = s_acc[i];
  }
}

}  // namespace base
}  // namespace vcnf

using namespace vcnf;

static int64_t plan_chunks(int64_t B, int64_t* rpc, int min_rows) {
  int64_t chunks = (B + min_rows - 1) / min_rows;
  if (chunks > base::kChunks) chunks = base::kChunks;
  if (chunks < 1) chunks = 1;
  *rpc = (B + chunks - 1) / chunks;
  return (B + *rpc - 1) / *rpc;
}

extern "C" {

int vcnf_gauss_fwd(const void* zin, const void* loc, const void* log_scale, int64_t j_stride, int64_t c_stride,
                   const int64_t* y, double log_temp, const void* accum, void* zout, void* logp, int64_t B, int n,
                   int mode, int dtype, void* stream) {
  if (B == 0) return 0;
  if (!zin || !loc || !log_scale || !logp || B < 0 || n <= 0 || mode < 0 || mode > 1 || (mode == 1 && !zout) ||
      (dtype != VCNF_F32 && dtype != VCNF_F64))
    return VCNF_E_BADARG;
  const int warp_rows = n < 256;
  const int sms = sm_count() > 0 ? sm_count() : VCNF_SM_COUNT_DEFAULT;
  int64_t blocks = warp_rows ? (B + 7) / 8 : B;
  if (blocks > static_cast<int64_t>(sms) * 8) blocks = static_cast<int64_t>(sms) * 8;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
#define GO(T)                                                                                                         \
  base::gauss_fwd_kernel<T><<<static_cast<unsigned>(blocks), base::kThreads, 0, st>>>(                                \
      static_cast<const T*>(zin), static_cast<const T*>(loc), static_cast<const T*>(log_scale), j_stride, c_stride, y, \
      static_cast<T>(log_temp), static_cast<const T*>(accum), static_cast<T*>(zout), static_cast<T*>(logp), B, n, mode, \
      warp_rows)
  if (dtype == VCNF_F32) GO(float); else GO(double);
#undef GO
  return launch_status();
}

int64_t vcnf_gauss_bwd_workspace(int n, int classes, int dtype) {
  return static_cast<int64_t>(base::kChunks) * 2 * classes * n * (dtype == VCNF_F64 ? 8 : 4);
}

int vcnf_gauss_bwd(const void* zin, const void* loc, const void* log_scale, int64_t j_stride, int64_t c_stride,
                   const int64_t* y, double log_temp, const void* g_logp, const void* g_zout, void* gz, void* g_loc,
                   void* g_log_scale, void* workspace, int64_t B, int n, int classes, int mode, int dtype, void* stream) {
  if (!zin || !loc || !log_scale || B < 0 || n <= 0 || classes <= 0 || mode < 0 || mode > 1 ||
      ((g_loc == nullptr) != (g_log_scale == nullptr)) || (g_loc && !workspace) ||
      (dtype != VCNF_F32 && dtype != VCNF_F64))
    return VCNF_E_BADARG;
  const size_t eb = dtype == VCNF_F64 ? 8 : 4;
  const int64_t tsize = static_cast<int64_t>(classes) * n;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (B == 0) {
    if (g_loc) { cudaMemsetAsync(g_loc, 0, tsize * eb, st); cudaMemsetAsync(g_log_scale, 0, tsize * eb, st); }
    return launch_status();
  }
  const int sms = sm_count() > 0 ? sm_count() : VCNF_SM_COUNT_DEFAULT;
  if (!g_loc) {
    // input gradient only: one flat elementwise launch
    if (mode != 0 || !gz) return 0;
    if (!g_logp) { cudaMemsetAsync(gz, 0, static_cast<size_t>(B) * n * eb, st); return launch_status(); }
    const int64_t total = B * n;
    int64_t blocks = (total + base::kThreads - 1) / base::kThreads;
    if (blocks > static_cast<int64_t>(sms) * 16) blocks = static_cast<int64_t>(sms) * 16;
    if (dtype == VCNF_F32)
      base::gauss_gz_kernel<float><<<static_cast<unsigned>(blocks), base::kThreads, 0, st>>>(
          static_cast<const float*>(zin), static_cast<const float*>(loc), static_cast<const float*>(log_scale), j_stride,
          c_stride, y, static_cast<float>(log_temp), static_cast<const float*>(g_logp), static_cast<float*>(gz), total, n);
    else
      base::gauss_gz_kernel<double><<<static_cast<unsigned>(blocks), base::kThreads, 0, st>>>(
          static_cast<const double*>(zin), static_cast<const double*>(loc), static_cast<const double*>(log_scale),
          j_stride, c_stride, y, log_temp, static_cast<const double*>(g_logp), static_cast<double*>(gz), total, n);
    return launch_status();
  }
  // row groups: as many as fit 256 threads and 200 KB of shared table
  int G = n >= base::kThreads ? 1 : base::kThreads / n;
  while (G > 1 && static_cast<size_t>(2 * tsize) * G * eb > 200 * 1024) G >>= 1;
  if (G > 32) G = 32;
  const size_t smem = static_cast<size_t>(2 * tsize) * G * eb;
  if (smem > 200 * 1024) return VCNF_E_UNSUPPORTED;
  int64_t rpc = 0;
  const int64_t chunks = plan_chunks(B, &rpc, 4 * G);
#define GO(T)                                                                                                          \
  {                                                                                                                    \
    static bool attr = false;                                                                                          \
    if (!attr) { cudaFuncSetAttribute(base::gauss_bwd_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024); attr = true; } \
    base::gauss_bwd_kernel<T><<<static_cast<unsigned>(chunks), base::kThreads, smem, st>>>(                             \
        static_cast<const T*>(zin), static_cast<const T*>(loc), static_cast<const T*>(log_scale), j_stride, c_stride, y, \
        static_cast<T>(log_temp), static_cast<const T*>(g_logp), static_cast<const T*>(g_zout), static_cast<T*>(gz),   \
        static_cast<T*>(workspace), B, n, classes, mode, rpc, G);                                                      \
    int rc = launch_status();                                                                                          \
    if (rc) return rc;                                                                                                 \
    /* partial layout [chunk][2][classes][n]: the two halves are the dense [classes][n] gradients */                    \
    const T* pw = static_cast<const T*>(workspace);                                                                    \
    launch_warp_reduce<T>(pw, static_cast<int>(chunks), 2 * tsize, tsize, static_cast<T*>(g_loc), st);                 \
    launch_warp_reduce<T>(pw + tsize, static_cast<int>(chunks), 2 * tsize, tsize, static_cast<T*>(g_log_scale), st);   \
  }
  if (dtype == VCNF_F32) GO(float) else GO(double)
#undef GO
  return launch_status();
}

static size_t gmm_smem(int K, int d, int bwd, size_t eb) {
  return (static_cast<size_t>(K) + base::kGmmRows * K + base::kGmmRows * d + base::kGmmRows + 8 +
          static_cast<size_t>(K) * d + K + (bwd ? static_cast<size_t>(K) * d * 2 + K : 0)) * eb;
}

int64_t vcnf_gmm_bwd_workspace(int K, int d, int dtype) {
  return static_cast<int64_t>(base::kChunks + 1) * (static_cast<int64_t>(K) * d * 2 + K) * (dtype == VCNF_F64 ? 8 : 4);
}

int vcnf_gmm_logprob(const void* z, const void* mu, const void* log_scale, const void* weight_scores,
                     const void* accum, void* logp, int64_t B, int K, int d, int dtype, void* stream) {
  if (B == 0) return 0;
  if (!z || !mu || !log_scale || !weight_scores || !logp || B < 0 || K <= 0 || d <= 0 ||
      (dtype != VCNF_F32 && dtype != VCNF_F64))
    return VCNF_E_BADARG;
  const size_t eb = dtype == VCNF_F64 ? 8 : 4;
  const size_t smem = gmm_smem(K, d, 0, eb);
  if (smem > 200 * 1024) return VCNF_E_UNSUPPORTED;
  int64_t rpc = 0;
  int64_t chunks = (B + base::kGmmRows - 1) / base::kGmmRows;
  const int sms = sm_count() > 0 ? sm_count() : VCNF_SM_COUNT_DEFAULT;
  if (chunks > static_cast<int64_t>(sms) * 4) chunks = static_cast<int64_t>(sms) * 4;
  rpc = ((B + chunks - 1) / chunks + base::kGmmRows - 1) / base::kGmmRows * base::kGmmRows;
  chunks = (B + rpc - 1) / rpc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
#define GO(T)                                                                                                          \
  {                                                                                                                    \
    static bool attr = false;                                                                                          \
    if (!attr) { cudaFuncSetAttribute(base::gmm_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024); attr = true; } \
    base::gmm_kernel<T><<<static_cast<unsigned>(chunks), base::kThreads, smem, st>>>(                                   \
        static_cast<const T*>(z), static_cast<const T*>(mu), static_cast<const T*>(log_scale),                         \
        static_cast<const T*>(weight_scores), static_cast<const T*>(accum), static_cast<T*>(logp), nullptr, nullptr,   \
        nullptr, B, K, d, rpc, 0);                                                                                     \
  }
  if (dtype == VCNF_F32) GO(float) else GO(double)
#undef GO
  return launch_status();
}

int vcnf_gmm_logprob_bwd(const void* z, const void* mu, const void* log_scale, const void* weight_scores,
                         const void* g_logp, void* gz, void* g_params, void* workspace, int64_t B, int K, int d,
                         int dtype, void* stream) {
  if (!mu || !log_scale || !weight_scores || !g_params || !workspace || B < 0 || K <= 0 || d <= 0 ||
      (dtype != VCNF_F32 && dtype != VCNF_F64))
    return VCNF_E_BADARG;
  const size_t eb = dtype == VCNF_F64 ? 8 : 4;
  const int64_t nacc = static_cast<int64_t>(K) * d * 2 + K;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (B == 0) {
    cudaMemsetAsync(g_params, 0, nacc * eb, st);
    return launch_status();
  }
  if (!z) return VCNF_E_BADARG;
  const size_t smem = gmm_smem(K, d, 1, eb);
  if (smem > 200 * 1024) return VCNF_E_UNSUPPORTED;
  int64_t chunks = (B + base::kGmmRows - 1) / base::kGmmRows;
  if (chunks > base::kChunks) chunks = base::kChunks;
  int64_t rpc = ((B + chunks - 1) / chunks + base::kGmmRows - 1) / base::kGmmRows * base::kGmmRows;
  chunks = (B + rpc - 1) / rpc;
#define GO(T)                                                                                                          \
  {                                                                                                                    \
    static bool attr = false;                                                                                          \
    if (!attr) { cudaFuncSetAttribute(base::gmm_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024); attr = true; } \
    base::gmm_kernel<T><<<static_cast<unsigned>(chunks), base::kThreads, smem, st>>>(                                   \
        static_cast<const T*>(z), static_cast<const T*>(mu), static_cast<const T*>(log_scale),                         \
        static_cast<const T*>(weight_scores), nullptr, nullptr, static_cast<const T*>(g_logp), static_cast<T*>(gz),    \
        static_cast<T*>(workspace), B, K, d, rpc, 1);                                                                  \
    int rc = launch_status();                                                                                          \
    if (rc) return rc;                                                                                                 \
    launch_warp_reduce<T>(static_cast<const T*>(workspace), static_cast<int>(chunks), nacc, nacc,                       \
                          static_cast<T*>(g_params), st);                                                              \
  }
  if (dtype == VCNF_F32) GO(float) else GO(double)
#undef GO
  return launch_status();
}

}  // extern "C"
