// Variational dequantisation of the categorical columns (reference: normflow/nets/varquant.py:22-83 driven per column
// from core.py:42-54) as ONE launch per direction for all columns.
//
// On a [B, 1] column every layer of the dequantiser's flow stack -- ActNorm / AffineConstFlow (coupling.py:37-53) and
// MaskedAffineFlow with a one-element mask (coupling.py:202-222: b = 1 passes through, b = 0 makes the conditioner
// input b*z identically zero, so scale and shift are the constants s(0), t(0)) -- is an elementwise affine map
// v <- v * exp(a_l) + c_l with log-det a_l.  The host evaluates the (a_l, c_l) table [n_cols, L, 2] from the
// parameters (under autograd); the kernels do everything that touches the batch:
//   u ~ U(0,1) fp64 (given) -> logit(u (1-alpha) + alpha/2) + log-det -> L affine layers -> sigmoid + log-det ->
//   (x + w) / quants, log-det -log(quants) -> logit(z (1-alpha) + alpha/2) + log-det,
// with the per-column log-dets accumulated the way core.py:46-53 does: the running accumulator is handed to every
// dequantiser, mutated in place and added again, so column k's log-det counts (n - k) times (SURVEY.md hazard A.2),
// in fp32 (hazard A.3).  All arithmetic on the noise chain is fp64 like the reference's (varquant.py:70).
// HBM-bound: 8 n B bytes of noise + 2 n s B column bytes + 4 B; the table sits in shared memory.
#include "common.cuh"
#include "../../include/vcnf_b200.h"

namespace vcnf {
namespace vdeq {

constexpr int kThreads = 128;
constexpr int kMaxL = 64;
constexpr int kBwdChunks = 148;

__device__ __forceinline__ double softplus_t(double x) { return x > 20.0 ? x : log1p(exp(x)); }   // torch threshold 20

template <typename XT>
__global__ void __launch_bounds__(kThreads) vdeq_fwd_kernel(const XT* __restrict__ x, int64_t ldx,
                                                            const double* __restrict__ noise,
                                                            const double* __restrict__ tab,
                                                            const double* __restrict__ quants, double alpha, int64_t B,
                                                            int n, int L, XT* __restrict__ zc, float* __restrict__ logq_tot,
                                                            float* __restrict__ logq_run) {
  extern __shared__ double s_tab[];     // [n][L][2] then exp(a) [n][L]
  double* s_ea = s_tab + static_cast<size_t>(n) * L * 2;
  for (int i = threadIdx.x; i < n * L * 2; i += blockDim.x) s_tab[i] = tab[i];
  __syncthreads();
  for (int i = threadIdx.x; i < n * L; i += blockDim.x) s_ea[i] = exp(s_tab[2 * i]);
  __syncthreads();
  const double l1a = log(1.0 - alpha);
  for (int64_t b = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; b < B;
       b += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    float run = 0.f, tot = 0.f;
    for (int k = 0; k < n; ++k) {
      const double q = quants[k];
      const double xf = static_cast<double>(static_cast<float>(x[b * ldx + k]));   // z.to(torch.float32), varquant.py:65
      const double up = noise[b * n + k] * (1.0 - alpha) + 0.5 * alpha;
      const double lu = log(up), l1u = log(1.0 - up);
      double ld = l1a - lu - l1u;
      double v = lu - l1u;
      for (int l = 0; l < L; ++l) {
        v = v * s_ea[k * L + l] + s_tab[(k * L + l) * 2 + 1];
        ld += s_tab[(k * L + l) * 2];
      }
      ld += -v - 2.0 * softplus_t(-v);
      const double w = 1.0 / (1.0 + exp(-v));
      const double z = (xf + w) / q;
      ld -= log(q);
      const double zp = z * (1.0 - alpha) + 0.5 * alpha;
      const double lz = log(zp), l1z = log(1.0 - zp);
      ld += l1a - lz - l1z;
      zc[b * n + k] = static_cast<XT>(lz - l1z);
      run += static_cast<float>(ld);      // the accumulator every dequantiser mutates in place (fp32, core.py:39-40)
      tot += run;                         // log_qt += log_det with log_det aliasing that accumulator (core.py:47-52)
    }
    logq_tot[b] = tot;
    if (logq_run) logq_run[b] = run;
  }
}

// Backward: gradients of the table.  grid = (chunks, n): one column per blockIdx.y, thread-private accumulators,
// fixed-order block reduction, per-chunk partials [chunks][n][L][2] finished by vdeq_reduce_kernel.
template <typename XT>
__global__ void __launch_bounds__(kThreads) vdeq_bwd_kernel(const XT* __restrict__ x, int64_t ldx,
                                                            const double* __restrict__ noise,
                                                            const double* __restrict__ tab,
                                                            const double* __restrict__ quants, double alpha, int64_t B,
                                                            int n, int L, const XT* __restrict__ g_zc,
                                                            const float* __restrict__ g_logq, int64_t rows_per_chunk,
                                                            double* __restrict__ partial) {
  // thread-private columns of three [L][kThreads] shared tables (accumulators of d/da, d/dc and the layer inputs of
  // the current row): indexed arrays in registers would live in local memory (measured: 510 MB of spill traffic)
  extern __shared__ double s_dyn[];
  double* s_ga = s_dyn;
  double* s_gc = s_ga + static_cast<size_t>(L) * kThreads;
  double* s_vs = s_gc + static_cast<size_t>(L) * kThreads;
  __shared__ double s_a[kMaxL], s_c[kMaxL], s_ea[kMaxL];
  __shared__ double red[kThreads / 32][2];
  const int k = blockIdx.y, chunk = blockIdx.x, tid = threadIdx.x;
  for (int l = tid; l < L; l += blockDim.x) {
    s_a[l] = tab[(k * L + l) * 2];
    s_c[l] = tab[(k * L + l) * 2 + 1];
    s_ea[l] = exp(s_a[l]);
  }
  for (int l = 0; l < L; ++l) { s_ga[l * kThreads + tid] = 0.0; s_gc[l * kThreads + tid] = 0.0; }
  __syncthreads();
  const double q = quants[k];
  const int64_t b0 = chunk * rows_per_chunk;
  const int64_t b1 = b0 + rows_per_chunk < B ? b0 + rows_per_chunk : B;
  for (int64_t b = b0 + tid; b < b1; b += blockDim.x) {
    const double g_out = g_zc ? static_cast<double>(g_zc[b * n + k]) : 0.0;
    const double g_ld = g_logq ? static_cast<double>(g_logq[b]) * (n - k) : 0.0;
    const double xf = static_cast<double>(static_cast<float>(x[b * ldx + k]));
    const double up = noise[b * n + k] * (1.0 - alpha) + 0.5 * alpha;
    double v = log(up) - log(1.0 - up);
    for (int l = 0; l < L; ++l) {
      s_vs[l * kThreads + tid] = v;
      v = v * s_ea[l] + s_c[l];
    }
    const double w = 1.0 / (1.0 + exp(-v));
    const double zp = (xf + w) / q * (1.0 - alpha) + 0.5 * alpha;
    const double g_zp = g_out * (1.0 / zp + 1.0 / (1.0 - zp)) + g_ld * (-1.0 / zp + 1.0 / (1.0 - zp));
    const double g_w = g_zp * (1.0 - alpha) / q;
    double g_v = g_w * w * (1.0 - w) + g_ld * (1.0 - 2.0 * w);
    for (int l = L - 1; l >= 0; --l) {
      s_gc[l * kThreads + tid] += g_v;
      s_ga[l * kThreads + tid] += g_v * s_vs[l * kThreads + tid] * s_ea[l] + g_ld;
      g_v *= s_ea[l];
    }
  }
  const int warp = tid >> 5, lane = tid & 31;
  for (int l = 0; l < L; ++l) {
    const double sa = warp_sum(s_ga[l * kThreads + tid]), sc = warp_sum(s_gc[l * kThreads + tid]);
    if (lane == 0) { red[warp][0] = sa; red[warp][1] = sc; }
    __syncthreads();
    if (tid == 0) {
      double ta = 0.0, tc = 0.0;
      for (int w2 = 0; w2 < kThreads / 32; ++w2) { ta += red[w2][0]; tc += red[w2][1]; }
      double* dst = partial + ((static_cast<int64_t>(chunk) * n + k) * L + l) * 2;
      dst[0] = ta; dst[1] = tc;
    }
    __syncthreads();
  }
}

__global__ void vdeq_reduce_kernel(const double* __restrict__ partial, int chunks, int total, double* __restrict__ g_tab) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  double acc = 0.0;
  for (int c = 0; c < chunks; ++c) acc += partial[static_cast<int64_t>(c) * total + i];
  g_tab[i] = acc;
}

// Sampling direction of Dequantization.forward (varquant.py:26-30) for all columns of a row:
//   ldj += sum(-z - 2 softplus(-z)); z = sigmoid(z) * quants; ldj += log(quants); z = clamp(floor(z), 0, quants - 1)
// with core.py:160-164's accumulation: the accumulator handed in IS the returned log-det, so `log_q += log_det`
// doubles it after every column.
template <typename T>
__global__ void __launch_bounds__(kThreads) vdeq_decode_kernel(const T* __restrict__ z, int64_t ldz,
                                                               const double* __restrict__ quants, int64_t B, int n,
                                                               T* __restrict__ out, T* __restrict__ logq) {
  for (int64_t b = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; b < B;
       b += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    T lq = logq[b];
    for (int k = 0; k < n; ++k) {
      const T v = z[b * ldz + k];
      const T q = static_cast<T>(quants[k]);
      const T sp = -v > T(20) ? -v : log1p(exp(-v));
      lq += (-v - T(2) * sp);
      T y = (T(1) / (T(1) + exp(-v))) * q;
      lq += static_cast<T>(log(quants[k]));
      y = floor(y);
      y = y < T(0) ? T(0) : (y > q - T(1) ? q - T(1) : y);
      out[b * n + k] = y;
      lq += lq;
    }
    logq[b] = lq;
  }
}

}  // namespace vdeq
}  // namespace vcnf

using namespace vcnf;

extern "C" {

int vcnf_vdeq_fwd(const void* x, int64_t ldx, const void* noise, const void* tab, const void* quants, double alpha,
                  int64_t B, int n, int L, void* zc, void* logq_tot, void* logq_run, int dtype, void* stream) {
  if (B == 0) return 0;
  if (!x || !noise || !quants || !zc || !logq_tot || B < 0 || n <= 0 || L < 0 || (L > 0 && !tab) || ldx < n ||
      (dtype != VCNF_F32 && dtype != VCNF_F64))
    return VCNF_E_BADARG;
  if (L > vdeq::kMaxL) return VCNF_E_UNSUPPORTED;
  const size_t smem = static_cast<size_t>(n) * (L > 0 ? L : 1) * 3 * sizeof(double);
  if (smem > 48 * 1024) return VCNF_E_UNSUPPORTED;
  const int sms = sm_count() > 0 ? sm_count() : VCNF_SM_COUNT_DEFAULT;
  int64_t blocks = (B + vdeq::kThreads - 1) / vdeq::kThreads;
  if (blocks > sms * 8) blocks = sms * 8;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == VCNF_F32)
    vdeq::vdeq_fwd_kernel<float><<<static_cast<unsigned>(blocks), vdeq::kThreads, smem, st>>>(
        static_cast<const float*>(x), ldx, static_cast<const double*>(noise), static_cast<const double*>(tab),
        static_cast<const double*>(quants), alpha, B, n, L, static_cast<float*>(zc), static_cast<float*>(logq_tot),
        static_cast<float*>(logq_run));
  else
    vdeq::vdeq_fwd_kernel<double><<<static_cast<unsigned>(blocks), vdeq::kThreads, smem, st>>>(
        static_cast<const double*>(x), ldx, static_cast<const double*>(noise), static_cast<const double*>(tab),
        static_cast<const double*>(quants), alpha, B, n, L, static_cast<double*>(zc), static_cast<float*>(logq_tot),
        static_cast<float*>(logq_run));
  return launch_status();
}

int64_t vcnf_vdeq_bwd_workspace(int n, int L) {
  return static_cast<int64_t>(vdeq::kBwdChunks) * n * (L > 0 ? L : 1) * 2 * 8;
}

int vcnf_vdeq_bwd(const void* x, int64_t ldx, const void* noise, const void* tab, const void* quants, double alpha,
                  int64_t B, int n, int L, const void* g_zc, const void* g_logq, void* g_tab, void* workspace,
                  int dtype, void* stream) {
  if (L == 0) return 0;
  if (!g_tab || !workspace || n <= 0 || L < 0 || B < 0 || (dtype != VCNF_F32 && dtype != VCNF_F64)) return VCNF_E_BADARG;
  if (L > vdeq::kMaxL || n > 65535) return VCNF_E_UNSUPPORTED;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int total = n * L * 2;
  if (B == 0) {
    cudaMemsetAsync(g_tab, 0, static_cast<size_t>(total) * 8, st);
    return launch_status();
  }
  if (!x || !noise || !tab || !quants || ldx < n) return VCNF_E_BADARG;
  int64_t chunks = (B + vdeq::kThreads - 1) / vdeq::kThreads;
  if (chunks > vdeq::kBwdChunks) chunks = vdeq::kBwdChunks;
  const int64_t rpc = (B + chunks - 1) / chunks;
  chunks = (B + rpc - 1) / rpc;
  dim3 grid(static_cast<unsigned>(chunks), n);
  const size_t dsm = static_cast<size_t>(3) * L * vdeq::kThreads * sizeof(double);      // <= 196 KB at L = 64
  static bool attr = false;
  if (!attr) {
    cudaFuncSetAttribute(vdeq::vdeq_bwd_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaFuncSetAttribute(vdeq::vdeq_bwd_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    attr = true;
  }
  if (dtype == VCNF_F32)
    vdeq::vdeq_bwd_kernel<float><<<grid, vdeq::kThreads, dsm, st>>>(
        static_cast<const float*>(x), ldx, static_cast<const double*>(noise), static_cast<const double*>(tab),
        static_cast<const double*>(quants), alpha, B, n, L, static_cast<const float*>(g_zc),
        static_cast<const float*>(g_logq), rpc, static_cast<double*>(workspace));
  else
    vdeq::vdeq_bwd_kernel<double><<<grid, vdeq::kThreads, dsm, st>>>(
        static_cast<const double*>(x), ldx, static_cast<const double*>(noise), static_cast<const double*>(tab),
        static_cast<const double*>(quants), alpha, B, n, L, static_cast<const double*>(g_zc),
        static_cast<const float*>(g_logq), rpc, static_cast<double*>(workspace));
  int rc = launch_status();
  if (rc) return rc;
  launch_warp_reduce<double>(static_cast<const double*>(workspace), static_cast<int>(chunks), total, total,
                             static_cast<double*>(g_tab), st);
  return launch_status();
}

int vcnf_vdeq_decode(const void* z, int64_t ldz, const void* quants, int64_t B, int n, void* out, void* logq, int dtype,
                     void* stream) {
  if (B == 0) return 0;
  if (!z || !quants || !out || !logq || B < 0 || n <= 0 || ldz < n || (dtype != VCNF_F32 && dtype != VCNF_F64))
    return VCNF_E_BADARG;
  const int sms = sm_count() > 0 ? sm_count() : VCNF_SM_COUNT_DEFAULT;
  int64_t blocks = (B + vdeq::kThreads - 1) / vdeq::kThreads;
  if (blocks > sms * 8) blocks = sms * 8;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == VCNF_F32)
    vdeq::vdeq_decode_kernel<float><<<static_cast<unsigned>(blocks), vdeq::kThreads, 0, st>>>(
        static_cast<const float*>(z), ldz, static_cast<const double*>(quants), B, n, static_cast<float*>(out),
        static_cast<float*>(logq));
  else
    vdeq::vdeq_decode_kernel<double><<<static_cast<unsigned>(blocks), vdeq::kThreads, 0, st>>>(
        static_cast<const double*>(z), ldz, static_cast<const double*>(quants), B, n, static_cast<double*>(out),
        static_cast<double*>(logq));
  return launch_status();
}

}  // extern "C"
