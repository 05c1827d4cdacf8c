// Small MLP conditioners evaluated in registers (sm_100a): normflow/nets/mlp.py:7-58, Linear + LeakyReLU(slope)
// ... Linear, every width <= 32 -- the [1, 32, 32, 2] net of the README Real NVP (cfg 1) and the [15, 30, 15] /
// [1, 4, 1] nets of the dequantisation flows (cfg 4).
//
// These layers are 1 120 MACs per sample on 20-byte rows: as library GEMMs every layer round-trips a [B, 32]
// activation through HBM (and a 128-wide tensor-core tile is 4-32x padding), so the whole net runs as ONE
// kernel on the FP32 / FP64 FMA pipe (~100 FLOP/B, above the SIMT ridge):
//   forward : one thread per sample, the activation vector lives in registers, weights are read from shared
//             memory as broadcast 128-bit words;
//   backward: the same thread recomputes the activations, back-propagates through the layers in registers, and
//             the weight gradients are formed per warp: the 32 samples of a warp park (g, a) in shared memory,
//             then lane j accumulates column j of g^T a over the 32 samples.  Per-warp sums are combined in a
//             fixed order per CTA, per-CTA partials by mlp_partial_reduce_kernel: bit-reproducible.
#include "common.cuh"
#include "../../include/vcnf_b200.h"

namespace vcnf {

constexpr int kMlpThreads = 128;
constexpr int kMlpW = 32;            // maximum layer width
constexpr int kMlpMaxLayers = 4;     // linear layers
constexpr int kMlpMaxGrid = 148 * 4;

template <typename T>
struct MlpArgs {
  const T* x; int64_t ldx;
  T* y; int64_t ldy;
  const T* W[kMlpMaxLayers]; const T* b[kMlpMaxLayers];
  int dims[kMlpMaxLayers + 1]; int nl; T slope; int64_t B;
  // backward
  const T* gy; int64_t ldgy; T* gx; int64_t ldgx; T* partial; int psize;
};

// Weights of layer l in shared memory: Ws[l][o][i] (i padded to 32 with zeros), biases bs[l][o].
template <typename T>
__device__ __forceinline__ void mlp_load_weights(const MlpArgs<T>& a, T* Ws, T* bs, T* WsT) {
  for (int l = 0; l < a.nl; ++l) {
    const int din = a.dims[l], dout = a.dims[l + 1];
    for (int e = threadIdx.x; e < kMlpW * kMlpW; e += blockDim.x) {
      const int o = e / kMlpW, i = e - o * kMlpW;
      const T w = (o < dout && i < din) ? a.W[l][o * din + i] : T(0);
      Ws[(l * kMlpW + o) * kMlpW + i] = w;
      if (WsT) WsT[(l * kMlpW + i) * kMlpW + o] = w;
    }
    for (int o = threadIdx.x; o < kMlpW; o += blockDim.x) bs[l * kMlpW + o] = (o < dout && a.b[l]) ? a.b[l][o] : T(0);
  }
}

// out[o] = act(bias[o] + sum_i W[o][i] h[i]) for o < dout (zeros above); fully unrolled so that the vectors
// stay in registers.
template <typename T>
__device__ __forceinline__ void mlp_layer(const T* W, const T* bias, int dout, const T* h, T* out, bool act, T slope) {
#pragma unroll
  for (int o = 0; o < kMlpW; ++o) {
    T acc = T(0);
    if (o < dout) {
      const T* w = W + o * kMlpW;
      T a0 = bias[o], a1 = T(0);
#pragma unroll
      for (int i = 0; i < kMlpW; i += 2) { a0 += w[i] * h[i]; a1 += w[i + 1] * h[i + 1]; }
      acc = a0 + a1;
      if (act) acc = acc > T(0) ? acc : acc * slope;
    }
    out[o] = acc;
  }
}

template <typename T>
__global__ void __launch_bounds__(kMlpThreads) mlp_fwd_kernel(const MlpArgs<T> a) {
  extern __shared__ __align__(16) unsigned char mlp_smem[];
  T* Ws = reinterpret_cast<T*>(mlp_smem);
  T* bs = Ws + kMlpMaxLayers * kMlpW * kMlpW;
  mlp_load_weights(a, Ws, bs, static_cast<T*>(nullptr));
  __syncthreads();
  const int din = a.dims[0], dlast = a.dims[a.nl];
  for (int64_t r = blockIdx.x * static_cast<int64_t>(kMlpThreads) + threadIdx.x; r < a.B;
       r += static_cast<int64_t>(gridDim.x) * kMlpThreads) {
    T h[kMlpW], t[kMlpW];
    const T* xr = a.x + r * a.ldx;
#pragma unroll
    for (int i = 0; i < kMlpW; ++i) h[i] = i < din ? xr[i] : T(0);
#pragma unroll
    for (int l = 0; l < kMlpMaxLayers; ++l) {
      if (l < a.nl) {
        mlp_layer(Ws + l * kMlpW * kMlpW, bs + l * kMlpW, a.dims[l + 1], h, t, l + 1 < a.nl, a.slope);
#pragma unroll
        for (int i = 0; i < kMlpW; ++i) h[i] = t[i];
      }
    }
    T* yr = a.y + r * a.ldy;
#pragma unroll
    for (int o = 0; o < kMlpW; ++o) if (o < dlast) yr[o] = h[o];
  }
}

// Parameter-gradient vector layout (also the layout of `partial` rows and of the reduce's output):
// for l = 0 .. nl-1: [gW_l (dout x din, row-major) | gb_l (dout)].
__host__ __device__ inline int mlp_psize(const int* dims, int nl) {
  int p = 0;
  for (int l = 0; l < nl; ++l) p += dims[l + 1] * dims[l] + dims[l + 1];
  return p;
}

template <typename T>
__global__ void __launch_bounds__(kMlpThreads) mlp_bwd_kernel(const MlpArgs<T> a) {
  extern __shared__ __align__(16) unsigned char mlp_smem[];
  constexpr int kWarps = kMlpThreads / 32;
  T* Ws = reinterpret_cast<T*>(mlp_smem);
  T* bs = Ws + kMlpMaxLayers * kMlpW * kMlpW;
  T* WsT = bs + kMlpMaxLayers * kMlpW;
  T* tiles = WsT + kMlpMaxLayers * kMlpW * kMlpW;                 // per warp: gT[32][32] | at[32][33]
  T* accs = tiles + kWarps * (kMlpW * kMlpW + kMlpW * (kMlpW + 1));   // per warp: psize accumulators
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  T* gT = tiles + warp * (kMlpW * kMlpW + kMlpW * (kMlpW + 1));
  T* at = gT + kMlpW * kMlpW;
  T* acc = accs + warp * a.psize;
  mlp_load_weights(a, Ws, bs, WsT);
  for (int p = tid; p < kWarps * a.psize; p += kMlpThreads) accs[p] = T(0);
  __syncthreads();
  const int nl = a.nl, din0 = a.dims[0], dlast = a.dims[nl];
  int off[kMlpMaxLayers];
  {
    int p = 0;
    for (int l = 0; l < kMlpMaxLayers; ++l) { off[l] = p; if (l < nl) p += a.dims[l + 1] * a.dims[l] + a.dims[l + 1]; }
  }
  const int64_t n_tiles = (a.B + kMlpThreads - 1) / kMlpThreads;
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int64_t r = tile * kMlpThreads + tid;
    const bool valid = r < a.B;
    // forward recompute: act[l] = input of layer l (act[0] = x is re-read when layer 0 needs it)
    T act[kMlpMaxLayers][kMlpW];
    {
      const T* xr = a.x + (valid ? r : 0) * a.ldx;
#pragma unroll
      for (int i = 0; i < kMlpW; ++i) act[1][i] = (valid && i < din0) ? xr[i] : T(0);      // staged in act[1]
    }
#pragma unroll
    for (int l = 0; l < kMlpMaxLayers - 1; ++l) {
      if (l + 1 < nl) {
        T t[kMlpW];
        mlp_layer(Ws + l * kMlpW * kMlpW, bs + l * kMlpW, a.dims[l + 1], l == 0 ? act[1] : act[l], t, true, a.slope);
#pragma unroll
        for (int i = 0; i < kMlpW; ++i) act[l + 1][i] = t[i];
      }
    }
    T g[kMlpW];
    {
      const T* gr = a.gy + (valid ? r : 0) * a.ldgy;
#pragma unroll
      for (int o = 0; o < kMlpW; ++o) g[o] = (valid && o < dlast) ? gr[o] : T(0);
    }
#pragma unroll
    for (int ll = 0; ll < kMlpMaxLayers; ++ll) {
      const int l = kMlpMaxLayers - 1 - ll;
      if (l < nl) {
        const int din = a.dims[l], dout = a.dims[l + 1];
        // park this warp's (g, a_l) pairs: gT[o][s], at[s][i]
#pragma unroll
        for (int o = 0; o < kMlpW; ++o) if (o < dout) gT[o * kMlpW + lane] = g[o];
        if (l == 0) {
          const T* xr = a.x + (valid ? r : 0) * a.ldx;
          for (int i = 0; i < din; ++i) at[lane * (kMlpW + 1) + i] = valid ? xr[i] : T(0);
        } else {
#pragma unroll
          for (int i = 0; i < kMlpW; ++i) if (i < din) at[lane * (kMlpW + 1) + i] = act[l][i];
        }
        __syncwarp();
        if (lane < din) {                                  // column `lane` of gW_l = g^T a over the warp's 32 samples
          T col[kMlpW];
#pragma unroll
          for (int s = 0; s < kMlpW; ++s) col[s] = at[s * (kMlpW + 1) + lane];
          T* aw = acc + off[l];
          for (int o = 0; o < dout; ++o) {
            const T* grow = gT + o * kMlpW;
            T d0 = T(0), d1 = T(0);
#pragma unroll
            for (int s = 0; s < kMlpW; s += 2) { d0 += grow[s] * col[s]; d1 += grow[s + 1] * col[s + 1]; }
            aw[o * din + lane] += d0 + d1;
          }
        }
        if (lane < dout) {
          const T* grow = gT + lane * kMlpW;
          T d0 = T(0);
#pragma unroll
          for (int s = 0; s < kMlpW; ++s) d0 += grow[s];
          acc[off[l] + dout * din + lane] += d0;
        }
        __syncwarp();
        if (l > 0 || a.gx) {                               // input gradient of the layer, then the activation's slope
          T gp[kMlpW];
          const T* wt = WsT + l * kMlpW * kMlpW;
#pragma unroll
          for (int i = 0; i < kMlpW; ++i) {
            T v = T(0);
            if (i < din) {
              const T* w = wt + i * kMlpW;
              T a0 = T(0), a1 = T(0);
#pragma unroll
              for (int o = 0; o < kMlpW; o += 2) { a0 += w[o] * g[o]; a1 += w[o + 1] * g[o + 1]; }
              v = a0 + a1;
              if (l > 0) v *= act[l][i] > T(0) ? T(1) : a.slope;
            }
            gp[i] = v;
          }
#pragma unroll
          for (int i = 0; i < kMlpW; ++i) g[i] = gp[i];
        }
      }
    }
    if (a.gx && valid) {
      T* gxr = a.gx + r * a.ldgx;
#pragma unroll
      for (int i = 0; i < kMlpW; ++i) if (i < din0) gxr[i] = g[i];
    }
  }
  __syncthreads();
  T* mine = a.partial + static_cast<size_t>(blockIdx.x) * a.psize;
  for (int p = tid; p < a.psize; p += kMlpThreads) {
    T s = T(0);
    for (int w = 0; w < kWarps; ++w) s += accs[w * a.psize + p];
    mine[p] = s;
  }
}

template <typename T>
__global__ void mlp_partial_reduce_kernel(const T* partial, int nblocks, int n, T* out) {
  const int j = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (j >= n) return;
  T acc = T(0);
  for (int b = lane; b < nblocks; b += 32) acc += partial[static_cast<size_t>(b) * n + j];
  acc = warp_sum(acc);
  if (lane == 0) out[j] = acc;
}

template <typename T>
static bool mlp_fill(MlpArgs<T>& a, const void* const* W, const void* const* b, const int32_t* dims, int nl, double slope) {
  if (nl < 1 || nl > kMlpMaxLayers || !W || !dims || slope < 0.0) return false;
  for (int l = 0; l <= nl; ++l) {
    if (dims[l] < 1 || dims[l] > kMlpW) return false;
    a.dims[l] = dims[l];
  }
  for (int l = nl + 1; l <= kMlpMaxLayers; ++l) a.dims[l] = 0;
  for (int l = 0; l < kMlpMaxLayers; ++l) {
    a.W[l] = l < nl ? static_cast<const T*>(W[l]) : nullptr;
    a.b[l] = (l < nl && b) ? static_cast<const T*>(b[l]) : nullptr;
    if (l < nl && !a.W[l]) return false;
  }
  a.nl = nl;
  a.slope = static_cast<T>(slope);
  return true;
}

template <typename T>
static int launch_mlp_fwd(MlpArgs<T>& a, cudaStream_t st) {
  const size_t smem = (static_cast<size_t>(kMlpMaxLayers) * kMlpW * kMlpW + kMlpMaxLayers * kMlpW) * sizeof(T);
  const int sms = sm_count() > 0 ? sm_count() : VCNF_SM_COUNT_DEFAULT;
  const int64_t tiles = (a.B + kMlpThreads - 1) / kMlpThreads;
  const int64_t cap = static_cast<int64_t>(sms) * 8;
  mlp_fwd_kernel<T><<<static_cast<unsigned>(tiles < cap ? tiles : cap), kMlpThreads, smem, st>>>(a);
  return launch_status();
}

template <typename T>
static int launch_mlp_bwd(MlpArgs<T>& a, T* gparams, cudaStream_t st) {
  constexpr int kWarps = kMlpThreads / 32;
  const size_t elems = 2 * static_cast<size_t>(kMlpMaxLayers) * kMlpW * kMlpW + kMlpMaxLayers * kMlpW +
                       kWarps * (kMlpW * kMlpW + kMlpW * (kMlpW + 1)) + static_cast<size_t>(kWarps) * a.psize;
  const size_t smem = elems * sizeof(T);
  if (smem > 200 * 1024) return VCNF_E_UNSUPPORTED;
  auto kern = mlp_bwd_kernel<T>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    attr_set = true;
  }
  const int sms = sm_count() > 0 ? sm_count() : VCNF_SM_COUNT_DEFAULT;
  const int64_t tiles = (a.B + kMlpThreads - 1) / kMlpThreads;
  int per_sm = static_cast<int>((227 * 1024) / (smem + 1024));
  per_sm = per_sm < 1 ? 1 : (per_sm > 4 ? 4 : per_sm);
  int64_t grid = static_cast<int64_t>(sms) * per_sm;
  if (grid > kMlpMaxGrid) grid = kMlpMaxGrid;
  if (grid > tiles) grid = tiles;
  kern<<<static_cast<unsigned>(grid), kMlpThreads, smem, st>>>(a);
  int rc = launch_status();
  if (rc) return rc;
  mlp_partial_reduce_kernel<T><<<(a.psize + 3) / 4, 128, 0, st>>>(a.partial, static_cast<int>(grid), a.psize, gparams);
  return launch_status();
}

}  // namespace vcnf

using namespace vcnf;

extern "C" int vcnf_mlp_fwd(const void* x, int64_t ldx, void* y, int64_t ldy, const void* const* W, const void* const* b,
                            const int32_t* dims, int nl, double slope, int64_t B, int dtype, void* stream) {
  if (B == 0) return 0;
  if (!x || !y || B < 0 || (dtype != VCNF_F32 && dtype != VCNF_F64)) return VCNF_E_BADARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == VCNF_F32) {
    MlpArgs<float> a{};
    if (!mlp_fill(a, W, b, dims, nl, slope)) return VCNF_E_BADARG;
    a.x = static_cast<const float*>(x); a.ldx = ldx; a.y = static_cast<float*>(y); a.ldy = ldy; a.B = B;
    return launch_mlp_fwd(a, st);
  }
  MlpArgs<double> a{};
  if (!mlp_fill(a, W, b, dims, nl, slope)) return VCNF_E_BADARG;
  a.x = static_cast<const double*>(x); a.ldx = ldx; a.y = static_cast<double*>(y); a.ldy = ldy; a.B = B;
  return launch_mlp_fwd(a, st);
}

extern "C" int64_t vcnf_mlp_bwd_workspace(const int32_t* dims, int nl, int dtype) {
  if (!dims || nl < 1 || nl > kMlpMaxLayers) return 0;
  int d[kMlpMaxLayers + 1];
  for (int l = 0; l <= nl; ++l) d[l] = dims[l];
  return static_cast<int64_t>(kMlpMaxGrid) * mlp_psize(d, nl) * (dtype == VCNF_F64 ? 8 : 4);
}

extern "C" int vcnf_mlp_bwd(const void* x, int64_t ldx, const void* gy, int64_t ldgy, const void* const* W,
                            const void* const* b, const int32_t* dims, int nl, double slope, void* gx, int64_t ldgx,
                            void* gparams, void* workspace, int64_t B, int dtype, void* stream) {
  if (!gparams || B < 0 || (dtype != VCNF_F32 && dtype != VCNF_F64)) return VCNF_E_BADARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
#define VCNF_MLP_BWD(T)                                                                                   \
  {                                                                                                       \
    MlpArgs<T> a{};                                                                                       \
    if (!mlp_fill(a, W, b, dims, nl, slope)) return VCNF_E_BADARG;                                        \
    a.psize = mlp_psize(a.dims, nl);                                                                      \
    if (B == 0) { cudaMemsetAsync(gparams, 0, sizeof(T) * a.psize, st); return launch_status(); }         \
    if (!x || !gy || !workspace) return VCNF_E_BADARG;                                                    \
    a.x = static_cast<const T*>(x); a.ldx = ldx; a.gy = static_cast<const T*>(gy); a.ldgy = ldgy;         \
    a.gx = static_cast<T*>(gx); a.ldgx = ldgx; a.partial = static_cast<T*>(workspace); a.B = B;           \
    return launch_mlp_bwd(a, static_cast<T*>(gparams), st);                                               \
  }
  if (dtype == VCNF_F32) VCNF_MLP_BWD(float) else VCNF_MLP_BWD(double)
#undef VCNF_MLP_BWD
}
