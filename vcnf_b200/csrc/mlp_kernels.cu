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
constexpr int64_t kMlpQuadMaxRows = 8192;   // at or below: the four-lanes-per-sample kernels (see mlp_fwd_quad_kernel)

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
  // Every global load of every layer is issued before the first store (blockDim.x == kMlpThreads): with a load -> store
  // pair per loop iteration the prologue was 8 x nl serialised L2 round trips -- 15 of the 15.6 us of a 1 024-row
  // forward launch, and the same again in the backward (cfg 1 / cfg 4 steps are chains of such launches).
  constexpr int kPer = kMlpW * kMlpW / kMlpThreads;
  T w[kMlpMaxLayers][kPer];
  T bv[kMlpMaxLayers];
#pragma unroll
  for (int l = 0; l < kMlpMaxLayers; ++l) {
    bv[l] = T(0);
#pragma unroll
    for (int k = 0; k < kPer; ++k) w[l][k] = T(0);
    if (l < a.nl) {
      const int din = a.dims[l], dout = a.dims[l + 1];
#pragma unroll
      for (int k = 0; k < kPer; ++k) {
        const int e = threadIdx.x + k * kMlpThreads, o = e / kMlpW, i = e - o * kMlpW;
        if (o < dout && i < din) w[l][k] = __ldg(a.W[l] + o * din + i);
      }
      if (static_cast<int>(threadIdx.x) < dout && a.b[l]) bv[l] = __ldg(a.b[l] + threadIdx.x);
    }
  }
#pragma unroll
  for (int l = 0; l < kMlpMaxLayers; ++l) {
    if (l < a.nl) {
#pragma unroll
      for (int k = 0; k < kPer; ++k) {
        const int e = threadIdx.x + k * kMlpThreads, o = e / kMlpW, i = e - o * kMlpW;
        if (Ws) Ws[(l * kMlpW + o) * kMlpW + i] = w[l][k];
        if (WsT) WsT[(l * kMlpW + i) * kMlpW + o] = w[l][k];
      }
      if (threadIdx.x < kMlpW) bs[l * kMlpW + threadIdx.x] = bv[l];
    }
  }
}

// out[o] = act(bias[o] + sum_i W[o][i] h[i]) for o < dout (zeros above); fully unrolled so that the vectors
// stay in registers.
template <typename T>
__device__ __forceinline__ void mlp_layer(const T* W, const T* bias, int din, int dout, const T* h, T* out, bool act,
                                          T slope) {
  // Two outputs at a time, two accumulator chains each (four independent FMA chains), inputs in groups of 8 up to
  // din: a launch on a few thousand rows is one serial pass per thread, so chain length and padded work are what it
  // costs (13.7 us for [1, 32, 32, 2] on 1 024 rows with 32-wide loops and two chains).
#pragma unroll
  for (int o = 0; o < kMlpW; o += 2) {
    T r0 = T(0), r1 = T(0);
    if (o < dout) {
      const T* w0 = W + o * kMlpW;
      const T* w1 = w0 + kMlpW;
      T a0 = bias[o], a1 = T(0), c0 = bias[o + 1], c1 = T(0);
#pragma unroll
      for (int ib = 0; ib < kMlpW; ib += 8) {
        if (ib < din) {
#pragma unroll
          for (int i = ib; i < ib + 8; i += 2) {
            a0 += w0[i] * h[i]; a1 += w0[i + 1] * h[i + 1];
            c0 += w1[i] * h[i]; c1 += w1[i + 1] * h[i + 1];
          }
        }
      }
      r0 = a0 + a1;
      r1 = c0 + c1;                                    // row o + 1 >= dout is all zeros (weights and bias): r1 = 0
      if (act) {
        r0 = r0 > T(0) ? r0 : r0 * slope;
        r1 = r1 > T(0) ? r1 : r1 * slope;
      }
    }
    out[o] = r0;
    out[o + 1] = r1;
  }
}

template <typename T>
__global__ void __launch_bounds__(kMlpThreads) mlp_fwd_kernel(const MlpArgs<T> a) {
  extern __shared__ __align__(16) unsigned char mlp_smem[];
  T* Ws = reinterpret_cast<T*>(mlp_smem);
  T* bs = Ws + kMlpMaxLayers * kMlpW * kMlpW;
  __shared__ int sdims[kMlpMaxLayers + 1];
  if (threadIdx.x <= kMlpMaxLayers) sdims[threadIdx.x] = a.dims[threadIdx.x];
  mlp_load_weights(a, Ws, bs, static_cast<T*>(nullptr));
  __syncthreads();
  const int nl = a.nl, din = sdims[0], dlast = sdims[nl];
  for (int64_t r = blockIdx.x * static_cast<int64_t>(kMlpThreads) + threadIdx.x; r < a.B;
       r += static_cast<int64_t>(gridDim.x) * kMlpThreads) {
    T h[kMlpW], t[kMlpW];
    const T* xr = a.x + r * a.ldx;
#pragma unroll
    for (int i = 0; i < kMlpW; ++i) h[i] = i < din ? xr[i] : T(0);
    // ONE copy of the layer code, looped: unrolled over the layers the kernel was 13 700 instructions that every warp
    // executes once -- instruction fetch (ncu: no_instruction 10.6 of 13 stall cycles per issue), not the FMA pipe,
    // set its time (26 us for 1 024 rows).
#pragma unroll 1
    for (int l = 0; l < nl; ++l) {
      mlp_layer(Ws + l * kMlpW * kMlpW, bs + l * kMlpW, sdims[l], sdims[l + 1], h, t, l + 1 < nl, a.slope);
#pragma unroll
      for (int i = 0; i < kMlpW; ++i) h[i] = t[i];
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

// Shared-memory elements of the backward kernel (host and device agree through this function).
__host__ __device__ inline size_t mlp_bwd_elems(int nl, int psize) {
  constexpr int kWarps = kMlpThreads / 32;
  return 2 * static_cast<size_t>(nl) * kMlpW * kMlpW + static_cast<size_t>(nl) * kMlpW +
         kWarps * (kMlpW * kMlpW + kMlpW * (kMlpW + 1)) + static_cast<size_t>(kWarps) * psize +
         static_cast<size_t>(nl > 1 ? nl - 1 : 0) * kMlpW * kMlpThreads;
}

template <typename T>
__global__ void __launch_bounds__(kMlpThreads) mlp_bwd_kernel(const MlpArgs<T> a) {
  extern __shared__ __align__(16) unsigned char mlp_smem[];
  constexpr int kWarps = kMlpThreads / 32;
  const int nl = a.nl;
  T* Ws = reinterpret_cast<T*>(mlp_smem);                         // [nl][32][32]
  T* WsT = Ws + nl * kMlpW * kMlpW;                               // [nl][32][32]
  T* bs = WsT + nl * kMlpW * kMlpW;                               // [nl][32]
  T* tiles = bs + nl * kMlpW;                                     // per warp: gT[32][32] | at[32][33]
  T* accs = tiles + kWarps * (kMlpW * kMlpW + kMlpW * (kMlpW + 1));   // per warp: psize accumulators
  T* acts = accs + kWarps * a.psize;                              // [nl - 1][32][threads]: inputs of layers 1 .. nl-1
  __shared__ int sdims[kMlpMaxLayers + 1], soff[kMlpMaxLayers];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  T* gT = tiles + warp * (kMlpW * kMlpW + kMlpW * (kMlpW + 1));
  T* at = gT + kMlpW * kMlpW;
  T* acc = accs + warp * a.psize;
  if (tid == 0) {
    int p = 0;
    for (int l = 0; l <= kMlpMaxLayers; ++l) sdims[l] = a.dims[l];
    for (int l = 0; l < kMlpMaxLayers; ++l) { soff[l] = p; if (l < nl) p += a.dims[l + 1] * a.dims[l] + a.dims[l + 1]; }
  }
  mlp_load_weights(a, Ws, bs, WsT);
  for (int p = tid; p < kWarps * a.psize; p += kMlpThreads) accs[p] = T(0);
  __syncthreads();
  const int din0 = sdims[0], dlast = sdims[nl];
  const int64_t n_tiles = (a.B + kMlpThreads - 1) / kMlpThreads;
  // The layer loops are real loops over ONE copy of the layer code (see mlp_fwd_kernel: unrolled, the kernel was
  // instruction-fetch bound); the hidden activations therefore live in shared memory ([layer][unit][thread]:
  // conflict-free), not in a register array indexed by the layer.
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int64_t r = tile * kMlpThreads + tid;
    const bool valid = r < a.B;
    const T* xr = a.x + (valid ? r : 0) * a.ldx;
    {
      T h[kMlpW], t[kMlpW];
#pragma unroll
      for (int i = 0; i < kMlpW; ++i) h[i] = (valid && i < din0) ? xr[i] : T(0);
#pragma unroll 1
      for (int l = 0; l + 1 < nl; ++l) {                          // forward recompute up to the last layer's input
        mlp_layer(Ws + l * kMlpW * kMlpW, bs + l * kMlpW, sdims[l], sdims[l + 1], h, t, true, a.slope);
        T* al = acts + (static_cast<size_t>(l) * kMlpW) * kMlpThreads + tid;
#pragma unroll
        for (int i = 0; i < kMlpW; ++i) { h[i] = t[i]; al[i * kMlpThreads] = t[i]; }
      }
    }
    T g[kMlpW];
    {
      const T* gr = a.gy + (valid ? r : 0) * a.ldgy;
#pragma unroll
      for (int o = 0; o < kMlpW; ++o) g[o] = (valid && o < dlast) ? gr[o] : T(0);
    }
#pragma unroll 1
    for (int l = nl - 1; l >= 0; --l) {
      const int din = sdims[l], dout = sdims[l + 1];
      const T* al = acts + (static_cast<size_t>(l > 0 ? l - 1 : 0) * kMlpW) * kMlpThreads + tid;   // input of layer l (l > 0)
      // park this warp's (g, a_l) pairs: gT[o][s], at[s][i]
#pragma unroll
      for (int o = 0; o < kMlpW; ++o) if (o < dout) gT[o * kMlpW + lane] = g[o];
      if (l == 0) {
        for (int i = 0; i < din; ++i) at[lane * (kMlpW + 1) + i] = valid ? xr[i] : T(0);
      } else {
#pragma unroll
        for (int i = 0; i < kMlpW; ++i) if (i < din) at[lane * (kMlpW + 1) + i] = al[i * kMlpThreads];
      }
      __syncwarp();
      if (lane < din) {                                    // column `lane` of gW_l = g^T a over the warp's 32 samples
        T col[kMlpW];
#pragma unroll
        for (int s = 0; s < kMlpW; ++s) col[s] = at[s * (kMlpW + 1) + lane];
        T* aw = acc + soff[l];
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
        acc[soff[l] + dout * din + lane] += d0;
      }
      __syncwarp();
      if (l > 0 || a.gx) {                                 // input gradient of the layer, then the activation's slope
        T gp[kMlpW];
        const T* wt = WsT + l * kMlpW * kMlpW;
#pragma unroll
        for (int i = 0; i < kMlpW; i += 2) {
          T v0 = T(0), v1 = T(0);
          if (i < din) {
            const T* w0 = wt + i * kMlpW;
            const T* w1 = w0 + kMlpW;                      // row i + 1 >= din of WsT is zero
            T a0 = T(0), a1 = T(0), c0 = T(0), c1 = T(0);
#pragma unroll
            for (int ob = 0; ob < kMlpW; ob += 8) {
              if (ob < dout) {
#pragma unroll
                for (int o = ob; o < ob + 8; o += 2) {
                  a0 += w0[o] * g[o]; a1 += w0[o + 1] * g[o + 1];
                  c0 += w1[o] * g[o]; c1 += w1[o + 1] * g[o + 1];
                }
              }
            }
            v0 = a0 + a1;
            v1 = c0 + c1;
            if (l > 0) {
              v0 *= al[i * kMlpThreads] > T(0) ? T(1) : a.slope;
              v1 *= al[(i + 1) * kMlpThreads] > T(0) ? T(1) : a.slope;
            }
          }
          gp[i] = v0;
          gp[i + 1] = v1;
        }
#pragma unroll
        for (int i = 0; i < kMlpW; ++i) g[i] = gp[i];
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

// ---------------------------------------------------------------------------------------------------
// Small batches (cfg 1 trains on 1 024 rows, cfg 4 on 2 048): with one thread per sample a launch is ONE serial pass
// of ~4 000 dependent instructions per thread on a handful of SMs (9 us forward, 27 us backward whatever the batch).
// The quad kernels give every sample FOUR lanes: lane q of a quad owns outputs 8q .. 8q+7 of every layer (a quarter
// of the FMAs and weight loads) and the full activation vector is re-assembled inside the quad with shuffles.
// Same per-output summation order as mlp_layer, so both families agree bit for bit.
constexpr int kQuadSamples = kMlpThreads / 4;          // samples per CTA tile

// out8[j] = act(bias[8q + j] + sum_i Wt[i][8q + j] h[i]), j = 0..7, i < din rounded up to 8.  Wt is the TRANSPOSED
// weight ([input][output], 32 x 32): the four lanes of a quad read 32 consecutive elements of row i (conflict-free, the
// eight quads of the warp read the same words), where rows of the untransposed matrix 8 apart share their banks.
template <typename T>
__device__ __forceinline__ void quad_layer(const T* Wt, const T* bias, int din, int dout, int q, const T* h, T* out8,
                                           bool act, T slope) {
  T a0[8], a1[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { a0[j] = T(0); a1[j] = T(0); }
  if (q * 8 < dout) {
#pragma unroll
    for (int j = 0; j < 8; ++j) a0[j] = bias[q * 8 + j];
#pragma unroll
    for (int ib = 0; ib < kMlpW; ib += 8) {
      if (ib < din) {
#pragma unroll
        for (int i = ib; i < ib + 8; i += 2) {
          const T* r0 = Wt + i * kMlpW + q * 8;
          const T* r1 = r0 + kMlpW;
#pragma unroll
          for (int j = 0; j < 8; ++j) { a0[j] += r0[j] * h[i]; a1[j] += r1[j] * h[i + 1]; }
        }
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    T r = a0[j] + a1[j];
    if (act) r = r > T(0) ? r : r * slope;
    out8[j] = r;
  }
}

// full[8 qq + j] = the value out8[j] of lane qq of this quad (all four lanes end with the same 32-vector).
template <typename T>
__device__ __forceinline__ void quad_gather(const T* out8, T* full, int lane) {
  const int base = lane & ~3;
#pragma unroll
  for (int qq = 0; qq < 4; ++qq) {
#pragma unroll
    for (int j = 0; j < 8; ++j) full[qq * 8 + j] = __shfl_sync(0xffffffffu, out8[j], base + qq);
  }
}

template <typename T>
__global__ void __launch_bounds__(kMlpThreads) mlp_fwd_quad_kernel(const MlpArgs<T> a) {
  extern __shared__ __align__(16) unsigned char mlp_smem[];
  T* WsT = reinterpret_cast<T*>(mlp_smem);              // [layer][input][output]
  T* bs = WsT + kMlpMaxLayers * kMlpW * kMlpW;
  __shared__ int sdims[kMlpMaxLayers + 1];
  if (threadIdx.x <= kMlpMaxLayers) sdims[threadIdx.x] = a.dims[threadIdx.x];
  mlp_load_weights(a, static_cast<T*>(nullptr), bs, WsT);
  __syncthreads();
  const int nl = a.nl, din = sdims[0], dlast = sdims[nl];
  const int lane = threadIdx.x & 31, q = lane & 3;
  const int64_t n_tiles = (a.B + kQuadSamples - 1) / kQuadSamples;
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int64_t r = tile * kQuadSamples + (threadIdx.x >> 2);
    const bool valid = r < a.B;
    T h[kMlpW], o8[8];
    const T* xr = a.x + (valid ? r : 0) * a.ldx;
#pragma unroll
    for (int i = 0; i < kMlpW; ++i) h[i] = (valid && i < din) ? xr[i] : T(0);
#pragma unroll 1
    for (int l = 0; l < nl; ++l) {
      quad_layer(WsT + l * kMlpW * kMlpW, bs + l * kMlpW, sdims[l], sdims[l + 1], q, h, o8, l + 1 < nl, a.slope);
      quad_gather(o8, h, lane);
    }
    if (valid) {
      T* yr = a.y + r * a.ldy;
#pragma unroll
      for (int j = 0; j < 8; ++j) if (q * 8 + j < dlast) yr[q * 8 + j] = o8[j];
    }
  }
}

// Backward with four lanes per sample (see mlp_fwd_quad_kernel).  A warp holds 8 samples: the forward is recomputed
// (hidden activations of the tile in shared memory, [layer][unit][sample]), the input gradient of a layer is the same
// quad_layer on the transposed weights, and the weight gradients are formed per warp from the parked (g, a) pairs of
// its 8 samples (lane i accumulates column i of g^T a).  Per-warp sums are combined per CTA, per-CTA partials by
// mlp_partial_reduce_kernel: fixed order, bit-reproducible.
__host__ __device__ inline size_t mlp_bwd_quad_elems(int nl, int psize) {
  constexpr int kWarps = kMlpThreads / 32;
  return 2 * static_cast<size_t>(nl) * kMlpW * kMlpW + static_cast<size_t>(nl) * kMlpW + kMlpW +
         kWarps * (kMlpW * 9 + 8 * (kMlpW + 1)) + static_cast<size_t>(kWarps) * psize +
         static_cast<size_t>(nl > 1 ? nl - 1 : 0) * kQuadSamples * (kMlpW + 1);
}

template <typename T>
__global__ void __launch_bounds__(kMlpThreads) mlp_bwd_quad_kernel(const MlpArgs<T> a) {
  extern __shared__ __align__(16) unsigned char mlp_smem[];
  constexpr int kWarps = kMlpThreads / 32;
  constexpr int kTile = kMlpW * 9 + 8 * (kMlpW + 1);   // gT[32][9] | at[8][33]: odd strides keep the quad's stores apart
  constexpr int kAct = kMlpW + 1;                      // acts[layer][sample][33]
  const int nl = a.nl;
  T* Ws = reinterpret_cast<T*>(mlp_smem);                         // [nl][32][32]
  T* WsT = Ws + nl * kMlpW * kMlpW;                               // [nl][32][32]
  T* bs = WsT + nl * kMlpW * kMlpW;                               // [nl][32]
  T* zb = bs + nl * kMlpW;                                        // [32] zeros (bias of the transposed products)
  T* tiles = zb + kMlpW;                                          // per warp: gT[32][9] | at[8][33]
  T* accs = tiles + kWarps * kTile;                               // per warp: psize accumulators
  T* acts = accs + kWarps * a.psize;                              // [nl - 1][kQuadSamples][33]
  __shared__ int sdims[kMlpMaxLayers + 1], soff[kMlpMaxLayers];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, q = lane & 3, s8 = lane >> 2;
  const int sc = tid >> 2;                                        // sample within the CTA tile
  T* gT = tiles + warp * kTile;
  T* at = gT + kMlpW * 9;
  T* acc = accs + warp * a.psize;
  if (tid == 0) {
    int p = 0;
    for (int l = 0; l <= kMlpMaxLayers; ++l) sdims[l] = a.dims[l];
    for (int l = 0; l < kMlpMaxLayers; ++l) { soff[l] = p; if (l < nl) p += a.dims[l + 1] * a.dims[l] + a.dims[l + 1]; }
  }
  if (tid < kMlpW) zb[tid] = T(0);
  mlp_load_weights(a, Ws, bs, WsT);
  for (int p = tid; p < kWarps * a.psize; p += kMlpThreads) accs[p] = T(0);
  __syncthreads();
  const int din0 = sdims[0], dlast = sdims[nl];
  const int64_t n_tiles = (a.B + kQuadSamples - 1) / kQuadSamples;
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int64_t r = tile * kQuadSamples + sc;
    const bool valid = r < a.B;
    const T* xr = a.x + (valid ? r : 0) * a.ldx;
    T h[kMlpW], o8[8];
#pragma unroll
    for (int i = 0; i < kMlpW; ++i) h[i] = (valid && i < din0) ? xr[i] : T(0);
#pragma unroll 1
    for (int l = 0; l + 1 < nl; ++l) {                            // forward recompute up to the last layer's input
      quad_layer(WsT + l * kMlpW * kMlpW, bs + l * kMlpW, sdims[l], sdims[l + 1], q, h, o8, true, a.slope);
      T* al = acts + (static_cast<size_t>(l) * kQuadSamples + sc) * kAct + q * 8;
#pragma unroll
      for (int j = 0; j < 8; ++j) al[j] = o8[j];
      quad_gather(o8, h, lane);
    }
    T g[kMlpW];
    {
      const T* gr = a.gy + (valid ? r : 0) * a.ldgy;
#pragma unroll
      for (int o = 0; o < kMlpW; ++o) g[o] = (valid && o < dlast) ? gr[o] : T(0);
    }
#pragma unroll 1
    for (int l = nl - 1; l >= 0; --l) {
      const int din = sdims[l], dout = sdims[l + 1];
      const T* al = acts + (static_cast<size_t>(l > 0 ? l - 1 : 0) * kQuadSamples + sc) * kAct;   // input of layer l
      // park this warp's (g, a_l) pairs: gT[o][s], at[s][i]; lane q of a quad writes entries 8q .. 8q+7
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int e = q * 8 + j;
        gT[e * 9 + s8] = g[e];
        T av = T(0);
        if (e < din) av = l == 0 ? (valid ? xr[e] : T(0)) : al[e];
        at[s8 * (kMlpW + 1) + e] = av;
      }
      __syncwarp();
      if (lane < din) {                                    // column `lane` of gW_l = g^T a over the warp's 8 samples
        T col[8];
#pragma unroll
        for (int s = 0; s < 8; ++s) col[s] = at[s * (kMlpW + 1) + lane];
        T* aw = acc + soff[l];
        for (int o = 0; o < dout; ++o) {
          const T* grow = gT + o * 9;
          T d0 = T(0), d1 = T(0);
#pragma unroll
          for (int s = 0; s < 8; s += 2) { d0 += grow[s] * col[s]; d1 += grow[s + 1] * col[s + 1]; }
          aw[o * din + lane] += d0 + d1;
        }
      }
      if (lane < dout) {
        const T* grow = gT + lane * 9;
        T d0 = T(0);
#pragma unroll
        for (int s = 0; s < 8; ++s) d0 += grow[s];
        acc[soff[l] + dout * din + lane] += d0;
      }
      __syncwarp();
      if (l > 0 || a.gx) {                                 // input gradient of the layer, then the activation's slope
        quad_layer(Ws + l * kMlpW * kMlpW, zb, dout, din, q, g, o8, false, a.slope);   // W is the transpose of W^T
        if (l > 0) {
#pragma unroll
          for (int j = 0; j < 8; ++j) o8[j] *= al[q * 8 + j] > T(0) ? T(1) : a.slope;
        }
        quad_gather(o8, g, lane);
      }
    }
    if (a.gx && valid) {
      T* gxr = a.gx + r * a.ldgx;
#pragma unroll
      for (int j = 0; j < 8; ++j) if (q * 8 + j < din0) gxr[q * 8 + j] = g[q * 8 + j];
    }
    __syncwarp();
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
  const int64_t cap = static_cast<int64_t>(sms) * 8;
  if (a.B <= kMlpQuadMaxRows) {                        // few rows: four lanes per sample (latency, not throughput)
    const int64_t qt = (a.B + kQuadSamples - 1) / kQuadSamples;
    mlp_fwd_quad_kernel<T><<<static_cast<unsigned>(qt < cap ? qt : cap), kMlpThreads, smem, st>>>(a);
    return launch_status();
  }
  const int64_t tiles = (a.B + kMlpThreads - 1) / kMlpThreads;
  mlp_fwd_kernel<T><<<static_cast<unsigned>(tiles < cap ? tiles : cap), kMlpThreads, smem, st>>>(a);
  return launch_status();
}

template <typename T>
static int launch_mlp_bwd(MlpArgs<T>& a, T* gparams, cudaStream_t st) {
  if (a.B <= kMlpQuadMaxRows) {                        // few rows: four lanes per sample
    const size_t qs = mlp_bwd_quad_elems(a.nl, a.psize) * sizeof(T);
    if (qs <= 224 * 1024) {
      auto qk = mlp_bwd_quad_kernel<T>;
      static bool qattr = false;
      if (!qattr) {
        cudaFuncSetAttribute(qk, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024);
        qattr = true;
      }
      const int64_t qt = (a.B + kQuadSamples - 1) / kQuadSamples;
      const int64_t qg = qt < kMlpMaxGrid ? qt : kMlpMaxGrid;
      qk<<<static_cast<unsigned>(qg), kMlpThreads, qs, st>>>(a);
      int qrc = launch_status();
      if (qrc) return qrc;
      mlp_partial_reduce_kernel<T><<<(a.psize + 3) / 4, 128, 0, st>>>(a.partial, static_cast<int>(qg), a.psize, gparams);
      return launch_status();
    }
  }
  const size_t smem = mlp_bwd_elems(a.nl, a.psize) * sizeof(T);
  if (smem > 224 * 1024) return VCNF_E_UNSUPPORTED;
  auto kern = mlp_bwd_kernel<T>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024);
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
