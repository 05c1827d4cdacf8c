// Autoregressive layers in the SAMPLING direction (reference: normflow/flows/affine/autoregressive.py:29-36: D full
// conditioner passes, each followed by the elementwise inverse; SURVEY.md a15 / K7 "sampling: 64x everything").
//
// Pass t can only fix the feature of degree t + 1.  Its parameters depend on hidden units of degree <= t, and a
// hidden unit of degree d depends on inputs of degree <= d alone (nets/made.py:56-75: hidden mask deg_out >= deg_in,
// output mask deg_out > deg_in).  So the hidden units of degree t become final in pass t and never change again: ONE
// evaluation of MADE in total is enough, spread over the D passes as "bands" of a few units per layer.  This kernel
// runs all D passes of one layer for a tile of 16 samples on chip:
//   * the hidden vectors of the tile (initial layer, both layers of each residual block) live in shared memory,
//     laid out [unit][sample] so a warp (lane = sample) reads them conflict-free;
//   * pass t evaluates, layer after layer, only the units of degree t, then the P output rows of feature order[t],
//     then the elementwise inverse of that one feature (RQ spline or the MAF affine map) and its log-det; the
//     weights of a pass are two contiguous host-packed blocks that a 1-D bulk copy (TMA engine) brings into shared
//     memory one pass ahead (the band block travels during the previous pass's output rows and vice versa): with
//     the weights read straight from L2 the kernel was bound by that latency, 8 warps per SM cannot hide it;
//   * each dot product is split over the 8 warps (segments of the contraction index) and summed in a fixed order.
// Rows are independent, so there is no communication between CTAs; weights stream from L2 (~60 KB per pass).
// Results equal the reference's D-pass loop up to summation order (tests: 1e-5 / 1e-4).
#include "common.cuh"
#include "rqs_math.cuh"
#include "rqs_fast.cuh"
#include "affine_math.cuh"
#include "../../include/vcnf_b200.h"

namespace vcnf {
namespace made {

constexpr int kThreads = 512;
constexpr int kWarps = 16;
constexpr int kR = 16;          // samples per CTA; lane = (sample r = lane & 15, half = lane >> 4: alternate j)
constexpr int kBU = 8;          // band units per degree, padded (sequential degrees give ceil(H / (D - 1)) <= 8)
constexpr int kMaxLayers = 5;   // initial layer + 2 layers per residual block, <= 2 blocks
constexpr int kPP = 32;         // output rows per feature, padded to a multiple of 4 (3K + 1 <= 31 for K <= 10; MAF: 2)

struct Args {
  const float* y; float* x; float* logdet; int64_t B;
  int D, H, P, Pp, n_layers;              // n_layers = 1 + 2 * blocks; Pp = P rounded up to a multiple of 4
  // per-pass operand blocks, each one contiguous span so that ONE bulk copy brings a pass's weights on chip:
  //   band[t]  = [Wb_0[t] (D x 8) | Wb_1[t] (H x 8) | ... | bb_0[t] (8) | bb_1[t] (8) | ...]   masked weights /
  //              biases of the hidden units of degree t (8 slots per degree), layer after layer
  //   final[t] = [Wf[t] (H x Pp) | bf[t] (Pp)]   the final layer's rows of feature order[t], transposed
  const float* band; int band_floats;
  const float* fin; int fin_floats;
  const int32_t* band_idx;                // [D][kBU] hidden unit index of band slot u, -1 = padding
  const int32_t* order;                   // [D] feature fixed in pass t
  const float* bounds;                    // [D] tail bound per feature (spline map)
  const uint8_t* modes;                   // [D] DerivMode per feature (spline map)
  int K, map, outside_zero;               // map 0: MAF affine, 1: RQ spline
  int32_t* err_flag;
};

// dst[unit][r] = bias + sum_j W[j][u] * act(src[j][r])  (+ residual[unit][r]) for the 8 slots of the current band.
// The contraction is split over the 8 warps (segments of j) and inside a warp over its two half-warps (alternate
// j): a warp reads 32 consecutive floats of `src` per step; the weights are 32-byte rows in shared memory.
template <bool RELU, int NB>
__device__ __forceinline__ void band_layer(const float* __restrict__ src, int Hin, const float* __restrict__ W,
                                           const float* __restrict__ bias, const int32_t* __restrict__ idx,
                                           float* __restrict__ dst, const float* __restrict__ res,
                                           float* __restrict__ part, int warp, int lane) {
  const int seg = Hin / kWarps, r = lane & 15, half = lane >> 4;
  float acc[NB];
#pragma unroll
  for (int u = 0; u < NB; ++u) acc[u] = 0.f;
#pragma unroll 4
  for (int i = half; i < seg; i += 2) {
    const int j = warp * seg + i;
    float hv = src[j * kR + r];
    if (RELU) hv = fmaxf(hv, 0.f);
    const float4 w0 = *reinterpret_cast<const float4*>(W + j * kBU);
    acc[0] += w0.x * hv; acc[1] += w0.y * hv; acc[2] += w0.z * hv; acc[3] += w0.w * hv;
    if (NB == 6) {                                  // slots 6, 7 are padding in every band: skip their products
      const float2 w1 = *reinterpret_cast<const float2*>(W + j * kBU + 4);
      acc[4] += w1.x * hv; acc[5] += w1.y * hv;
    } else {
      const float4 w1 = *reinterpret_cast<const float4*>(W + j * kBU + 4);
      acc[4] += w1.x * hv; acc[5] += w1.y * hv; acc[NB - 2] += w1.z * hv; acc[NB - 1] += w1.w * hv;
    }
  }
#pragma unroll
  for (int u = 0; u < NB; ++u) acc[u] += __shfl_xor_sync(0xffffffffu, acc[u], 16);
  if (half == 0) {
#pragma unroll
    for (int u = 0; u < NB; ++u) part[(warp * kBU + u) * kR + r] = acc[u];
  }
  __syncthreads();
  if (threadIdx.x < NB * kR) {
    const int u = threadIdx.x >> 4, rr = threadIdx.x & 15;
    const int unit = __ldg(idx + u);
    if (unit >= 0) {
      float s = bias[u];
#pragma unroll
      for (int w = 0; w < kWarps; ++w) s += part[(w * kBU + u) * kR + rr];
      if (res) s += res[unit * kR + rr];
      dst[unit * kR + rr] = s;
    }
  }
  __syncthreads();
}

// KMAX: spline bins the elementwise map is compiled for; PG: groups of four output rows per feature (ceil(P / 4));
// NB: band slots that can hold a unit (6 or 8).
template <int KMAX, int PG, int NB>
__global__ void __launch_bounds__(kThreads, 1) made_inverse_kernel(const Args a) {
  extern __shared__ __align__(128) float sm[];
  const int D = a.D, H = a.H, P = a.P;
  constexpr int Pp = PG * 4;
  float* bbuf = sm;                                 // band operands of the current pass   (bulk-copy target)
  float* fbuf = bbuf + a.band_floats;               // final-layer operands of the current pass
  float* xs = fbuf + a.fin_floats;                  // [D][kR]   features fixed so far (zeros before)
  float* ys = xs + D * kR;                          // [D][kR]   the layer's input tile
  float* hb = ys + D * kR;                          // [n_layers][H][kR]
  float* part = hb + static_cast<size_t>(a.n_layers) * H * kR;      // [kWarps][kPP][kR]
  float* prm = part + kWarps * kPP * kR;            // [kPP][kR]
  __shared__ uint64_t bars[2];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    mbar_init_fence();
  }
  __syncthreads();
  uint32_t pb = 0, pf = 0;                          // phase parities of the two barriers
  int flag = 0;
  for (int64_t r0 = static_cast<int64_t>(blockIdx.x) * kR; r0 < a.B; r0 += static_cast<int64_t>(gridDim.x) * kR) {
    const int rows = static_cast<int>(a.B - r0 < kR ? a.B - r0 : kR);
    if (tid == 0) {
      // pass 0 needs only the final-layer block; the band block of pass 1 travels meanwhile
      fence_async_smem();
      mbar_expect_tx(&bars[1], static_cast<uint32_t>(a.fin_floats) * 4);
      bulk_g2s(fbuf, a.fin, static_cast<uint32_t>(a.fin_floats) * 4, &bars[1]);
      if (D > 1) {
        mbar_expect_tx(&bars[0], static_cast<uint32_t>(a.band_floats) * 4);
        bulk_g2s(bbuf, a.band + static_cast<size_t>(a.band_floats), static_cast<uint32_t>(a.band_floats) * 4, &bars[0]);
      }
    }
    for (int e = tid; e < D * kR; e += kThreads) {
      const int r = e / D, c = e - r * D;           // coalesced read of the [rows, D] tile, transposed into smem
      ys[c * kR + r] = r < rows ? a.y[(r0 + r) * D + c] : 0.f;
      xs[c * kR + r] = 0.f;
    }
    for (int e = tid; e < a.n_layers * H * kR; e += kThreads) hb[e] = 0.f;
    __syncthreads();
    float ld_acc = 0.f;
    for (int t = 0; t < D; ++t) {
      if (t >= 1) {
        // the hidden units of degree t, layer after layer (a unit of degree t sees inputs / units of degree <= t)
        mbar_wait(&bars[0], pb);
        pb ^= 1;
        const int32_t* idx = a.band_idx + t * kBU;
        const float* bias = bbuf + (D + (a.n_layers - 1) * H) * kBU;
        band_layer<false, NB>(xs, D, bbuf, bias, idx, hb, nullptr, part, warp, lane);
        for (int l = 1; l < a.n_layers; l += 2) {
          float* hk = hb + static_cast<size_t>(l - 1) * H * kR;       // block input
          float* tk = hb + static_cast<size_t>(l) * H * kR;
          float* hn = hb + static_cast<size_t>(l + 1) * H * kR;
          const float* W0 = bbuf + (D + (l - 1) * H) * kBU;
          band_layer<true, NB>(hk, H, W0, bias + l * kBU, idx, tk, nullptr, part, warp, lane);
          band_layer<true, NB>(tk, H, W0 + H * kBU, bias + (l + 1) * kBU, idx, hn, hk, part, warp, lane);
        }
        // every thread is past its last read of bbuf (band_layer ends with a CTA barrier): fetch the next band block
        if (tid == 0 && t + 1 < D) {
          fence_async_smem();
          mbar_expect_tx(&bars[0], static_cast<uint32_t>(a.band_floats) * 4);
          bulk_g2s(bbuf, a.band + static_cast<size_t>(t + 1) * a.band_floats, static_cast<uint32_t>(a.band_floats) * 4,
                   &bars[0]);
        }
      }
      // the P parameters of the feature fixed in this pass (final masked layer, no activation: made.py:296-299)
      mbar_wait(&bars[1], pf);
      pf ^= 1;
      {
        const float* hl = hb + static_cast<size_t>(a.n_layers - 1) * H * kR;
        const int seg = H / kWarps, r = lane & 15, half = lane >> 4;
        float acc[PG * 4];
#pragma unroll
        for (int m = 0; m < PG * 4; ++m) acc[m] = 0.f;
        if (t >= 1) {
#pragma unroll 2
          for (int i = half; i < seg; i += 2) {
            const int j = warp * seg + i;
            const float hv = hl[j * kR + r];
            const float4* wp = reinterpret_cast<const float4*>(fbuf + static_cast<size_t>(j) * (PG * 4));
#pragma unroll
            for (int g = 0; g < PG; ++g) {
              const float4 w = wp[g];
              acc[g * 4] += w.x * hv; acc[g * 4 + 1] += w.y * hv; acc[g * 4 + 2] += w.z * hv; acc[g * 4 + 3] += w.w * hv;
            }
          }
        }
#pragma unroll
        for (int m = 0; m < PG * 4; ++m) {
          const float v = acc[m] + __shfl_xor_sync(0xffffffffu, acc[m], 16);
          if (half == 0) part[(warp * kPP + m) * kR + r] = v;
        }
        __syncthreads();
        const float* bf = fbuf + static_cast<size_t>(H) * Pp;
        for (int e = tid; e < P * kR; e += kThreads) {
          const int m = e >> 4, rr = e & 15;
          float s2 = bf[m];
#pragma unroll
          for (int w = 0; w < kWarps; ++w) s2 += part[(w * kPP + m) * kR + rr];
          prm[m * kR + rr] = s2;
        }
        __syncthreads();
        if (tid == 0 && t + 1 < D) {                // fbuf is free: the next pass's block travels during its bands
          fence_async_smem();
          mbar_expect_tx(&bars[1], static_cast<uint32_t>(a.fin_floats) * 4);
          bulk_g2s(fbuf, a.fin + static_cast<size_t>(t + 1) * a.fin_floats, static_cast<uint32_t>(a.fin_floats) * 4,
                   &bars[1]);
        }
      }
      // elementwise inverse of feature f (autoregressive.py:34-35 restricted to the one column this pass can fix)
      if (tid < kR) {
        const int f = __ldg(a.order + t);
        const float v = ys[f * kR + tid];
        float out, lad;
        if (a.map == 0) {
          affine_apply<float>(SCALE_MAF, 1, v, prm[1 * kR + tid], prm[0 * kR + tid], &out, &lad);
        } else {
          float p[kPP];
#pragma unroll
          for (int m = 0; m < kPP; ++m) p[m] = m < P ? prm[m * kR + tid] : 0.f;
          int bin;
          if (KMAX == 8 && a.K == 8) {
            // the packed-math element of the spline fast path (rqs_fast.cuh; what vcnf_rqs_layer_fwd runs for fp32,
            // K = 8, i.e. what the D-pass loop evaluates): a third of the generic element's instructions, and this
            // phase is serial -- 16 threads of the CTA while everything else waits
            fast::element<8, true>(p, v, a.bounds[f], 1.f, a.modes[f], a.outside_zero, &out, &lad, &bin, &flag);
          } else {
            RqsElemCfg c;
            c.K = a.K; c.mode = a.modes[f]; c.inverse = 1; c.outside_zero = a.outside_zero;
            rqs_element<float, KMAX>(p, v, a.bounds[f], 1.f, c, &out, &lad, &bin, &flag);
          }
        }
        xs[f * kR + tid] = out;
        ld_acc += lad;
      }
      __syncthreads();
    }
    for (int e = tid; e < rows * D; e += kThreads) {
      const int r = e / D, c = e - r * D;
      a.x[(r0 + r) * D + c] = xs[c * kR + r];
    }
    if (tid < rows) a.logdet[r0 + tid] = ld_acc;
    __syncthreads();
  }
  if (flag && a.err_flag) atomicOr(a.err_flag, flag);
}

template <int KMAX, int PG, int NB>
static void launch(const Args& a, unsigned grid, size_t smem, cudaStream_t st) {
  static bool attr = false;
  if (!attr) {
    cudaFuncSetAttribute(made_inverse_kernel<KMAX, PG, NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    attr = true;
  }
  made_inverse_kernel<KMAX, PG, NB><<<grid, kThreads, smem, st>>>(a);
}

}  // namespace made
}  // namespace vcnf

using namespace vcnf;

extern "C" int vcnf_made_inverse(const void* y, void* x, void* logdet, int64_t B, int D, int H, int P, int n_layers,
                                 const void* band, const void* fin, const int32_t* band_idx, const int32_t* order,
                                 int band_units, const void* bounds, const uint8_t* modes, int K, int map,
                                 int outside_zero, int32_t* err_flag, void* stream) {
  if (B == 0) return 0;
  if (!y || !x || !logdet || !band || !fin || !band_idx || !order || B < 0 || D <= 0 || H <= 0 || P <= 0 ||
      n_layers < 1 || (n_layers & 1) == 0 || band_units < 1 || band_units > made::kBU || map < 0 || map > 1 || (map == 1 && (!bounds || !modes)))
    return VCNF_E_BADARG;
  if (n_layers > made::kMaxLayers || P > made::kPP || D % made::kWarps != 0 || H % made::kWarps != 0 ||
      (map == 1 && (K < 1 || K > 16)) || (map == 0 && P != 2) || reinterpret_cast<uintptr_t>(band) % 16 != 0 ||
      reinterpret_cast<uintptr_t>(fin) % 16 != 0)
    return VCNF_E_UNSUPPORTED;
  made::Args a{};
  a.y = static_cast<const float*>(y); a.x = static_cast<float*>(x); a.logdet = static_cast<float*>(logdet); a.B = B;
  a.D = D; a.H = H; a.P = P; a.Pp = (P + 3) / 4 * 4; a.n_layers = n_layers;
  a.band = static_cast<const float*>(band);
  a.band_floats = (D + (n_layers - 1) * H) * made::kBU + n_layers * made::kBU;
  a.fin = static_cast<const float*>(fin);
  a.fin_floats = (H + 1) * a.Pp;
  a.band_idx = band_idx; a.order = order;
  a.bounds = static_cast<const float*>(bounds); a.modes = modes; a.K = K; a.map = map; a.outside_zero = outside_zero;
  a.err_flag = err_flag;
  const size_t smem = (static_cast<size_t>(a.band_floats) + a.fin_floats + static_cast<size_t>(2) * D * made::kR +
                       static_cast<size_t>(n_layers) * H * made::kR +
                       static_cast<size_t>(made::kWarps) * made::kPP * made::kR + made::kPP * made::kR) * sizeof(float);
  if (smem > 220 * 1024) return VCNF_E_UNSUPPORTED;
  const int sms = sm_count() > 0 ? sm_count() : VCNF_SM_COUNT_DEFAULT;
  const int64_t tiles = (B + made::kR - 1) / made::kR;
  const unsigned grid = static_cast<unsigned>(tiles < sms ? tiles : sms);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int pg = a.Pp / 4;
  const bool wide = map == 1 && K > 8;
  if (wide && pg < 7) return VCNF_E_BADARG;          // K >= 9 has P >= 26
  bool done = false;
#define VCNF_MADE_CASE(KM, PGV)                                                                                     \
  if (!done && wide == (KM == 16) && pg == PGV) {                                                                   \
    done = true;                                                                                                    \
    if (band_units <= 6) made::launch<KM, PGV, 6>(a, grid, smem, st);                                               \
    else made::launch<KM, PGV, 8>(a, grid, smem, st);                                                               \
  }
  VCNF_MADE_CASE(8, 1) VCNF_MADE_CASE(8, 2) VCNF_MADE_CASE(8, 3) VCNF_MADE_CASE(8, 4) VCNF_MADE_CASE(8, 5)
  VCNF_MADE_CASE(8, 6) VCNF_MADE_CASE(8, 7) VCNF_MADE_CASE(16, 7) VCNF_MADE_CASE(16, 8)
#undef VCNF_MADE_CASE
  if (!done) return VCNF_E_UNSUPPORTED;
  return launch_status();
}
