// Small dense per-sample linear maps (D <= 64) for sm_100a:
//   y[r, o] = sum_i W[o, i] * x[r, in_idx[i]] + bias[o]      (written to column out_idx[o])
// used by LULinearPermute (normflow/flows/mixing.py:398-429,484-492: permutation + L*U matvec,
// or the explicit inverse in the sampling direction) and Invertible1x1Conv (mixing.py:101-128:
// a C x C matvec per pixel).  A "row" is a sample (S == 1, row-major [B, D]) or a pixel of an
// NCHW tensor (S = H*W, feature stride S).
//
// These are D x D contractions with D in {2, 12, 24, 29, 48, 64}: far too small for a tensor-core
// tile per sample and HBM-bound at s*(2D) bytes per row, so they stay on the FP32/FP64 FMA pipe:
// a CTA stages a tile of rows in shared memory (coalesced), W lives in shared memory padded to DP
// columns and is read as broadcast 128-bit words, each thread keeps its row in registers.
#include "common.cuh"
#include "../../include/vcnf_b200.h"

namespace vcnf {

constexpr int kLinRows = 128;   // rows per tile = threads per CTA

template <typename T>
struct LinArgs {
  const T* x; T* y; const T* W; const T* bias;
  const int32_t* in_idx; const int32_t* out_idx;
  int64_t rows; int D; int S; int sub_bias_first;   // sub_bias_first: y = W (x - bias) (sampling dir)
};

template <typename T, int DP>
__global__ void __launch_bounds__(kLinRows) small_linear_kernel(const LinArgs<T> a) {
  extern __shared__ __align__(16) unsigned char smem_lin[];
  T* Ws = reinterpret_cast<T*>(smem_lin);              // [D][DP], zero padded
  T* bs = Ws + static_cast<size_t>(a.D) * DP;          // [DP]
  T* xs = bs + DP;                                     // [kLinRows][D + 1]
  const int D = a.D, ldt = D + 1;
  __shared__ int idx[2 * DP];                          // input / output feature maps (identity when absent)
  for (int e = threadIdx.x; e < DP; e += blockDim.x) {
    idx[e] = e < D ? (a.in_idx ? a.in_idx[e] : e) : 0;
    idx[DP + e] = e < D ? (a.out_idx ? a.out_idx[e] : e) : 0;
  }
#pragma unroll 8
  for (int e = threadIdx.x; e < D * DP; e += blockDim.x) {            // (unrolled: 8 independent loads in flight)
    const int o = e / DP, i = e - o * DP;
    Ws[e] = i < D ? a.W[o * D + i] : T(0);
  }
  for (int e = threadIdx.x; e < DP; e += blockDim.x) bs[e] = (e < D && a.bias) ? a.bias[e] : T(0);
  const int64_t n_tiles = (a.rows + kLinRows - 1) / kLinRows;
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int64_t r0 = tile * kLinRows;
    const int nr = static_cast<int>(a.rows - r0 < kLinRows ? a.rows - r0 : kLinRows);
    __syncthreads();
    if (a.S != 1) {
      // image layout [B, D, S]: thread r owns pixel r0 + r.  Consecutive threads hold consecutive s, so each of the D
      // strided loads / stores is one coalesced access of the warp: no shared-memory staging, no index arithmetic per
      // element (the staged version spent 46 us per launch in dependent division + load chains at 16 CTAs).
      if (static_cast<int>(threadIdx.x) < nr) {
        const int64_t row = r0 + threadIdx.x;
        const int64_t b = row / a.S, sp = row - b * a.S;
        const T* px = a.x + b * D * a.S + sp;
        T* py = a.y + b * D * a.S + sp;
        T xr[DP];
        // all D loads first, nothing that consumes them in between (with the bias subtraction inside this loop the
        // compiler kept load -> use pairs: 48 serialised L2 round trips, 34 us per CTA under ncu)
#pragma unroll
        for (int i = 0; i < DP; ++i) xr[i] = i < D ? __ldg(px + static_cast<int64_t>(idx[i]) * a.S) : T(0);
        if (a.sub_bias_first) {
#pragma unroll
          for (int i = 0; i < DP; ++i) xr[i] -= bs[i];
        }
        for (int o = 0; o < D; ++o) {
          T acc0 = a.sub_bias_first ? T(0) : bs[o], acc1 = T(0);
          const float4* w = reinterpret_cast<const float4*>(Ws + o * DP);   // rows are 16 B aligned (DP % 4 == 0)
          if (sizeof(T) == 4) {
#pragma unroll
            for (int i = 0; i < DP / 4; ++i) {
              const float4 wv = w[i];
              acc0 += static_cast<T>(wv.x) * xr[4 * i] + static_cast<T>(wv.z) * xr[4 * i + 2];
              acc1 += static_cast<T>(wv.y) * xr[4 * i + 1] + static_cast<T>(wv.w) * xr[4 * i + 3];
            }
          } else {
            const T* wd = Ws + o * DP;
#pragma unroll
            for (int i = 0; i < DP; i += 2) { acc0 += wd[i] * xr[i]; acc1 += wd[i + 1] * xr[i + 1]; }
          }
          py[static_cast<int64_t>(idx[DP + o]) * a.S] = acc0 + acc1;
        }
      }
      continue;
    }
    // row-major rows that did not fit the fast path below: stage the input tile (feature gather applied here)
    for (int e = threadIdx.x; e < nr * D; e += blockDim.x) {
      const int r = e / D, i = e - r * D;
      xs[r * ldt + i] = a.x[(r0 + r) * D + (a.in_idx ? __ldg(a.in_idx + i) : i)];
    }
    __syncthreads();
    const int r = threadIdx.x;
    T xr[DP];
#pragma unroll
    for (int i = 0; i < DP; ++i) {
      T v = (i < D && r < nr) ? xs[r * ldt + i] : T(0);
      if (a.sub_bias_first) v -= bs[i];
      xr[i] = v;
    }
    __syncthreads();
    if (r < nr) {
      for (int o = 0; o < D; ++o) {
        T acc = a.sub_bias_first ? T(0) : bs[o];
        const T* w = Ws + o * DP;
#pragma unroll
        for (int i = 0; i < DP; ++i) acc += w[i] * xr[i];
        xs[r * ldt + o] = acc;
      }
    }
    __syncthreads();
    for (int e = threadIdx.x; e < nr * D; e += blockDim.x) {
      const int rr = e / D, o = e - rr * D;
      a.y[(r0 + rr) * D + (a.out_idx ? __ldg(a.out_idx + o) : o)] = xs[rr * ldt + o];
    }
  }
}

// Row-major [rows, D] fast path (S == 1: LULinearPermute on tabular data).  The tile is one contiguous
// span in HBM, so it is staged and written back as a flat coalesced copy (no index arithmetic per
// element); the feature gather / scatter happens on the shared-memory side, where a row pitch of D
// (odd for the shapes in use) is conflict-free.  Two tiles per CTA iteration share one W load.
constexpr int kRowThreads = 128;   // 128-row tiles: 512 CTAs at B = 65 536, ~4 resident per SM, whose load / compute / store phases overlap

template <typename T, int DP>
__global__ void __launch_bounds__(kRowThreads) small_linear_rows_kernel(const LinArgs<T> a) {
  extern __shared__ __align__(16) unsigned char smem_lin[];
  const int D = a.D;
  T* Ws = reinterpret_cast<T*>(smem_lin);              // [D][DP], zero padded
  T* bs = Ws + static_cast<size_t>(D) * DP;            // [DP]
  int* iin = reinterpret_cast<int*>(bs + DP);          // [DP]
  int* iout = iin + DP;                                // [DP]
  T* xs = reinterpret_cast<T*>(iout + DP);             // [kRowThreads * D]
  T* ys = xs + static_cast<size_t>(kRowThreads) * D;   // [kRowThreads * D]
#pragma unroll 8
  for (int e = threadIdx.x; e < D * DP; e += kRowThreads) {            // (unrolled: independent loads in flight)
    const int o = e / DP, i = e - o * DP;
    Ws[e] = i < D ? __ldg(a.W + o * D + i) : T(0);
  }
  for (int e = threadIdx.x; e < DP; e += kRowThreads) {
    bs[e] = (e < D && a.bias) ? a.bias[e] : T(0);
    iin[e] = e < D ? (a.in_idx ? a.in_idx[e] : e) : 0;
    iout[e] = e < D ? (a.out_idx ? a.out_idx[e] : e) : 0;
  }
  const int64_t n_tiles = (a.rows + kRowThreads - 1) / kRowThreads;
  const int r = threadIdx.x;
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int64_t r0 = tile * kRowThreads;
    const int nr = static_cast<int>(a.rows - r0 < kRowThreads ? a.rows - r0 : kRowThreads);
    const int n = nr * D;
    const T* src = a.x + r0 * D;
    // the tile is one contiguous span: 16-byte loads, all of a thread's loads issued before its first store (a
    // load -> store pair per iteration serialises one memory latency per element)
    const bool vec = (static_cast<size_t>(n) * sizeof(T)) % 16 == 0 && reinterpret_cast<uintptr_t>(src) % 16 == 0 &&
                     reinterpret_cast<uintptr_t>(a.y + r0 * D) % 16 == 0;
    constexpr int kPerVec = 16 / sizeof(T);
    if (vec) {
      const int nv = n / kPerVec;
      const uint4* sv = reinterpret_cast<const uint4*>(src);
      uint4* xv = reinterpret_cast<uint4*>(xs);
      for (int e0 = threadIdx.x; e0 < nv; e0 += 8 * kRowThreads) {
        uint4 v[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int e = e0 + k * kRowThreads;
          if (e < nv) v[k] = __ldg(sv + e);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int e = e0 + k * kRowThreads;
          if (e < nv) xv[e] = v[k];
        }
      }
    } else {
#pragma unroll 8
      for (int e = threadIdx.x; e < n; e += kRowThreads) xs[e] = __ldg(src + e);
    }
    __syncthreads();                                   // (also orders the W / index loads on the first tile)
    if (r < nr) {
      T xr[DP];
#pragma unroll
      for (int i = 0; i < DP; ++i) {
        T v = i < D ? xs[r * D + iin[i]] : T(0);
        if (a.sub_bias_first) v -= bs[i];
        xr[i] = v;
      }
      // two outputs per step: four independent FMA chains (same per-output summation order as one output per step)
      int o = 0;
      for (; o + 1 < D; o += 2) {
        T a0 = a.sub_bias_first ? T(0) : bs[o], a1 = T(0), c0 = a.sub_bias_first ? T(0) : bs[o + 1], c1 = T(0);
        const T* w0 = Ws + o * DP;
        const T* w1 = w0 + DP;
#pragma unroll
        for (int i = 0; i < DP; i += 2) {
          a0 += w0[i] * xr[i]; a1 += w0[i + 1] * xr[i + 1];
          c0 += w1[i] * xr[i]; c1 += w1[i + 1] * xr[i + 1];
        }
        ys[r * D + iout[o]] = a0 + a1;
        ys[r * D + iout[o + 1]] = c0 + c1;
      }
      if (o < D) {
        T acc0 = a.sub_bias_first ? T(0) : bs[o], acc1 = T(0);
        const T* w = Ws + o * DP;
#pragma unroll
        for (int i = 0; i < DP; i += 2) { acc0 += w[i] * xr[i]; acc1 += w[i + 1] * xr[i + 1]; }
        ys[r * D + iout[o]] = acc0 + acc1;
      }
    }
    __syncthreads();
    T* dst = a.y + r0 * D;
    if (vec) {
      const int nv = n / kPerVec;
      const uint4* yv = reinterpret_cast<const uint4*>(ys);
      uint4* dv = reinterpret_cast<uint4*>(dst);
      for (int e = threadIdx.x; e < nv; e += kRowThreads) dv[e] = yv[e];
    } else {
      for (int e = threadIdx.x; e < n; e += kRowThreads) dst[e] = ys[e];
    }
  }
}

template <typename T, int DP>
static int launch_lin(const LinArgs<T>& a, cudaStream_t st) {
  const size_t smem_rows = (static_cast<size_t>(a.D) * DP + DP) * sizeof(T) + 2 * DP * sizeof(int) +
                           2 * static_cast<size_t>(kRowThreads) * a.D * sizeof(T);
  if (a.S == 1 && smem_rows <= 160 * 1024) {
    const size_t smem = smem_rows;
    auto kern = small_linear_rows_kernel<T, DP>;
    static bool attr_set = false;
    if (!attr_set) {
      cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
      attr_set = true;
    }
    const int sms = sm_count() > 0 ? sm_count() : VCNF_SM_COUNT_DEFAULT;
    const int64_t tiles = (a.rows + kRowThreads - 1) / kRowThreads;
    const int64_t cap = static_cast<int64_t>(sms) * 6;
    kern<<<static_cast<unsigned>(tiles < cap ? tiles : cap), kRowThreads, smem, st>>>(a);
    return launch_status();
  }
  const size_t smem = (static_cast<size_t>(a.D) * DP + DP + static_cast<size_t>(kLinRows) * (a.D + 1)) * sizeof(T);
  auto kern = small_linear_kernel<T, DP>;
  if (smem > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
  const int sms = sm_count() > 0 ? sm_count() : VCNF_SM_COUNT_DEFAULT;
  const int64_t tiles = (a.rows + kLinRows - 1) / kLinRows;
  const int64_t cap = static_cast<int64_t>(sms) * 8;
  kern<<<static_cast<unsigned>(tiles < cap ? tiles : cap), kLinRows, smem, st>>>(a);
  return launch_status();
}

template <typename T>
static int dispatch_lin(const LinArgs<T>& a, cudaStream_t st) {
  if (a.D <= 4) return launch_lin<T, 4>(a, st);
  if (a.D <= 8) return launch_lin<T, 8>(a, st);
  if (a.D <= 16) return launch_lin<T, 16>(a, st);
  if (a.D <= 32) return launch_lin<T, 32>(a, st);
  if (a.D <= 48) return launch_lin<T, 48>(a, st);
  if (a.D <= 64) return launch_lin<T, 64>(a, st);
  return VCNF_E_UNSUPPORTED;
}


// ---------------------------------------------------------------------------------------------------
// Backward of the row-major map above in ONE pass over (x, gy):
//   gx[r, in_idx[i]] = sum_o W[o, i] gy[r, out_idx[o]]
//   gW[o, i]         = sum_r gy[r, out_idx[o]] x[r, in_idx[i]]        gb[o] = sum_r gy[r, out_idx[o]]
// (autograd of mixing.py:398-410 after the permutation).  It replaces an index_select, a transposed
// small_linear launch and a [D, rows] x [rows, D] weight-gradient GEMM (36 us per cfg-2 layer) and reads each
// operand once.  A tile of rows is staged flat in shared memory; thread r back-propagates row r, then the
// same threads re-map to entries of gW (entry e = tid + 256 k: lanes share o and walk i, so the gy read is a
// broadcast and the x read conflict-free) and accumulate over the tile's rows in registers.  Per-CTA partial
// sums go to the workspace and are summed in a fixed order by lin_partial_reduce_kernel (bit-reproducible).
constexpr int kBwdThreads = 256;
constexpr int kBwdMaxGrid = 148 * 4;

template <typename T>
struct LinBwdArgs {
  const T* x; const T* gy; const T* W; const int32_t* in_idx; const int32_t* out_idx;
  T* gx; T* partial; int64_t rows; int D; int tile_rows; int bulk;
};

// Data movement: the raw tile (two contiguous spans, x rows and gy rows) is pulled into a 2-stage ring by the
// TMA engine (1-D bulk copies on an mbarrier), so the load of tile i+1 overlaps the math of tile i (the first
// version loaded synchronously: 39 % of its stall samples were the load -> store dependency).  From the ring
// the tile is re-laid with the permutations applied (xs[r][i] = x[r, in_idx[i]], gs[r][o] = gy[r, out_idx[o]])
// at an even row pitch, so every later access is a direct, pair-aligned index:
//   * gW: thread <-> 2 x 2 blocks of (o, i); per row two 2-element loads feed four FMAs;
//   * gx: two threads per row, each back-propagates half of the input features from the row in registers.
template <typename T, int DP, int NB>
__global__ void __launch_bounds__(kBwdThreads, 2) small_linear_bwd_kernel(const LinBwdArgs<T> a) {
  extern __shared__ __align__(128) unsigned char smem_lin[];
  const int D = a.D, PD = (D + 1) & ~1, TR = a.tile_rows;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_lin);                    // [2]
  T* raw = reinterpret_cast<T*>(smem_lin + 128);       // [2 stages][x | gy][TR * D]
  T* Wt = raw + 4 * static_cast<size_t>(TR) * D;       // [D][DP]: Wt[i][o] = W[o][i], zero padded
  int* inv_in = reinterpret_cast<int*>(Wt + static_cast<size_t>(D) * DP);   // inv_in[in_idx[i]] = i
  int* inv_out = inv_in + DP;
  T* xs = reinterpret_cast<T*>(inv_out + DP);          // [TR][PD]
  T* gs = xs + static_cast<size_t>(TR) * PD;
  T* os = gs + static_cast<size_t>(TR) * PD;
  const int tid = threadIdx.x;
  const int64_t n_tiles = (a.rows + TR - 1) / TR;
  auto issue = [&](int64_t t, int st) {                // thread 0 only
    const int64_t r0 = t * TR;
    const int nr = static_cast<int>(a.rows - r0 < TR ? a.rows - r0 : TR);
    const uint32_t bytes = static_cast<uint32_t>(static_cast<size_t>(nr) * D * sizeof(T));
    if (!a.bulk || bytes % 16 != 0) return;
    T* dst = raw + static_cast<size_t>(st) * 2 * TR * D;
    mbar_expect_tx(&bars[st], 2 * bytes);
    bulk_g2s(dst, a.x + r0 * D, bytes, &bars[st]);
    bulk_g2s(dst + static_cast<size_t>(TR) * D, a.gy + r0 * D, bytes, &bars[st]);
  };
  if (tid == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    mbar_init_fence();
    if (static_cast<int64_t>(blockIdx.x) < n_tiles) issue(blockIdx.x, 0);
  }
  for (int e = tid; e < D * DP; e += kBwdThreads) {
    const int i = e / DP, o = e - i * DP;
    Wt[e] = o < D ? a.W[o * D + i] : T(0);
  }
  for (int e = tid; e < D; e += kBwdThreads) {
    inv_in[a.in_idx ? a.in_idx[e] : e] = e;
    inv_out[a.out_idx ? a.out_idx[e] : e] = e;
  }
  for (int e = tid; e < TR * PD; e += kBwdThreads) { xs[e] = T(0); gs[e] = T(0); }   // pad column stays 0
  const int nbi = (D + 1) >> 1, nblk = nbi * nbi;
  int bo[NB], bi[NB];
  T acc[NB][4];
#pragma unroll
  for (int k = 0; k < NB; ++k) {
    const int blk = tid + k * kBwdThreads;
    const int q = blk < nblk ? blk : 0;
    bo[k] = 2 * (q / nbi); bi[k] = 2 * (q - (q / nbi) * nbi);
    acc[k][0] = acc[k][1] = acc[k][2] = acc[k][3] = T(0);
  }
  T accb = T(0);
  uint32_t par0 = 0u, par1 = 0u;
  int it = 0;
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
    const int st = it & 1;
    const int64_t r0 = tile * TR;
    const int nr = static_cast<int>(a.rows - r0 < TR ? a.rows - r0 : TR);
    const int n = nr * D;
    __syncthreads();                                   // previous tile's readers (ring stage st^1, xs, gs, os) are done
    if (tid == 0 && tile + gridDim.x < n_tiles) issue(tile + gridDim.x, st ^ 1);
    const bool bulk = a.bulk && (static_cast<size_t>(n) * sizeof(T)) % 16 == 0;
    if (bulk) {
      mbar_wait(&bars[st], st ? par1 : par0);
      if (st) par1 ^= 1u; else par0 ^= 1u;
      const T* rx = raw + static_cast<size_t>(st) * 2 * TR * D;
      const T* rg = rx + static_cast<size_t>(TR) * D;
      if (D <= 32) {                                   // lane = column (its permuted slots are loop invariants), warp = row:
        const int c = tid & 31;                        // no division and no index lookup per element
        if (c < D) {
          const int ci = inv_in[c], co = inv_out[c];
          for (int r = tid >> 5; r < nr; r += kBwdThreads / 32) {
            xs[r * PD + ci] = rx[r * D + c];
            gs[r * PD + co] = rg[r * D + c];
          }
        }
      } else {
        for (int e = tid; e < n; e += kBwdThreads) {
          const int r = e / D, c = e - r * D;
          xs[r * PD + inv_in[c]] = rx[e];
          gs[r * PD + inv_out[c]] = rg[e];
        }
      }
    } else {
      const T* sx = a.x + r0 * D;
      const T* sg = a.gy + r0 * D;
      for (int e = tid; e < n; e += kBwdThreads) {
        const int r = e / D, c = e - r * D;
        xs[r * PD + inv_in[c]] = sx[e];
        gs[r * PD + inv_out[c]] = sg[e];
      }
    }
    __syncthreads();
    if (a.gx) {
      const int r = tid >> 1;
      if (r < nr) {
        T gr[DP];
#pragma unroll
        for (int o = 0; o < DP; ++o) gr[o] = o < D ? gs[r * PD + o] : T(0);
        const int half = (D + 1) >> 1;
        const int i0 = (tid & 1) * half, i1 = i0 + half < D ? i0 + half : D;
        for (int i = i0; i < i1; ++i) {
          T a0 = T(0), a1 = T(0);
          const T* w = Wt + i * DP;
#pragma unroll
          for (int o = 0; o < DP; o += 2) { a0 += w[o] * gr[o]; a1 += w[o + 1] * gr[o + 1]; }
          os[r * PD + i] = a0 + a1;
        }
      }
      for (int r2 = r + kBwdThreads / 2; r2 < nr; r2 += kBwdThreads / 2) {   // tiles taller than 128 rows
        for (int i = (tid & 1); i < D; i += 2) {
          T a0 = T(0);
          for (int o = 0; o < D; ++o) a0 += Wt[i * DP + o] * gs[r2 * PD + o];
          os[r2 * PD + i] = a0;
        }
      }
    }
#pragma unroll 4
    for (int r = 0; r < nr; ++r) {
      const T* grow = gs + r * PD;
      const T* xrow = xs + r * PD;
#pragma unroll
      for (int k = 0; k < NB; ++k) {
        const T g0 = grow[bo[k]], g1 = grow[bo[k] + 1];
        const T x0 = xrow[bi[k]], x1 = xrow[bi[k] + 1];
        acc[k][0] += g0 * x0; acc[k][1] += g0 * x1; acc[k][2] += g1 * x0; acc[k][3] += g1 * x1;
      }
    }
    if (tid < D)
      for (int r = 0; r < nr; ++r) accb += gs[r * PD + tid];
    if (a.gx) {
      __syncthreads();
      T* dst = a.gx + r0 * D;
      if (D <= 32) {
        const int c = tid & 31;
        if (c < D) {
          const int ci = inv_in[c];
          for (int r = tid >> 5; r < nr; r += kBwdThreads / 32) dst[r * D + c] = os[r * PD + ci];
        }
      } else {
        for (int e = tid; e < n; e += kBwdThreads) {
          const int r = e / D, c = e - r * D;
          dst[e] = os[r * PD + inv_in[c]];
        }
      }
    }
  }
  T* mine = a.partial + static_cast<size_t>(blockIdx.x) * (D * D + D);
#pragma unroll
  for (int k = 0; k < NB; ++k) {
    if (tid + k * kBwdThreads < nblk) {
      const int o = bo[k], i = bi[k];
      mine[o * D + i] = acc[k][0];
      if (i + 1 < D) mine[o * D + i + 1] = acc[k][1];
      if (o + 1 < D) {
        mine[(o + 1) * D + i] = acc[k][2];
        if (i + 1 < D) mine[(o + 1) * D + i + 1] = acc[k][3];
      }
    }
  }
  if (tid < D) mine[D * D + tid] = accb;
}

// out[j] = sum_b partial[b, j] in a fixed order (one warp per output, fixed xor tree); j < nw -> gW, else gb
template <typename T>
__global__ void lin_partial_reduce_kernel(const T* partial, int nblocks, int nw, int nb, T* gW, T* gb) {
  const int j = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (j >= nw + nb) return;
  T acc = T(0);
  for (int b = lane; b < nblocks; b += 32) acc += partial[static_cast<size_t>(b) * (nw + nb) + j];
  acc = warp_sum(acc);
  if (lane == 0) {
    if (j < nw) { if (gW) gW[j] = acc; }
    else if (gb) gb[j - nw] = acc;
  }
}

template <typename T>
static size_t lin_bwd_smem(int D, int DP, int rows) {
  const int PD = (D + 1) & ~1;
  return 128 + 4 * static_cast<size_t>(rows) * D * sizeof(T) + static_cast<size_t>(D) * DP * sizeof(T) +
         2 * DP * sizeof(int) + 3 * static_cast<size_t>(rows) * PD * sizeof(T);
}

template <typename T, int DP, int NB>
static int launch_lin_bwd(LinBwdArgs<T>& a, T* gW, T* gb, cudaStream_t st) {
  a.tile_rows = 128;                                   // two CTAs per SM when the tile allows it
  while (a.tile_rows > 16 && lin_bwd_smem<T>(a.D, DP, a.tile_rows) > 110 * 1024) a.tile_rows >>= 1;
  const size_t smem = lin_bwd_smem<T>(a.D, DP, a.tile_rows);
  if (smem > 200 * 1024) return VCNF_E_UNSUPPORTED;
  a.bulk = (reinterpret_cast<uintptr_t>(a.x) % 16 == 0) && (reinterpret_cast<uintptr_t>(a.gy) % 16 == 0) &&
           (static_cast<size_t>(a.tile_rows) * a.D * sizeof(T)) % 16 == 0;
  auto kern = small_linear_bwd_kernel<T, DP, NB>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    attr_set = true;
  }
  const int sms = sm_count() > 0 ? sm_count() : VCNF_SM_COUNT_DEFAULT;
  const int64_t tiles = (a.rows + a.tile_rows - 1) / a.tile_rows;
  int per_sm = static_cast<int>((227 * 1024) / (smem + 1024));
  per_sm = per_sm < 1 ? 1 : (per_sm > 2 ? 2 : per_sm);
  int64_t grid = static_cast<int64_t>(sms) * per_sm;
  if (grid > kBwdMaxGrid) grid = kBwdMaxGrid;
  if (grid > tiles) grid = tiles;
  kern<<<static_cast<unsigned>(grid), kBwdThreads, smem, st>>>(a);
  int rc = launch_status();
  if (rc) return rc;
  const int nw = a.D * a.D, nb = a.D;
  lin_partial_reduce_kernel<T><<<(nw + nb + 3) / 4, 128, 0, st>>>(a.partial, static_cast<int>(grid), nw, nb, gW, gb);
  return launch_status();
}

template <typename T>
static int dispatch_lin_bwd(LinBwdArgs<T>& a, T* gW, T* gb, cudaStream_t st) {
  if (a.D <= 16) return launch_lin_bwd<T, 16, 1>(a, gW, gb, st);     // NB = ceil(ceil(D/2)^2 / 256) blocks per thread
  if (a.D <= 32) return launch_lin_bwd<T, 32, 1>(a, gW, gb, st);
  if (a.D <= 48) return launch_lin_bwd<T, 48, 3>(a, gW, gb, st);
  if (a.D <= 64) return launch_lin_bwd<T, 64, 4>(a, gW, gb, st);
  return VCNF_E_UNSUPPORTED;
}

// ---------------------------------------------------------------------------------------------------
// Parameter gradients of the per-pixel map on image tensors [B, D, S] (Invertible1x1Conv, mixing.py:101-128):
//   gW[o, i] = sum_{b, s} gy[b, out_idx[o], s] x[b, in_idx[i], s],   gb[o] = sum gy[b, out_idx[o], s].
// Both operands are read as they lie (pixel index contiguous): no [pixels, D] transposed copies and no library GEMM
// (the previous route: two full-tensor transposes + the tabular weight-gradient kernel = 7 launches per layer).
// A CTA walks tiles of 64 pixels; thread e, e + 256, ... own entries of gW and accumulate in registers; per-CTA
// partials are summed in a fixed order by warp_reduce_chunks_kernel.
constexpr int kImgTile = 64;
constexpr int kImgGrid = 148 * 2;

template <typename T>
__global__ void __launch_bounds__(256) small_linear_wgrad_img_kernel(const T* __restrict__ x, const T* __restrict__ gy,
                                                                    const int32_t* __restrict__ in_idx,
                                                                    const int32_t* __restrict__ out_idx, int64_t P, int D,
                                                                    int S, T* __restrict__ partial) {
  extern __shared__ __align__(16) unsigned char smem_lin[];
  const int ldt = kImgTile + 1;
  T* xs = reinterpret_cast<T*>(smem_lin);              // [D][65]
  T* gs = xs + static_cast<size_t>(D) * ldt;           // [D][65]
  const int n_out = D * D;
  T acc[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) acc[k] = T(0);
  T accb = T(0);
  const int sp = threadIdx.x & (kImgTile - 1), d0 = threadIdx.x >> 6;
  const int64_t n_tiles = (P + kImgTile - 1) / kImgTile;
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int64_t pix = tile * kImgTile + sp;
    const bool ok = pix < P;
    const int64_t b = ok ? pix / S : 0, s0 = ok ? pix - b * S : 0;
    const T* px = x + b * D * S + s0;
    const T* pg = gy + b * D * S + s0;
    __syncthreads();
#pragma unroll 4
    for (int d = d0; d < D; d += 4) {
      xs[d * ldt + sp] = ok ? __ldg(px + static_cast<int64_t>(in_idx ? __ldg(in_idx + d) : d) * S) : T(0);
      gs[d * ldt + sp] = ok ? __ldg(pg + static_cast<int64_t>(out_idx ? __ldg(out_idx + d) : d) * S) : T(0);
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      const int e = threadIdx.x + 256 * k;
      if (e < n_out) {
        const int o = e / D, i = e - o * D;
        const T* g = gs + o * ldt;
        const T* xx = xs + i * ldt;
        T a0 = T(0), a1 = T(0);
#pragma unroll 8
        for (int q = 0; q < kImgTile; q += 2) { a0 += g[q] * xx[q]; a1 += g[q + 1] * xx[q + 1]; }
        acc[k] += a0 + a1;
      }
    }
    if (static_cast<int>(threadIdx.x) < D) {
      const T* g = gs + threadIdx.x * ldt;
      T a0 = T(0);
      for (int q = 0; q < kImgTile; ++q) a0 += g[q];
      accb += a0;
    }
  }
  T* out = partial + static_cast<int64_t>(blockIdx.x) * (n_out + D);
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    const int e = threadIdx.x + 256 * k;
    if (e < n_out) out[e] = acc[k];
  }
  if (static_cast<int>(threadIdx.x) < D) out[n_out + threadIdx.x] = accb;
}

template <typename T>
static int launch_wgrad_img(const void* x, const void* gy, const int32_t* in_idx, const int32_t* out_idx, void* gW, void* gb,
                            void* workspace, int64_t B, int D, int S, cudaStream_t st) {
  const int64_t P = B * S;
  const int64_t tiles = (P + kImgTile - 1) / kImgTile;
  const int grid = static_cast<int>(tiles < kImgGrid ? tiles : kImgGrid);
  const size_t smem = 2 * static_cast<size_t>(D) * (kImgTile + 1) * sizeof(T);
  auto kern = small_linear_wgrad_img_kernel<T>;
  if (smem > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
  T* part = static_cast<T*>(workspace);
  kern<<<grid, 256, smem, st>>>(static_cast<const T*>(x), static_cast<const T*>(gy), in_idx, out_idx, P, D, S, part);
  int rc = launch_status();
  if (rc) return rc;
  const int n_out = D * D;
  if (gW) launch_warp_reduce<T>(part, grid, n_out + D, n_out, static_cast<T*>(gW), st);
  if (gb) launch_warp_reduce<T>(part + n_out, grid, n_out + D, D, static_cast<T*>(gb), st);
  return launch_status();
}

}  // namespace vcnf

using namespace vcnf;

extern "C" int vcnf_small_linear(const void* x, void* y, const void* W, const void* bias, const int32_t* in_idx,
                                 const int32_t* out_idx, int64_t rows, int D, int S, int sub_bias_first,
                                 int dtype, void* stream) {
  if (rows == 0) return 0;   /* empty batch: pointers may be NULL */
  if (!x || !y || !W || rows < 0 || D <= 0 || S <= 0 || x == y || (dtype != VCNF_F32 && dtype != VCNF_F64))
    return VCNF_E_BADARG;
  if (rows == 0) return 0;
  if (dtype == VCNF_F32) {
    LinArgs<float> a{(const float*)x, (float*)y, (const float*)W, (const float*)bias, in_idx, out_idx, rows, D, S, sub_bias_first};
    return dispatch_lin<float>(a, (cudaStream_t)stream);
  }
  LinArgs<double> a{(const double*)x, (double*)y, (const double*)W, (const double*)bias, in_idx, out_idx, rows, D, S, sub_bias_first};
  return dispatch_lin<double>(a, (cudaStream_t)stream);
}

extern "C" int64_t vcnf_small_linear_bwd_workspace(int D, int dtype) {
  if (D <= 0) return 0;
  return static_cast<int64_t>(kBwdMaxGrid) * (static_cast<int64_t>(D) * D + D) * (dtype == VCNF_F64 ? 8 : 4);
}

extern "C" int vcnf_small_linear_bwd(const void* x, const void* gy, const void* W, const int32_t* in_idx,
                                     const int32_t* out_idx, void* gx, void* gW, void* gb, void* workspace,
                                     int64_t rows, int D, int dtype, void* stream) {
  if (D <= 0 || D > 64 || rows < 0 || (dtype != VCNF_F32 && dtype != VCNF_F64)) return VCNF_E_BADARG;
  cudaStream_t st = (cudaStream_t)stream;
  const size_t es = dtype == VCNF_F64 ? 8 : 4;
  if (rows == 0) {   /* empty batch: parameter gradients are zero */
    if (gW) cudaMemsetAsync(gW, 0, es * D * D, st);
    if (gb) cudaMemsetAsync(gb, 0, es * D, st);
    return launch_status();
  }
  if (!x || !gy || !W || !workspace || (!gW && !gb && !gx)) return VCNF_E_BADARG;
  if (dtype == VCNF_F32) {
    LinBwdArgs<float> a{(const float*)x, (const float*)gy, (const float*)W, in_idx, out_idx, (float*)gx,
                        (float*)workspace, rows, D, 0, 0};
    return dispatch_lin_bwd<float>(a, (float*)gW, (float*)gb, st);
  }
  LinBwdArgs<double> a{(const double*)x, (const double*)gy, (const double*)W, in_idx, out_idx, (double*)gx,
                       (double*)workspace, rows, D, 0, 0};
  return dispatch_lin_bwd<double>(a, (double*)gW, (double*)gb, st);
}

extern "C" int64_t vcnf_small_linear_wgrad_img_workspace(int D, int dtype) {
  if (D <= 0) return 0;
  return static_cast<int64_t>(kImgGrid) * (static_cast<int64_t>(D) * D + D) * (dtype == VCNF_F64 ? 8 : 4);
}

extern "C" int vcnf_small_linear_wgrad_img(const void* x, const void* gy, const int32_t* in_idx, const int32_t* out_idx,
                                           void* gW, void* gb, void* workspace, int64_t B, int D, int S, int dtype,
                                           void* stream) {
  if (D <= 0 || D > 64 || B < 0 || S <= 0 || (dtype != VCNF_F32 && dtype != VCNF_F64) || (!gW && !gb)) return VCNF_E_BADARG;
  cudaStream_t st = (cudaStream_t)stream;
  const size_t es = dtype == VCNF_F64 ? 8 : 4;
  if (B == 0) {
    if (gW) cudaMemsetAsync(gW, 0, es * D * D, st);
    if (gb) cudaMemsetAsync(gb, 0, es * D, st);
    return launch_status();
  }
  if (!x || !gy || !workspace) return VCNF_E_BADARG;
  if (dtype == VCNF_F32) return launch_wgrad_img<float>(x, gy, in_idx, out_idx, gW, gb, workspace, B, D, S, st);
  return launch_wgrad_img<double>(x, gy, in_idx, out_idx, gW, gb, workspace, B, D, S, st);
}
