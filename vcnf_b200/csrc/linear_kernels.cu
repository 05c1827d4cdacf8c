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
  for (int e = threadIdx.x; e < D * DP; e += blockDim.x) {
    const int o = e / DP, i = e - o * DP;
    Ws[e] = i < D ? a.W[o * D + i] : T(0);
  }
  for (int e = threadIdx.x; e < DP; e += blockDim.x) bs[e] = (e < D && a.bias) ? a.bias[e] : T(0);
  const int64_t n_tiles = (a.rows + kLinRows - 1) / kLinRows;
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int64_t r0 = tile * kLinRows;
    const int nr = static_cast<int>(a.rows - r0 < kLinRows ? a.rows - r0 : kLinRows);
    __syncthreads();
    // stage the input tile (feature gather applied here)
    if (a.S == 1) {
      for (int e = threadIdx.x; e < nr * D; e += blockDim.x) {
        const int r = e / D, i = e - r * D;
        xs[r * ldt + i] = a.x[(r0 + r) * D + (a.in_idx ? __ldg(a.in_idx + i) : i)];
      }
    } else {
      for (int e = threadIdx.x; e < nr * D; e += blockDim.x) {
        const int i = e / nr, r = e - i * nr;
        const int64_t row = r0 + r;
        const int64_t b = row / a.S, s = row - b * a.S;
        xs[r * ldt + i] = a.x[(b * D + (a.in_idx ? __ldg(a.in_idx + i) : i)) * a.S + s];
      }
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
    if (a.S == 1) {
      for (int e = threadIdx.x; e < nr * D; e += blockDim.x) {
        const int rr = e / D, o = e - rr * D;
        a.y[(r0 + rr) * D + (a.out_idx ? __ldg(a.out_idx + o) : o)] = xs[rr * ldt + o];
      }
    } else {
      for (int e = threadIdx.x; e < nr * D; e += blockDim.x) {
        const int o = e / nr, rr = e - o * nr;
        const int64_t row = r0 + rr;
        const int64_t b = row / a.S, s = row - b * a.S;
        a.y[(b * D + (a.out_idx ? __ldg(a.out_idx + o) : o)) * a.S + s] = xs[rr * ldt + o];
      }
    }
  }
}

// Row-major [rows, D] fast path (S == 1: LULinearPermute on tabular data).  The tile is one contiguous
// span in HBM, so it is staged and written back as a flat coalesced copy (no index arithmetic per
// element); the feature gather / scatter happens on the shared-memory side, where a row pitch of D
// (odd for the shapes in use) is conflict-free.  Two tiles per CTA iteration share one W load.
constexpr int kRowThreads = 256;

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
  for (int e = threadIdx.x; e < D * DP; e += kRowThreads) {
    const int o = e / DP, i = e - o * DP;
    Ws[e] = i < D ? a.W[o * D + i] : T(0);
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
    for (int e = threadIdx.x; e < n; e += kRowThreads) xs[e] = src[e];
    __syncthreads();                                   // (also orders the W / index loads on the first tile)
    if (r < nr) {
      T xr[DP];
#pragma unroll
      for (int i = 0; i < DP; ++i) {
        T v = i < D ? xs[r * D + iin[i]] : T(0);
        if (a.sub_bias_first) v -= bs[i];
        xr[i] = v;
      }
      for (int o = 0; o < D; ++o) {
        T acc0 = a.sub_bias_first ? T(0) : bs[o], acc1 = T(0);
        const T* w = Ws + o * DP;
#pragma unroll
        for (int i = 0; i < DP; i += 2) { acc0 += w[i] * xr[i]; acc1 += w[i + 1] * xr[i + 1]; }
        ys[r * D + iout[o]] = acc0 + acc1;
      }
    }
    __syncthreads();
    T* dst = a.y + r0 * D;
    for (int e = threadIdx.x; e < n; e += kRowThreads) dst[e] = ys[e];
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
    const int64_t cap = static_cast<int64_t>(sms) * 2;
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
