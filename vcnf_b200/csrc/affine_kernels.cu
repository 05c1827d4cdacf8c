// Affine bijector kernels for sm_100a: affine coupling block (split + apply + merge + log-det),
// masked affine flow, MAF elementwise, constant affine (ActNorm) -- forward, inverse, backward.
// All are HBM-bound streaming kernels: every tensor is read/written exactly once,
// fully coalesced, with the per-sample log|det| reduced in-kernel by warp shuffles.
//
// Work decomposition: G threads cooperate on one sample (G = 1 for tiny rows such as the 2-D
// RealNVP of cfg 1, G = 32 one warp per row for tabular widths, G = 256 one CTA per image).
#include <mutex>
#include "common.cuh"
#include "affine_math.cuh"
#include "../../include/vcnf_b200.h"

namespace vcnf {

constexpr int kAffThreads = 256;

template <typename T, int G>
__device__ __forceinline__ T group_sum(T v, T* red) {
  if (G == 1) return v;
  v = warp_sum(v);
  if (G == 32) return v;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) red[w] = v;
  __syncthreads();
  T acc = T(0);
#pragma unroll
  for (int i = 0; i < G / 32; ++i) acc += red[i];
  return acc;
}

template <typename T>
struct CouplingArgs {
  const T* z; const T* param; T* out; T* logdet;
  const T* gout; const T* gld; T* gz; T* gparam;
  int64_t B; int C; int S; int c2_start; int C2; int map; int inverse; int swap_ts;
};

// z, out, gout, gz: [B, C, S]; param, gparam: [B, np*C2, S] (np = 1 for SCALE_NONE else 2);
// shift = param[:, 0::2], raw scale = param[:, 1::2] (coupling.py:122-123) unless swap_ts (MAF).
template <typename T, int G, bool BWD>
__global__ void __launch_bounds__(kAffThreads) coupling_kernel(const CouplingArgs<T> a) {
  __shared__ T red[kAffThreads / 32];
  const int per_block = kAffThreads / G;
  const int sub = threadIdx.x / G, lane = threadIdx.x % G;
  const int n = a.C * a.S;
  const int np = a.map == SCALE_NONE ? 1 : 2;
  const int64_t n_iter = (a.B + per_block - 1) / per_block;
  for (int64_t it = blockIdx.x; it < n_iter; it += gridDim.x) {
    const int64_t b = it * per_block + sub;
    const bool live = b < a.B;
    T ld = T(0);
    if (live) {
      const T* zb = a.z + b * n;
      const T* pb = a.param + b * static_cast<int64_t>(np) * a.C2 * a.S;
      const T gl = BWD ? a.gld[b] : T(0);
      for (int e = lane; e < n; e += G) {
        const int c = e / a.S, s = e - c * a.S;
        const int j = c - a.c2_start;
        const T zv = zb[e];
        if (j < 0 || j >= a.C2) {
          if (BWD) a.gz[b * n + e] = a.gout[b * n + e];
          else a.out[b * n + e] = zv;
          continue;
        }
        T t, sr = T(0);
        int64_t it_ = 0, is_ = 0;
        if (np == 1) { it_ = static_cast<int64_t>(j) * a.S + s; t = pb[it_]; }
        else {
          it_ = (static_cast<int64_t>(2 * j + (a.swap_ts ? 1 : 0))) * a.S + s;
          is_ = (static_cast<int64_t>(2 * j + (a.swap_ts ? 0 : 1))) * a.S + s;
          t = pb[it_]; sr = pb[is_];
        }
        if (!BWD) {
          T y, l;
          affine_apply(a.map, a.inverse, zv, t, sr, &y, &l);
          a.out[b * n + e] = y;
          ld += l;
        } else {
          T gz, gt, gs;
          affine_adjoint(a.map, a.inverse, zv, t, sr, a.gout[b * n + e], gl, &gz, &gt, &gs);
          a.gz[b * n + e] = gz;
          T* gp = a.gparam + b * static_cast<int64_t>(np) * a.C2 * a.S;
          gp[it_] = gt;
          if (np == 2) gp[is_] = gs;
        }
      }
    }
    if (!BWD) {
      ld = group_sum<T, G>(ld, red);
      if (live && lane == 0) a.logdet[b] = ld;
    }
  }
}

template <typename T>
struct MaskedArgs {
  const T* z; const T* mask; const T* s; const T* t; T* out; T* logdet;
  const T* gout; const T* gld; T* gz; T* gs; T* gt;
  int64_t B; int n; int inverse;
};

template <typename T, int G, bool BWD>
__global__ void __launch_bounds__(kAffThreads) masked_affine_kernel(const MaskedArgs<T> a) {
  __shared__ T red[kAffThreads / 32];
  const int per_block = kAffThreads / G;
  const int sub = threadIdx.x / G, lane = threadIdx.x % G;
  const int64_t n_iter = (a.B + per_block - 1) / per_block;
  for (int64_t it = blockIdx.x; it < n_iter; it += gridDim.x) {
    const int64_t b = it * per_block + sub;
    const bool live = b < a.B;
    T ld = T(0);
    if (live) {
      const T gl = BWD ? a.gld[b] : T(0);
      for (int e = lane; e < a.n; e += G) {
        const int64_t i = b * a.n + e;
        const T m = __ldg(a.mask + e);
        const T sv = a.s ? a.s[i] : T(0);
        const T tv = a.t ? a.t[i] : T(0);
        if (!BWD) {
          T y, l;
          masked_affine_apply(a.inverse, m, a.z[i], tv, sv, &y, &l);
          a.out[i] = y;
          ld += l;
        } else {
          T gz, gt, gs;
          masked_affine_adjoint(a.inverse, m, a.z[i], tv, sv, a.gout[i], gl, &gz, &gt, &gs);
          a.gz[i] = gz;
          if (a.gs) a.gs[i] = gs;
          if (a.gt) a.gt[i] = gt;
        }
      }
    }
    if (!BWD) {
      ld = group_sum<T, G>(ld, red);
      if (live && lane == 0) a.logdet[b] = ld;
    }
  }
}

// ---- constant affine (AffineConstFlow / ActNorm): z [B, Pn, S], s, t [Pn] ----------------------
template <typename T>
__global__ void __launch_bounds__(kAffThreads) const_affine_kernel(const T* z, const T* s, const T* t, T* out,
                                                                   int64_t total, int Pn, int S, int inverse) {
  for (int64_t e = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; e < total;
       e += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>((e / S) % Pn);
    const T sv = __ldg(s + c), tv = __ldg(t + c);
    out[e] = inverse ? (z[e] - tv) * vexp(-sv) : z[e] * vexp(sv) + tv;
  }
}

// Backward: gz elementwise plus per-chunk partial sums of (gs, gt) per channel:
//   forward : gs = sum g*z*exp(s), gt = sum g          inverse : gs = -sum g*(z-t)*exp(-s), gt = -sum g*exp(-s)
// partial layout [n_chunks, 2, Pn]; finished by const_affine_reduce_kernel in chunk order.
template <typename T>
__global__ void __launch_bounds__(kAffThreads) const_affine_bwd_kernel(const T* z, const T* s, const T* t,
                                                                       const T* g, T* gz, T* partial, int64_t B,
                                                                       int Pn, int S, int inverse,
                                                                       int64_t rows_per_chunk) {
  __shared__ T sm_s[kAffThreads], sm_t[kAffThreads];
  const int chunk = blockIdx.x;
  const int64_t b0 = chunk * rows_per_chunk;
  const int64_t b1 = b0 + rows_per_chunk < B ? b0 + rows_per_chunk : B;
  T acc_s = T(0), acc_t = T(0);
  if (S == 1) {
    // threads keep a fixed column: stride is a multiple of Pn
    const int active = (kAffThreads / Pn) * Pn;
    if (static_cast<int>(threadIdx.x) < active) {
      const int c = threadIdx.x % Pn;
      const T sv = __ldg(s + c), tv = __ldg(t + c);
      const T e_ = vexp(inverse ? -sv : sv);
      for (int64_t e = b0 * Pn + threadIdx.x; e < b1 * Pn; e += active) {
        const T gv = g[e], zv = z[e];
        gz[e] = gv * e_;
        if (!inverse) { acc_s += gv * zv * e_; acc_t += gv; }
        else { acc_s -= gv * (zv - tv) * e_; acc_t -= gv * e_; }
      }
    }
    sm_s[threadIdx.x] = acc_s; sm_t[threadIdx.x] = acc_t;
    __syncthreads();
    for (int c = threadIdx.x; c < Pn; c += blockDim.x) {
      T as = T(0), at = T(0);
      for (int j = c; j < active; j += Pn) { as += sm_s[j]; at += sm_t[j]; }
      partial[(static_cast<int64_t>(chunk) * 2 + 0) * Pn + c] = as;
      partial[(static_cast<int64_t>(chunk) * 2 + 1) * Pn + c] = at;
    }
  } else {
    // one channel per blockIdx.y: contiguous runs of S elements
    const int c = blockIdx.y;
    const T sv = __ldg(s + c), tv = __ldg(t + c);
    const T e_ = vexp(inverse ? -sv : sv);
    for (int64_t b = b0; b < b1; ++b) {
      const int64_t base = (b * Pn + c) * S;
      for (int i = threadIdx.x; i < S; i += blockDim.x) {
        const T gv = g[base + i], zv = z[base + i];
        gz[base + i] = gv * e_;
        if (!inverse) { acc_s += gv * zv * e_; acc_t += gv; }
        else { acc_s -= gv * (zv - tv) * e_; acc_t -= gv * e_; }
      }
    }
    sm_s[threadIdx.x] = acc_s; sm_t[threadIdx.x] = acc_t;
    __syncthreads();
    if (threadIdx.x == 0) {
      T as = T(0), at = T(0);
      for (int j = 0; j < kAffThreads; ++j) { as += sm_s[j]; at += sm_t[j]; }
      partial[(static_cast<int64_t>(chunk) * 2 + 0) * Pn + c] = as;
      partial[(static_cast<int64_t>(chunk) * 2 + 1) * Pn + c] = at;
    }
  }
}

template <typename T>
__global__ void const_affine_reduce_kernel(const T* partial, int n_chunks, int Pn, T* gs, T* gt) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= Pn) return;
  T as = T(0), at = T(0);
  for (int k = 0; k < n_chunks; ++k) {
    as += partial[(static_cast<int64_t>(k) * 2 + 0) * Pn + c];
    at += partial[(static_cast<int64_t>(k) * 2 + 1) * Pn + c];
  }
  gs[c] = as; gt[c] = at;
}

// ---- channel permutation / squeeze ------------------------------------------------------------
// dst[b, j, s] = src[b, perm[j], s]
template <typename T>
__global__ void permute_channels_kernel(const T* src, T* dst, int64_t total, int C, int S, const int32_t* perm) {
  for (int64_t e = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; e < total;
       e += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t bc = e / S;
    const int s = static_cast<int>(e - bc * S);
    const int64_t b = bc / C;
    const int j = static_cast<int>(bc - b * C);
    dst[e] = src[(b * C + __ldg(perm + j)) * S + s];
  }
}

// Squeeze (flows/reshape.py:93-117).  `forward`: in [B, 4c, h, w] -> out [B, c, 2h, 2w] with
// out[b, c, 2h+i, 2w+j] = in[b, 4c+2i+j, h, w]; inverse is the opposite gather.
template <typename T>
__global__ void squeeze_kernel(const T* src, T* dst, int64_t total, int Cs, int H, int W, int inverse) {
  // Cs = small channel count c, (H, W) = small spatial size h, w; total = B*4*Cs*H*W
  for (int64_t e = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; e < total;
       e += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    int64_t r = e;
    if (!inverse) {   // e indexes dst [B, Cs, 2H, 2W]
      const int x = static_cast<int>(r % (2 * W)); r /= 2 * W;
      const int y = static_cast<int>(r % (2 * H)); r /= 2 * H;
      const int c = static_cast<int>(r % Cs); const int64_t b = r / Cs;
      const int i = y & 1, j = x & 1;
      dst[e] = src[((b * 4 * Cs + 4 * c + 2 * i + j) * H + (y >> 1)) * W + (x >> 1)];
    } else {          // e indexes dst [B, 4Cs, H, W]
      const int x = static_cast<int>(r % W); r /= W;
      const int y = static_cast<int>(r % H); r /= H;
      const int cc = static_cast<int>(r % (4 * Cs)); const int64_t b = r / (4 * Cs);
      const int c = cc >> 2, i = (cc >> 1) & 1, j = cc & 1;
      dst[e] = src[((b * Cs + c) * 2 * H + 2 * y + i) * 2 * W + 2 * x + j];
    }
  }
}

static inline unsigned stream_grid(int64_t total, int threads) {
  const int sms = sm_count() > 0 ? sm_count() : VCNF_SM_COUNT_DEFAULT;
  int64_t g = (total + threads - 1) / threads;
  const int64_t cap = static_cast<int64_t>(sms) * 16;
  return static_cast<unsigned>(g < 1 ? 1 : (g > cap ? cap : g));
}

static inline int pick_group(int n) { return n <= 8 ? 1 : (n <= 1024 ? 32 : 256); }

template <typename T, bool BWD>
static int launch_coupling(const CouplingArgs<T>& a, cudaStream_t st) {
  const int G = pick_group(a.C * a.S);
  const int64_t blocks = (a.B + (kAffThreads / G) - 1) / (kAffThreads / G);
  const unsigned grid = stream_grid(blocks * kAffThreads, kAffThreads);
  if (G == 1) coupling_kernel<T, 1, BWD><<<grid, kAffThreads, 0, st>>>(a);
  else if (G == 32) coupling_kernel<T, 32, BWD><<<grid, kAffThreads, 0, st>>>(a);
  else coupling_kernel<T, 256, BWD><<<grid, kAffThreads, 0, st>>>(a);
  return launch_status();
}

template <typename T, bool BWD>
static int launch_masked(const MaskedArgs<T>& a, cudaStream_t st) {
  const int G = pick_group(a.n);
  const int64_t blocks = (a.B + (kAffThreads / G) - 1) / (kAffThreads / G);
  const unsigned grid = stream_grid(blocks * kAffThreads, kAffThreads);
  if (G == 1) masked_affine_kernel<T, 1, BWD><<<grid, kAffThreads, 0, st>>>(a);
  else if (G == 32) masked_affine_kernel<T, 32, BWD><<<grid, kAffThreads, 0, st>>>(a);
  else masked_affine_kernel<T, 256, BWD><<<grid, kAffThreads, 0, st>>>(a);
  return launch_status();
}

constexpr int kConstAffineChunks = 592;   // 4 x 148

}  // namespace vcnf

using namespace vcnf;

#define VCNF_BAD_DTYPE(d) ((d) != VCNF_F32 && (d) != VCNF_F64)

/* No-allocation variant of vcnf_const_affine_bwd (no counterpart in the reference; see include/vcnf_b200.h).
 * The per-chunk partial sums live in a file-scope device array instead of a caller workspace.  In host code the
 * array's name is only a handle: its device address is looked up once per device with cudaGetSymbolAddress
 * (that device current) and that pointer is what the kernels get. */
constexpr int kStaticMaxPn = VCNF_CONST_AFFINE_STATIC_MAX_PN;
static __device__ __align__(256) unsigned char g_const_affine_scratch[kConstAffineChunks * 2 * kStaticMaxPn * 8];

static int const_affine_static_scratch(void** out) {
  constexpr int kMaxDev = 64;
  static void* ptrs[kMaxDev] = {};
  static std::mutex mu;
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return static_cast<int>(e);
  if (dev < 0 || dev >= kMaxDev) return VCNF_E_UNSUPPORTED;
  std::lock_guard<std::mutex> lock(mu);
  if (!ptrs[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_const_affine_scratch);
    if (e != cudaSuccess) return static_cast<int>(e);
    if (!p) return static_cast<int>(cudaErrorInvalidSymbol);
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return 0;
}


extern "C" {

int vcnf_affine_coupling_fwd(const void* z, const void* param, void* out, void* logdet, int64_t B, int C, int S,
                             int c2_start, int C2, int scale_map, int inverse, int swap_ts, int dtype,
                             void* stream) {
  if (B == 0) return 0;   /* empty batch: pointers may be NULL */
  if (!z || !param || !out || !logdet || B < 0 || C <= 0 || S <= 0 || C2 < 0 || c2_start < 0 ||
      c2_start + C2 > C || scale_map < 0 || scale_map > 4 || VCNF_BAD_DTYPE(dtype))
    return VCNF_E_BADARG;
  if (B == 0) return 0;
#define GO(T)                                                                                         \
  {                                                                                                   \
    CouplingArgs<T> a{};                                                                              \
    a.z = (const T*)z; a.param = (const T*)param; a.out = (T*)out; a.logdet = (T*)logdet;             \
    a.B = B; a.C = C; a.S = S; a.c2_start = c2_start; a.C2 = C2; a.map = scale_map;                   \
    a.inverse = inverse; a.swap_ts = swap_ts;                                                         \
    return launch_coupling<T, false>(a, (cudaStream_t)stream);                                        \
  }
  if (dtype == VCNF_F32) GO(float) else GO(double)
#undef GO
}

int vcnf_affine_coupling_bwd(const void* z, const void* param, const void* gout, const void* glogdet, void* gz,
                             void* gparam, int64_t B, int C, int S, int c2_start, int C2, int scale_map,
                             int inverse, int swap_ts, int dtype, void* stream) {
  if (B == 0) return 0;   /* empty batch: pointers may be NULL */
  if (!z || !param || !gout || !glogdet || !gz || !gparam || B < 0 || C <= 0 || S <= 0 || C2 < 0 ||
      c2_start < 0 || c2_start + C2 > C || scale_map < 0 || scale_map > 4 || VCNF_BAD_DTYPE(dtype))
    return VCNF_E_BADARG;
  if (B == 0) return 0;
#define GO(T)                                                                                         \
  {                                                                                                   \
    CouplingArgs<T> a{};                                                                              \
    a.z = (const T*)z; a.param = (const T*)param; a.gout = (const T*)gout; a.gld = (const T*)glogdet; \
    a.gz = (T*)gz; a.gparam = (T*)gparam;                                                             \
    a.B = B; a.C = C; a.S = S; a.c2_start = c2_start; a.C2 = C2; a.map = scale_map;                   \
    a.inverse = inverse; a.swap_ts = swap_ts;                                                         \
    return launch_coupling<T, true>(a, (cudaStream_t)stream);                                         \
  }
  if (dtype == VCNF_F32) GO(float) else GO(double)
#undef GO
}

int vcnf_masked_affine_fwd(const void* z, const void* mask, const void* s, const void* t, void* out, void* logdet,
                           int64_t B, int n, int inverse, int dtype, void* stream) {
  if (B == 0) return 0;   /* empty batch: pointers may be NULL */
  if (!z || !mask || !out || !logdet || B < 0 || n <= 0 || VCNF_BAD_DTYPE(dtype)) return VCNF_E_BADARG;
  if (B == 0) return 0;
#define GO(T)                                                                                         \
  {                                                                                                   \
    MaskedArgs<T> a{};                                                                                \
    a.z = (const T*)z; a.mask = (const T*)mask; a.s = (const T*)s; a.t = (const T*)t;                 \
    a.out = (T*)out; a.logdet = (T*)logdet; a.B = B; a.n = n; a.inverse = inverse;                    \
    return launch_masked<T, false>(a, (cudaStream_t)stream);                                          \
  }
  if (dtype == VCNF_F32) GO(float) else GO(double)
#undef GO
}

int vcnf_masked_affine_bwd(const void* z, const void* mask, const void* s, const void* t, const void* gout,
                           const void* glogdet, void* gz, void* gs, void* gt, int64_t B, int n, int inverse,
                           int dtype, void* stream) {
  if (B == 0) return 0;   /* empty batch: pointers may be NULL */
  if (!z || !mask || !gout || !glogdet || !gz || B < 0 || n <= 0 || VCNF_BAD_DTYPE(dtype)) return VCNF_E_BADARG;
  if (B == 0) return 0;
#define GO(T)                                                                                         \
  {                                                                                                   \
    MaskedArgs<T> a{};                                                                                \
    a.z = (const T*)z; a.mask = (const T*)mask; a.s = (const T*)s; a.t = (const T*)t;                 \
    a.gout = (const T*)gout; a.gld = (const T*)glogdet; a.gz = (T*)gz; a.gs = (T*)gs; a.gt = (T*)gt;  \
    a.B = B; a.n = n; a.inverse = inverse;                                                            \
    return launch_masked<T, true>(a, (cudaStream_t)stream);                                           \
  }
  if (dtype == VCNF_F32) GO(float) else GO(double)
#undef GO
}

int vcnf_const_affine_fwd(const void* z, const void* s, const void* t, void* out, int64_t B, int Pn, int S,
                          int inverse, int dtype, void* stream) {
  if (B == 0) return 0;   /* empty batch: pointers may be NULL */
  if (!z || !s || !t || !out || B < 0 || Pn <= 0 || S <= 0 || VCNF_BAD_DTYPE(dtype)) return VCNF_E_BADARG;
  const int64_t total = B * Pn * S;
  if (total == 0) return 0;
  const unsigned grid = stream_grid(total, kAffThreads);
  if (dtype == VCNF_F32)
    const_affine_kernel<float><<<grid, kAffThreads, 0, (cudaStream_t)stream>>>((const float*)z, (const float*)s,
                                                                              (const float*)t, (float*)out, total, Pn, S, inverse);
  else
    const_affine_kernel<double><<<grid, kAffThreads, 0, (cudaStream_t)stream>>>((const double*)z, (const double*)s,
                                                                               (const double*)t, (double*)out, total, Pn, S, inverse);
  return launch_status();
}

int64_t vcnf_const_affine_bwd_workspace(int Pn, int dtype) {
  return static_cast<int64_t>(kConstAffineChunks) * 2 * Pn * (dtype == VCNF_F64 ? 8 : 4);
}

int vcnf_const_affine_bwd(const void* z, const void* s, const void* t, const void* gout, void* gz, void* gs,
                          void* gt, void* workspace, int64_t B, int Pn, int S, int inverse, int dtype,
                          void* stream) {
  if (B == 0) return 0;   /* empty batch: pointers may be NULL */
  if (!z || !s || !t || !gout || !gz || !gs || !gt || !workspace || B < 0 || Pn <= 0 || S <= 0 ||
      (S == 1 && Pn > kAffThreads) || (S > 1 && Pn > 65535) || VCNF_BAD_DTYPE(dtype))
    return VCNF_E_BADARG;
  cudaStream_t st = (cudaStream_t)stream;
  const size_t eb = dtype == VCNF_F64 ? 8 : 4;
  if (B == 0) {
    cudaMemsetAsync(gs, 0, Pn * eb, st);
    cudaMemsetAsync(gt, 0, Pn * eb, st);
    return launch_status();
  }
  int64_t chunks = B < kConstAffineChunks ? B : kConstAffineChunks;
  const int64_t rpc = (B + chunks - 1) / chunks;
  chunks = (B + rpc - 1) / rpc;
  dim3 grid(static_cast<unsigned>(chunks), S == 1 ? 1 : Pn);
#define GO(T)                                                                                            \
  {                                                                                                      \
    const_affine_bwd_kernel<T><<<grid, kAffThreads, 0, st>>>((const T*)z, (const T*)s, (const T*)t,      \
                                                            (const T*)gout, (T*)gz, (T*)workspace, B, Pn, S, inverse, rpc); \
    int rc = launch_status();                                                                            \
    if (rc) return rc;                                                                                   \
    /* partial layout [chunk][2][Pn]: one warp per output sums the chunks (fixed order) */                \
    launch_warp_reduce<T>((const T*)workspace, (int)chunks, 2 * Pn, Pn, (T*)gs, st);                      \
    launch_warp_reduce<T>((const T*)workspace + Pn, (int)chunks, 2 * Pn, Pn, (T*)gt, st);                 \
    return launch_status();                                                                              \
  }
  if (dtype == VCNF_F32) GO(float) else GO(double)
#undef GO
}

/* No-allocation variant of vcnf_const_affine_bwd: scratch = g_const_affine_scratch (file scope, above). */
int vcnf_const_affine_bwd_static(const void* z, const void* s, const void* t, const void* gout, void* gz, void* gs,
                                 void* gt, int64_t B, int Pn, int S, int inverse, int dtype, void* stream) {
  if (B == 0) return 0;   /* empty batch: pointers may be NULL */
  if (!z || !s || !t || !gout || !gz || !gs || !gt || B < 0 || Pn <= 0 || S <= 0 || VCNF_BAD_DTYPE(dtype))
    return VCNF_E_BADARG;
  if (Pn > kStaticMaxPn) return VCNF_E_UNSUPPORTED;   /* the static scratch is sized for Pn <= 256 */
  void* scratch = nullptr;
  int rc = const_affine_static_scratch(&scratch);
  if (rc) return rc;
  /* every partial the reduce reads is written by this call's first kernel: nothing needs clearing */
  return vcnf_const_affine_bwd(z, s, t, gout, gz, gs, gt, scratch, B, Pn, S, inverse, dtype, stream);
}

int vcnf_permute_channels(const void* src, void* dst, int64_t B, int C, int S, const int32_t* perm, int dtype,
                          void* stream) {
  if (B == 0) return 0;   /* empty batch: pointers may be NULL */
  if (!src || !dst || !perm || B < 0 || C <= 0 || S <= 0 || src == dst || VCNF_BAD_DTYPE(dtype)) return VCNF_E_BADARG;
  const int64_t total = B * C * S;
  if (total == 0) return 0;
  const unsigned grid = stream_grid(total, 256);
  if (dtype == VCNF_F32)
    permute_channels_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>((const float*)src, (float*)dst, total, C, S, perm);
  else
    permute_channels_kernel<double><<<grid, 256, 0, (cudaStream_t)stream>>>((const double*)src, (double*)dst, total, C, S, perm);
  return launch_status();
}

int vcnf_squeeze(const void* src, void* dst, int64_t B, int c_small, int h_small, int w_small, int inverse,
                 int dtype, void* stream) {
  if (B == 0) return 0;   /* empty batch: pointers may be NULL */
  if (!src || !dst || B < 0 || c_small <= 0 || h_small <= 0 || w_small <= 0 || src == dst || VCNF_BAD_DTYPE(dtype))
    return VCNF_E_BADARG;
  const int64_t total = B * 4 * c_small * h_small * w_small;
  if (total == 0) return 0;
  const unsigned grid = stream_grid(total, 256);
  if (dtype == VCNF_F32)
    squeeze_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>((const float*)src, (float*)dst, total, c_small, h_small, w_small, inverse);
  else
    squeeze_kernel<double><<<grid, 256, 0, (cudaStream_t)stream>>>((const double*)src, (double*)dst, total, c_small, h_small, w_small, inverse);
  return launch_status();
}

}  // extern "C"
