// Rational-quadratic spline LAYER kernels (forward / inverse / backward) for sm_100a.
//
// One launch = one whole coupling or autoregressive spline layer after its conditioner:
// split (index_select), per-element spline with conditional or batch-shared parameters,
// merge (scatter) and the per-sample log|det| reduction -- the ~600 ATen ops / 16 `nonzero`
// syncs of normflow/utils/splines.py:20-193 + flows/neural_spline/coupling.py:70-159.
//
// Data movement (HBM-bound, SURVEY.md 8d: s*(T*P + 2D + 1) bytes per sample):
//   * the per-sample parameter block [R rows x Tc*Pc] of a tile is one CONTIGUOUS span in HBM;
//     it is pulled into shared memory by the TMA engine with a 1-D bulk async copy
//     (cp.async.bulk ... mbarrier::complete_tx) into a 2-stage ring, so the copy of tile i+1
//     overlaps the math of tile i and no registers are spent on staging;
//   * threads then read their element's P parameters from shared memory at stride P (P = 3K-1
//     is odd -> bank-conflict free), x / y go straight to/from HBM fully coalesced
//     (consecutive threads = consecutive columns of a row);
//   * persistent grid: #CTAs = SMs x resident CTAs, each walks tiles blockIdx.x, +gridDim.x, ...
//   * log|det|: per-element terms go through shared memory and are summed per row in feature
//     order by one thread (deterministic).
#include "common.cuh"
#include "rqs_math.cuh"
#include "../../include/vcnf_b200.h"

namespace vcnf {

constexpr int kThreads = 256;

template <typename T>
struct RqsLayerArgs {
  const T* x; int64_t ldx;
  T* y; int64_t ldy;
  T* logdet; int accumulate;
  int64_t B; int D; int K;
  const int32_t* col_map;
  const T* cond; int64_t ld_cond; int Tc; int Pc;
  const uint8_t* cmodes; const T* cbounds; T wh_scale;
  const T* shared; int Ts; int Ps;
  const uint8_t* smodes; const T* sbounds;
  int inverse; int outside_zero;
  int8_t* bin_out; int32_t* err_flag;
  int R; int bulk_in; int64_t num_tiles;
  // backward only
  const T* gy; int64_t ldgy; const T* gld;
  T* gx; int64_t ldgx;
  T* gcond; int64_t ld_gcond; int bulk_out;
  T* partial;     // [gridDim.x, Ts*Ps]
};

__host__ __device__ inline size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
__host__ __device__ inline int64_t imin64(int64_t a, int64_t b) { return a < b ? a : b; }

// Fill one parameter stage for tile `t`.  Full tiles of a contiguous, 16 B-aligned parameter
// matrix go through the TMA engine; anything else is copied cooperatively.
template <typename T>
__device__ __forceinline__ void stage_issue(const RqsLayerArgs<T>& a, int64_t t, T* stage, uint64_t* bar) {
  const int64_t rows = imin64(a.R, a.B - t * a.R);
  if (a.Tc > 0 && a.bulk_in && rows == a.R && threadIdx.x == 0) {
    const uint32_t bytes = static_cast<uint32_t>(static_cast<size_t>(a.R) * a.Tc * a.Pc * sizeof(T));
    mbar_expect_tx(bar, bytes);
    bulk_g2s(stage, a.cond + t * a.R * a.ld_cond, bytes, bar);
  }
}

template <typename T>
__device__ __forceinline__ void stage_wait(const RqsLayerArgs<T>& a, int64_t t, T* stage, uint64_t* bar,
                                           uint32_t parity) {
  if (a.Tc <= 0) return;
  const int64_t rows = imin64(a.R, a.B - t * a.R);
  if (a.bulk_in && rows == a.R) {
    mbar_wait(bar, parity);
  } else {
    const int n = a.Tc * a.Pc;
    for (int64_t e = threadIdx.x; e < rows * n; e += blockDim.x) {
      const int64_t r = e / n;
      stage[e] = a.cond[(t * a.R + r) * a.ld_cond + (e - r * n)];
    }
  }
}

template <typename T, int KMAX>
__global__ void __launch_bounds__(kThreads) rqs_layer_fwd_kernel(const RqsLayerArgs<T> a) {
  extern __shared__ __align__(128) unsigned char smem[];
  const size_t stage_bytes = round_up(static_cast<size_t>(a.R) * a.Tc * a.Pc * sizeof(T), 128);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
  T* stage0 = reinterpret_cast<T*>(smem + 128);
  T* stage1 = reinterpret_cast<T*>(smem + 128 + stage_bytes);
  T* sh = reinterpret_cast<T*>(smem + 128 + 2 * stage_bytes);
  T* lad_s = sh + round_up(static_cast<size_t>(a.Ts) * a.Ps, 32);

  const int tid = threadIdx.x;
  if (tid == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    mbar_init_fence();
  }
  for (int i = tid; i < a.Ts * a.Ps; i += blockDim.x) sh[i] = a.shared[i];
  __syncthreads();

  int flag = 0;
  int64_t tile = blockIdx.x;
  if (tile < a.num_tiles) stage_issue(a, tile, stage0, &bars[0]);
  for (int it = 0; tile < a.num_tiles; tile += gridDim.x, ++it) {
    const int s = it & 1;
    T* cur = s ? stage1 : stage0;
    const int64_t next = tile + gridDim.x;
    if (next < a.num_tiles) stage_issue(a, next, s ? stage0 : stage1, &bars[s ^ 1]);
    const int64_t row0 = tile * a.R;
    const int rows = static_cast<int>(imin64(a.R, a.B - row0));
    stage_wait(a, tile, cur, &bars[s], (it >> 1) & 1);
    __syncthreads();

    const int n_el = rows * a.D;
    for (int e = tid; e < n_el; e += blockDim.x) {
      const int r = e / a.D;
      const int c = e - r * a.D;
      const int cm = __ldg(a.col_map + c);
      const int kind = cm >> 16, idx = cm & 0xffff;
      const T v = a.x[(row0 + r) * a.ldx + c];
      T out = v, lad = T(0);
      int bin = -2;
      if (kind != VCNF_COL_PASS) {
        const T* prm;
        T scale, bound;
        RqsElemCfg cfg{a.K, 0, a.inverse, a.outside_zero};
        if (kind == VCNF_COL_COND) {
          prm = cur + (static_cast<size_t>(r) * a.Tc + idx) * a.Pc;
          scale = a.wh_scale;
          cfg.mode = __ldg(a.cmodes + idx);
          bound = __ldg(a.cbounds + idx);
        } else {
          prm = sh + static_cast<size_t>(idx) * a.Ps;
          scale = T(1);
          cfg.mode = __ldg(a.smodes + idx);
          bound = __ldg(a.sbounds + idx);
        }
        rqs_element<T, KMAX>(prm, v, bound, scale, cfg, &out, &lad, &bin, &flag);
      }
      a.y[(row0 + r) * a.ldy + c] = out;
      lad_s[e] = lad;
      if (a.bin_out) a.bin_out[(row0 + r) * a.D + c] = static_cast<int8_t>(bin);
    }
    __syncthreads();
    for (int r = tid; r < rows; r += blockDim.x) {
      T acc = T(0);
      for (int c = 0; c < a.D; ++c) acc += lad_s[r * a.D + c];
      if (a.accumulate) acc += a.logdet[row0 + r];
      a.logdet[row0 + r] = acc;
    }
    __syncthreads();
  }
  if (flag && a.err_flag) atomicOr(a.err_flag, flag);
}

// Backward: recomputes the bin record from (x, params) and applies the hand-derived adjoint
// (rqs_math.cuh: rqs_element_bwd).  Parameter gradients of a tile are assembled in shared memory
// and leave through one bulk store (TMA engine) per tile; batch-shared parameter gradients are
// reduced per CTA in a fixed order and finished by rqs_partial_reduce_kernel (deterministic).
template <typename T, int KMAX>
__global__ void __launch_bounds__(kThreads) rqs_layer_bwd_kernel(const RqsLayerArgs<T> a) {
  extern __shared__ __align__(128) unsigned char smem[];
  const size_t stage_bytes = round_up(static_cast<size_t>(a.R) * a.Tc * a.Pc * sizeof(T), 128);
  const size_t sgrad_bytes = round_up(static_cast<size_t>(a.R) * a.Ts * a.Ps * sizeof(T), 128);
  const int n_sh = a.Ts * a.Ps;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
  T* stage0 = reinterpret_cast<T*>(smem + 128);
  T* stage1 = reinterpret_cast<T*>(smem + 128 + stage_bytes);
  T* gstage = reinterpret_cast<T*>(smem + 128 + 2 * stage_bytes);
  T* sstage = reinterpret_cast<T*>(smem + 128 + 3 * stage_bytes);
  T* sh = reinterpret_cast<T*>(smem + 128 + 3 * stage_bytes + sgrad_bytes);
  T* accum = sh + round_up(static_cast<size_t>(n_sh), 32);

  const int tid = threadIdx.x;
  if (tid == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    mbar_init_fence();
  }
  for (int i = tid; i < n_sh; i += blockDim.x) { sh[i] = a.shared[i]; accum[i] = T(0); }
  __syncthreads();

  int64_t tile = blockIdx.x;
  if (tile < a.num_tiles) stage_issue(a, tile, stage0, &bars[0]);
  for (int it = 0; tile < a.num_tiles; tile += gridDim.x, ++it) {
    const int s = it & 1;
    T* cur = s ? stage1 : stage0;
    const int64_t next = tile + gridDim.x;
    if (next < a.num_tiles) stage_issue(a, next, s ? stage0 : stage1, &bars[s ^ 1]);
    const int64_t row0 = tile * a.R;
    const int rows = static_cast<int>(imin64(a.R, a.B - row0));
    stage_wait(a, tile, cur, &bars[s], (it >> 1) & 1);
    __syncthreads();

    const int n_el = rows * a.D;
    for (int e = tid; e < n_el; e += blockDim.x) {
      const int r = e / a.D;
      const int c = e - r * a.D;
      const int cm = __ldg(a.col_map + c);
      const int kind = cm >> 16, idx = cm & 0xffff;
      const T g_out = a.gy[(row0 + r) * a.ldgy + c];
      T gv = g_out;
      if (kind != VCNF_COL_PASS) {
        const T v = a.x[(row0 + r) * a.ldx + c];
        const T g_lad = a.gld[row0 + r];
        const T* prm;
        T* gprm;
        T scale, bound;
        RqsElemCfg cfg{a.K, 0, a.inverse, a.outside_zero};
        if (kind == VCNF_COL_COND) {
          const size_t off = (static_cast<size_t>(r) * a.Tc + idx) * a.Pc;
          prm = cur + off;
          gprm = gstage + off;
          scale = a.wh_scale;
          cfg.mode = __ldg(a.cmodes + idx);
          bound = __ldg(a.cbounds + idx);
        } else {
          prm = sh + static_cast<size_t>(idx) * a.Ps;
          gprm = sstage + (static_cast<size_t>(r) * a.Ts + idx) * a.Ps;
          scale = T(1);
          cfg.mode = __ldg(a.smodes + idx);
          bound = __ldg(a.sbounds + idx);
        }
        rqs_element_bwd<T, KMAX>(prm, v, bound, scale, cfg, g_out, g_lad, &gv, gprm);
      }
      a.gx[(row0 + r) * a.ldgx + c] = gv;
    }
    const bool bulk_store = a.Tc > 0 && a.bulk_out && rows == a.R;
    if (bulk_store) fence_async_smem();
    __syncthreads();
    if (a.Tc > 0) {
      const int n = a.Tc * a.Pc;
      if (bulk_store) {
        if (tid == 0) {
          bulk_s2g(a.gcond + row0 * a.ld_gcond, gstage, static_cast<uint32_t>(static_cast<size_t>(a.R) * n * sizeof(T)));
          bulk_commit();
        }
      } else {
        for (int e = tid; e < rows * n; e += blockDim.x) {
          const int r = e / n;
          a.gcond[(row0 + r) * a.ld_gcond + (e - r * n)] = gstage[e];
        }
      }
    }
    for (int j = tid; j < n_sh; j += blockDim.x) {
      T acc = accum[j];
      for (int r = 0; r < rows; ++r) acc += sstage[static_cast<size_t>(r) * n_sh + j];
      accum[j] = acc;
    }
    if (bulk_store && tid == 0) bulk_wait_read0();
    __syncthreads();
  }
  for (int j = tid; j < n_sh; j += blockDim.x) a.partial[static_cast<size_t>(blockIdx.x) * n_sh + j] = accum[j];
}

template <typename T>
__global__ void rqs_partial_reduce_kernel(const T* partial, int nblocks, int n, T* out) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  T acc = T(0);
  for (int b = 0; b < nblocks; ++b) acc += partial[static_cast<size_t>(b) * n + j];
  out[j] = acc;
}

template <typename T>
__global__ void gather_cols_kernel(const T* src, int64_t lds, T* dst, int64_t ldd, int64_t B, int n,
                                   const int32_t* cols) {
  const int64_t total = B * n;
  for (int64_t e = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; e < total;
       e += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t b = e / n;
    const int j = static_cast<int>(e - b * n);
    dst[b * ldd + j] = src[b * lds + __ldg(cols + j)];
  }
}

// One thread per destination element so that zero fill and scatter never race:
// inv[c] = j with cols[j] == c, or -1.
template <typename T>
__global__ void scatter_cols_kernel(const T* src, int64_t lds, T* dst, int64_t ldd, int64_t B, int n,
                                    int D_dst, const int32_t* cols, int accumulate) {
  __shared__ int inv[1024];
  for (int c = threadIdx.x; c < D_dst; c += blockDim.x) inv[c] = -1;
  __syncthreads();
  for (int j = threadIdx.x; j < n; j += blockDim.x) inv[cols[j]] = j;
  __syncthreads();
  const int64_t total = B * D_dst;
  for (int64_t e = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; e < total;
       e += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t b = e / D_dst;
    const int c = static_cast<int>(e - b * D_dst);
    const int j = inv[c];
    if (j >= 0) {
      const T v = src[b * lds + j];
      dst[b * ldd + c] = accumulate ? dst[b * ldd + c] + v : v;
    } else if (!accumulate) {
      dst[b * ldd + c] = T(0);
    }
  }
}

// ------------------------------------------------------------------------------------------- host
constexpr size_t kSmemMax = 227 * 1024;

static int pick_rows(int Tc, int Pc, int Ts, int Ps, size_t elt, bool bwd, int D) {
  const size_t per_row = static_cast<size_t>(bwd ? (Tc * Pc > Ts * Ps ? Tc * Pc : Ts * Ps) : Tc * Pc) * elt;
  const size_t budget = bwd ? 20 * 1024 : 40 * 1024;
  int R = per_row ? static_cast<int>(budget / per_row) : 64;
  if (R > 64) R = 64;
  R = R / 4 * 4;
  if (R < 4) R = 4;
  (void)D;
  return R;
}

template <typename T, int KMAX>
static int launch_fwd(RqsLayerArgs<T>& a, cudaStream_t st) {
  const int sms = sm_count();
  if (sms <= 0) return static_cast<int>(cudaErrorNoDevice);
  a.R = pick_rows(a.Tc, a.Pc, a.Ts, a.Ps, sizeof(T), false, a.D);
  a.num_tiles = (a.B + a.R - 1) / a.R;
  const size_t tile_bytes = static_cast<size_t>(a.R) * a.Tc * a.Pc * sizeof(T);
  a.bulk_in = a.Tc > 0 && a.ld_cond == static_cast<int64_t>(a.Tc) * a.Pc && tile_bytes % 16 == 0 &&
              reinterpret_cast<uintptr_t>(a.cond) % 16 == 0;
  const size_t smem = 128 + 2 * round_up(tile_bytes, 128) +
                      (round_up(static_cast<size_t>(a.Ts) * a.Ps, 32) + static_cast<size_t>(a.R) * a.D) * sizeof(T);
  if (smem > kSmemMax) return VCNF_E_UNSUPPORTED;
  auto kern = rqs_layer_fwd_kernel<T, KMAX>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
  int per_sm = static_cast<int>(kSmemMax / (smem + 1024));
  per_sm = per_sm < 1 ? 1 : (per_sm > 4 ? 4 : per_sm);
  const int64_t grid = a.num_tiles < static_cast<int64_t>(sms) * per_sm ? a.num_tiles : static_cast<int64_t>(sms) * per_sm;
  kern<<<static_cast<unsigned>(grid), kThreads, smem, st>>>(a);
  return launch_status();
}

constexpr int kMaxBwdGrid = 148 * 2;

template <typename T, int KMAX>
static int launch_bwd(RqsLayerArgs<T>& a, T* g_shared, cudaStream_t st) {
  const int sms = sm_count();
  if (sms <= 0) return static_cast<int>(cudaErrorNoDevice);
  a.R = pick_rows(a.Tc, a.Pc, a.Ts, a.Ps, sizeof(T), true, a.D);
  a.num_tiles = (a.B + a.R - 1) / a.R;
  const size_t tile_bytes = static_cast<size_t>(a.R) * a.Tc * a.Pc * sizeof(T);
  a.bulk_in = a.Tc > 0 && a.ld_cond == static_cast<int64_t>(a.Tc) * a.Pc && tile_bytes % 16 == 0 &&
              reinterpret_cast<uintptr_t>(a.cond) % 16 == 0;
  a.bulk_out = a.Tc > 0 && a.ld_gcond == static_cast<int64_t>(a.Tc) * a.Pc && tile_bytes % 16 == 0 &&
               reinterpret_cast<uintptr_t>(a.gcond) % 16 == 0;
  const size_t smem = 128 + 3 * round_up(tile_bytes, 128) +
                      round_up(static_cast<size_t>(a.R) * a.Ts * a.Ps * sizeof(T), 128) +
                      2 * round_up(static_cast<size_t>(a.Ts) * a.Ps, 32) * sizeof(T);
  if (smem > kSmemMax) return VCNF_E_UNSUPPORTED;
  auto kern = rqs_layer_bwd_kernel<T, KMAX>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
  int per_sm = static_cast<int>(kSmemMax / (smem + 1024));
  per_sm = per_sm < 1 ? 1 : (per_sm > 2 ? 2 : per_sm);
  int64_t grid = static_cast<int64_t>(sms) * per_sm;
  if (grid > kMaxBwdGrid) grid = kMaxBwdGrid;
  if (grid > a.num_tiles) grid = a.num_tiles;
  kern<<<static_cast<unsigned>(grid), kThreads, smem, st>>>(a);
  int rc = launch_status();
  if (rc == 0 && a.Ts > 0) {
    const int n = a.Ts * a.Ps;
    rqs_partial_reduce_kernel<T><<<(n + 127) / 128, 128, 0, st>>>(a.partial, static_cast<int>(grid), n, g_shared);
    rc = launch_status();
  }
  return rc;
}

template <typename T>
static int dispatch_fwd(RqsLayerArgs<T>& a, cudaStream_t st) {
  if (a.K <= 8) return launch_fwd<T, 8>(a, st);
  if (a.K <= 16) return launch_fwd<T, 16>(a, st);
  if (a.K <= 32) return launch_fwd<T, 32>(a, st);
  return VCNF_E_UNSUPPORTED;
}
template <typename T>
static int dispatch_bwd(RqsLayerArgs<T>& a, T* g_shared, cudaStream_t st) {
  if (a.K <= 8) return launch_bwd<T, 8>(a, g_shared, st);
  if (a.K <= 16) return launch_bwd<T, 16>(a, g_shared, st);
  if (a.K <= 32) return launch_bwd<T, 32>(a, g_shared, st);
  return VCNF_E_UNSUPPORTED;
}

template <typename T>
static void fill_common(RqsLayerArgs<T>& a, const void* x, int64_t ldx, int64_t B, int D, int K,
                        const int32_t* col_map, const void* cond, int64_t ld_cond, int Tc, int Pc,
                        const uint8_t* cmodes, const void* cbounds, double wh_scale, const void* shared,
                        int Ts, int Ps, const uint8_t* smodes, const void* sbounds, int inverse,
                        int outside_zero) {
  a = RqsLayerArgs<T>{};
  a.x = static_cast<const T*>(x); a.ldx = ldx; a.B = B; a.D = D; a.K = K; a.col_map = col_map;
  a.cond = static_cast<const T*>(cond); a.ld_cond = ld_cond; a.Tc = Tc; a.Pc = Pc;
  a.cmodes = cmodes; a.cbounds = static_cast<const T*>(cbounds); a.wh_scale = static_cast<T>(wh_scale);
  a.shared = static_cast<const T*>(shared); a.Ts = Ts; a.Ps = Ps;
  a.smodes = smodes; a.sbounds = static_cast<const T*>(sbounds);
  a.inverse = inverse; a.outside_zero = outside_zero;
}

static bool bad_layer_args(const void* x, int64_t B, int D, int K, const int32_t* col_map, const void* cond,
                           int Tc, int Pc, const void* cm, const void* cb, const void* shared, int Ts, int Ps,
                           const void* sm, const void* sb, int dtype) {
  if (!x || B < 0 || D <= 0 || D > 65535 || K < 1 || !col_map || Tc < 0 || Ts < 0) return true;
  if (Tc > 0 && (!cond || !cm || !cb || Pc < 3 * K - 1 || Pc > 3 * K + 1)) return true;
  if (Ts > 0 && (!shared || !sm || !sb || Ps < 3 * K - 1 || Ps > 3 * K + 1)) return true;
  return dtype != VCNF_F32 && dtype != VCNF_F64;
}

}  // namespace vcnf

using namespace vcnf;

extern "C" {

int vcnf_version(void) { return 100; }
int vcnf_sm_count(void) { return sm_count(); }

int vcnf_rqs_layer_fwd(const void* x, int64_t ldx, void* y, int64_t ldy, void* logdet, int accumulate,
                       int64_t B, int D, int K, const int32_t* col_map, const void* cond_params,
                       int64_t ld_cond, int Tc, int Pc, const uint8_t* cond_modes, const void* cond_bounds,
                       double wh_scale, const void* shared_params, int Ts, int Ps,
                       const uint8_t* shared_modes, const void* shared_bounds, int inverse, int outside_zero,
                       int8_t* bin_out, int32_t* err_flag, int dtype, void* stream) {
  if (bad_layer_args(x, B, D, K, col_map, cond_params, Tc, Pc, cond_modes, cond_bounds, shared_params, Ts, Ps,
                     shared_modes, shared_bounds, dtype) || !y || !logdet || x == y)
    return VCNF_E_BADARG;
  if (B == 0) return 0;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
#define VCNF_FWD(T)                                                                                      \
  {                                                                                                      \
    RqsLayerArgs<T> a;                                                                                   \
    fill_common<T>(a, x, ldx, B, D, K, col_map, cond_params, ld_cond, Tc, Pc, cond_modes, cond_bounds,   \
                   wh_scale, shared_params, Ts, Ps, shared_modes, shared_bounds, inverse, outside_zero); \
    a.y = static_cast<T*>(y); a.ldy = ldy; a.logdet = static_cast<T*>(logdet); a.accumulate = accumulate; \
    a.bin_out = bin_out; a.err_flag = err_flag;                                                          \
    return dispatch_fwd<T>(a, st);                                                                       \
  }
  if (dtype == VCNF_F32) VCNF_FWD(float) else VCNF_FWD(double)
#undef VCNF_FWD
}

int64_t vcnf_rqs_layer_bwd_workspace(int Ts, int Ps, int dtype) {
  if (Ts <= 0) return 0;
  return static_cast<int64_t>(kMaxBwdGrid) * Ts * Ps * (dtype == VCNF_F64 ? 8 : 4);
}

int vcnf_rqs_layer_bwd(const void* x, int64_t ldx, const void* gy, int64_t ldgy, const void* glogdet, void* gx,
                       int64_t ldgx, int64_t B, int D, int K, const int32_t* col_map, const void* cond_params,
                       int64_t ld_cond, int Tc, int Pc, const uint8_t* cond_modes, const void* cond_bounds,
                       double wh_scale, void* g_cond, int64_t ld_gcond, const void* shared_params, int Ts,
                       int Ps, const uint8_t* shared_modes, const void* shared_bounds, void* g_shared,
                       void* workspace, int inverse, int outside_zero, int dtype, void* stream) {
  if (bad_layer_args(x, B, D, K, col_map, cond_params, Tc, Pc, cond_modes, cond_bounds, shared_params, Ts, Ps,
                     shared_modes, shared_bounds, dtype) || !gy || !glogdet || !gx || (Tc > 0 && !g_cond) ||
      (Ts > 0 && (!g_shared || !workspace)))
    return VCNF_E_BADARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (B == 0) {
    if (Ts > 0) cudaMemsetAsync(g_shared, 0, static_cast<size_t>(Ts) * Ps * (dtype == VCNF_F64 ? 8 : 4), st);
    return launch_status();
  }
#define VCNF_BWD(T)                                                                                      \
  {                                                                                                      \
    RqsLayerArgs<T> a;                                                                                   \
    fill_common<T>(a, x, ldx, B, D, K, col_map, cond_params, ld_cond, Tc, Pc, cond_modes, cond_bounds,   \
                   wh_scale, shared_params, Ts, Ps, shared_modes, shared_bounds, inverse, outside_zero); \
    a.gy = static_cast<const T*>(gy); a.ldgy = ldgy; a.gld = static_cast<const T*>(glogdet);             \
    a.gx = static_cast<T*>(gx); a.ldgx = ldgx; a.gcond = static_cast<T*>(g_cond); a.ld_gcond = ld_gcond; \
    a.partial = static_cast<T*>(workspace);                                                              \
    return dispatch_bwd<T>(a, static_cast<T*>(g_shared), st);                                            \
  }
  if (dtype == VCNF_F32) VCNF_BWD(float) else VCNF_BWD(double)
#undef VCNF_BWD
}

int vcnf_gather_cols(const void* src, int64_t lds, void* dst, int64_t ldd, int64_t B, int n, const int32_t* cols,
                     int dtype, void* stream) {
  if (!src || !dst || !cols || B < 0 || n <= 0 || n > 1024 || (dtype != VCNF_F32 && dtype != VCNF_F64))
    return VCNF_E_BADARG;
  if (B == 0) return 0;
  const int64_t total = B * n;
  const int sms = sm_count() > 0 ? sm_count() : VCNF_SM_COUNT_DEFAULT;
  int64_t grid = (total + 255) / 256;
  if (grid > static_cast<int64_t>(sms) * 16) grid = static_cast<int64_t>(sms) * 16;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == VCNF_F32)
    gather_cols_kernel<float><<<static_cast<unsigned>(grid), 256, 0, st>>>(static_cast<const float*>(src), lds,
                                                                         static_cast<float*>(dst), ldd, B, n, cols);
  else
    gather_cols_kernel<double><<<static_cast<unsigned>(grid), 256, 0, st>>>(static_cast<const double*>(src), lds,
                                                                          static_cast<double*>(dst), ldd, B, n, cols);
  return launch_status();
}

int vcnf_scatter_cols(const void* src, int64_t lds, void* dst, int64_t ldd, int64_t B, int n, int D_dst,
                      const int32_t* cols, int accumulate, int dtype, void* stream) {
  if (!src || !dst || !cols || B < 0 || n <= 0 || D_dst < n || D_dst > 1024 ||
      (dtype != VCNF_F32 && dtype != VCNF_F64))
    return VCNF_E_BADARG;
  if (B == 0) return 0;
  const int64_t total = B * D_dst;
  const int sms = sm_count() > 0 ? sm_count() : VCNF_SM_COUNT_DEFAULT;
  int64_t grid = (total + 255) / 256;
  if (grid > static_cast<int64_t>(sms) * 16) grid = static_cast<int64_t>(sms) * 16;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == VCNF_F32)
    scatter_cols_kernel<float><<<static_cast<unsigned>(grid), 256, 0, st>>>(
        static_cast<const float*>(src), lds, static_cast<float*>(dst), ldd, B, n, D_dst, cols, accumulate);
  else
    scatter_cols_kernel<double><<<static_cast<unsigned>(grid), 256, 0, st>>>(
        static_cast<const double*>(src), lds, static_cast<double*>(dst), ldd, B, n, D_dst, cols, accumulate);
  return launch_status();
}

}  // extern "C"
