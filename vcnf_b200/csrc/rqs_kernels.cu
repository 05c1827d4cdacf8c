// Rational-quadratic spline LAYER kernels (forward / inverse / backward) for sm_100a.
//
// One launch = one whole coupling or autoregressive spline layer after its conditioner:
// split (index_select), per-element spline with conditional or batch-shared parameters,
// merge (scatter) and the per-sample log|det| reduction -- the ~600 ATen ops / 16 `nonzero`
// syncs of normflow/utils/splines.py:20-193 + flows/neural_spline/coupling.py:70-159.
//
// Data movement (HBM-bound, SURVEY.md 8d: s*(T*P + 2D + 1) bytes per sample):
//   * a tile is R consecutive samples.  Its parameter block [R x Tc*Pc], its x rows [R x D] (and
//     in the backward its gy rows) are CONTIGUOUS spans in HBM; each is pulled into shared memory
//     by the TMA engine with one 1-D bulk async copy (cp.async.bulk ... mbarrier::complete_tx)
//     into a 2-stage ring, so the copies of tile i+1 overlap the math of tile i and no registers
//     are spent on staging;
//   * the math runs in passes in which EVERY thread does the same kind of work (no divergence):
//       pass A: conditional elements, index space rows x Tc; thread e reads its P parameters from
//               shared memory at stride P (P = 3K-1 odd -> bank-conflict free);
//       pass B: batch-shared / pass-through elements; the knots of batch-shared parameters are
//               built ONCE per CTA into a shared-memory table (rqs_math.cuh: rqs_build_table);
//   * outputs (y, per-element log-det terms; gx, parameter gradients) are assembled in shared
//     memory and leave as full coalesced rows / one bulk store per tile;
//   * log|det|: half-warp per sample, shuffle tree (fixed order -> deterministic);
//   * persistent grid: #CTAs = SMs x resident CTAs, each walks tiles blockIdx.x, +gridDim.x, ...
//   * batch-shared parameter gradients: every thread owns a private accumulator table in shared
//     memory (no atomics), reduced per CTA in a fixed order, chained back to the raw parameters
//     once per CTA and finished by rqs_partial_reduce_kernel (deterministic two-stage reduction).
#include <cstdlib>
#include "common.cuh"
#include "rqs_math.cuh"
#include "rqs_fast.cuh"
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
  int R; int64_t num_tiles;
  int bulk_p, bulk_x, bulk_gy, bulk_gp;   // base pointer + leading dimension allow 1-D bulk copies
  int prow, grow;   // row pitch (elements) of the staged parameter / parameter-gradient tiles: Tc*Pc, or the
                    // caller's padded leading dimension (a 16 B multiple pitch keeps rows TMA-copyable)
  // backward only
  const T* gy; int64_t ldgy; const T* gld;
  T* gx; int64_t ldgx;
  T* gcond; int64_t ld_gcond;
  T* partial;     // [gridDim.x, Ts*Ps]
};

__host__ __device__ inline size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
__host__ __device__ inline int64_t imin64(int64_t a, int64_t b) { return a < b ? a : b; }

// Shared-memory carve-up, computed identically on host (size) and device (pointers).
struct RqsSmem {
  size_t bars, pst, pst_stride, xst, xst_stride, gyst, outs, lads, gst, tab, acc, tabg, meta, total;
};
template <typename T>
__host__ __device__ inline RqsSmem rqs_smem_layout(int R, int D, int Tc, int Pc, int Ts, int K, bool bwd, int nthreads,
                                                   int prow, int grow) {
  RqsSmem m;
  size_t o = 0;
  m.bars = o; o += 128;
  m.pst_stride = round_up(static_cast<size_t>(R) * prow * sizeof(T), 128);
  m.pst = o; o += 2 * m.pst_stride;
  m.xst_stride = round_up(static_cast<size_t>(R) * D * sizeof(T), 128);
  m.xst = o; o += 2 * m.xst_stride;
  m.gyst = o; if (bwd) o += 2 * m.xst_stride;
  m.outs = o; o += m.xst_stride;                       // y tile (fwd) / gx tile (bwd)
  m.lads = o; if (!bwd) o += m.xst_stride;             // per-element log-det terms (fwd)
  m.gst = o; if (bwd) o += round_up(static_cast<size_t>(R) * grow * sizeof(T), 128);   // conditional parameter gradients (bwd)
  const size_t TK = 3 * (static_cast<size_t>(K) + 1);
  m.tab = o; o += round_up(Ts * TK * sizeof(T), 128);
  m.acc = o; if (bwd && Ts > 0) o += round_up(static_cast<size_t>(nthreads / Ts) * Ts * TK * sizeof(T), 128);
  m.tabg = o; if (bwd) o += round_up(Ts * TK * sizeof(T), 128);
  // meta: cond_col[Tc] | other_col[D] | other_map[D] | cmode[Tc] | smode[Ts] | shared_col[Ts] (int32),
  //       cbound[Tc] | sbound[Ts] (T)
  m.meta = o; o += round_up((2 * static_cast<size_t>(Tc) + 2 * D + 2 * Ts) * sizeof(int32_t), 16) +
                   round_up((static_cast<size_t>(Tc) + Ts) * sizeof(T), 16);
  m.total = o;
  return m;
}

template <typename T>
struct RqsMeta {
  int32_t* cond_col; int32_t* other_col; int32_t* other_map; int32_t* cmode; int32_t* smode;
  int32_t* shared_col;
  T* cbound; T* sbound;
  int n_other;
};

// Column lists, tail modes and bounds into shared memory; knot tables of the batch-shared features.
template <typename T, int KMAX>
__device__ __forceinline__ RqsMeta<T> rqs_setup(const RqsLayerArgs<T>& a, int K, unsigned char* smem,
                                                const RqsSmem& L, int* n_other_s) {
  RqsMeta<T> m;
  int32_t* ip = reinterpret_cast<int32_t*>(smem + L.meta);
  m.cond_col = ip; ip += a.Tc;
  m.other_col = ip; ip += a.D;
  m.other_map = ip; ip += a.D;
  m.cmode = ip; ip += a.Tc;
  m.smode = ip; ip += a.Ts;
  m.shared_col = ip; ip += a.Ts;
  T* tp = reinterpret_cast<T*>(smem + L.meta + round_up((2 * static_cast<size_t>(a.Tc) + 2 * a.D + 2 * a.Ts) * sizeof(int32_t), 16));
  m.cbound = tp; tp += a.Tc;
  m.sbound = tp;
  const int tid = threadIdx.x;
  if (tid == 0) {
    int n = 0;
    for (int j = 0; j < a.Ts; ++j) m.shared_col[j] = -1;
    for (int c = 0; c < a.D; ++c) {
      const int cm = a.col_map[c];
      if ((cm >> 16) == VCNF_COL_COND) m.cond_col[cm & 0xffff] = c;
      else {
        m.other_col[n] = c; m.other_map[n] = cm; ++n;
        if ((cm >> 16) == VCNF_COL_SHARED) m.shared_col[cm & 0xffff] = c;
      }
    }
    *n_other_s = n;
  }
  for (int j = tid; j < a.Tc; j += blockDim.x) { m.cmode[j] = a.cmodes[j]; m.cbound[j] = a.cbounds[j]; }
  for (int j = tid; j < a.Ts; j += blockDim.x) { m.smode[j] = a.smodes[j]; m.sbound[j] = a.sbounds[j]; }
  T* tab = reinterpret_cast<T*>(smem + L.tab);
  const int TK = 3 * (K + 1);
  for (int j = tid; j < a.Ts; j += blockDim.x)
    rqs_build_table<T, KMAX>(a.shared + static_cast<size_t>(j) * a.Ps, K, a.smodes[j], a.sbounds[j], tab + j * TK);
  __syncthreads();
  m.n_other = *n_other_s;
  return m;
}

// ---- tile staging: TMA bulk copies when the span is contiguous and 16 B granular, else cooperative
template <typename T>
__device__ __forceinline__ bool span_bulk(int flag, int64_t elems) {
  return flag && ((static_cast<size_t>(elems) * sizeof(T)) % 16 == 0);
}

template <typename T, bool BWD>
__device__ __forceinline__ void stage_issue(const RqsLayerArgs<T>& a, int64_t t, T* pst, T* xst, T* gyst,
                                            uint64_t* bar) {
  if (threadIdx.x != 0) return;
  const int64_t rows = imin64(a.R, a.B - t * a.R);
  const int64_t np = rows * a.prow, nx = rows * a.D;
  const bool bp = a.Tc > 0 && span_bulk<T>(a.bulk_p, np);
  const bool bx = span_bulk<T>(a.bulk_x, nx);
  const bool bg = BWD && span_bulk<T>(a.bulk_gy, nx);
  const uint32_t bytes = static_cast<uint32_t>(((bp ? np : 0) + (bx ? nx : 0) + (bg ? nx : 0)) * sizeof(T));
  if (bytes == 0) return;
  mbar_expect_tx(bar, bytes);
  if (bp) bulk_g2s(pst, a.cond + t * a.R * a.ld_cond, static_cast<uint32_t>(np * sizeof(T)), bar);
  if (bx) bulk_g2s(xst, a.x + t * a.R * a.ldx, static_cast<uint32_t>(nx * sizeof(T)), bar);
  if (bg) bulk_g2s(gyst, a.gy + t * a.R * a.ldgy, static_cast<uint32_t>(nx * sizeof(T)), bar);
}

// Returns true when the barrier of this stage was used (its phase flipped).
template <typename T, bool BWD>
__device__ __forceinline__ bool stage_wait(const RqsLayerArgs<T>& a, int64_t t, T* pst, T* xst, T* gyst,
                                           uint64_t* bar, uint32_t parity) {
  const int64_t row0 = t * a.R;
  const int64_t rows = imin64(a.R, a.B - row0);
  const int64_t np = rows * a.prow, nx = rows * a.D;
  const bool bp = a.Tc > 0 && span_bulk<T>(a.bulk_p, np);
  const bool bx = span_bulk<T>(a.bulk_x, nx);
  const bool bg = BWD && span_bulk<T>(a.bulk_gy, nx);
  if (a.Tc > 0 && !bp) {
    const int n = a.Tc * a.Pc;
    for (int64_t e = threadIdx.x; e < rows * n; e += blockDim.x) {
      const int64_t r = e / n;
      pst[r * a.prow + (e - r * n)] = a.cond[(row0 + r) * a.ld_cond + (e - r * n)];
    }
  }
  if (!bx) {
    for (int64_t e = threadIdx.x; e < nx; e += blockDim.x) {
      const int64_t r = e / a.D;
      xst[e] = a.x[(row0 + r) * a.ldx + (e - r * a.D)];
    }
  }
  if (BWD && !bg) {
    for (int64_t e = threadIdx.x; e < nx; e += blockDim.x) {
      const int64_t r = e / a.D;
      gyst[e] = a.gy[(row0 + r) * a.ldgy + (e - r * a.D)];
    }
  }
  const bool used = bp || bx || bg;
  if (used) mbar_wait(bar, parity);
  return used;
}

template <typename T, int KMAX, bool EXACT>
__global__ void __launch_bounds__(kThreads) rqs_layer_fwd_kernel(const RqsLayerArgs<T> a) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ int n_other_s;
  const int K = EXACT ? KMAX : a.K;
  const int TK = 3 * (K + 1);
  const RqsSmem L = rqs_smem_layout<T>(a.R, a.D, a.Tc, a.Pc, a.Ts, K, false, kThreads, a.prow, a.grow);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.bars);
  T* ys = reinterpret_cast<T*>(smem + L.outs);
  T* lads = reinterpret_cast<T*>(smem + L.lads);
  const T* tab = reinterpret_cast<const T*>(smem + L.tab);
  const int tid = threadIdx.x;
  if (tid == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    mbar_init_fence();
  }
  const RqsMeta<T> M = rqs_setup<T, KMAX>(a, K, smem, L, &n_other_s);   // ends with __syncthreads()
  const int D = a.D, Tc = a.Tc, Pc = a.Pc, No = M.n_other;

  int flag = 0;
  uint32_t par0 = 0u, par1 = 0u;
  int64_t tile = blockIdx.x;
  if (tile < a.num_tiles)
    stage_issue<T, false>(a, tile, reinterpret_cast<T*>(smem + L.pst), reinterpret_cast<T*>(smem + L.xst), nullptr, &bars[0]);
  for (int it = 0; tile < a.num_tiles; tile += gridDim.x, ++it) {
    const int s = it & 1;
    T* cur = reinterpret_cast<T*>(smem + L.pst + s * L.pst_stride);
    T* xs = reinterpret_cast<T*>(smem + L.xst + s * L.xst_stride);
    const int64_t next = tile + gridDim.x;
    if (next < a.num_tiles)
      stage_issue<T, false>(a, next, reinterpret_cast<T*>(smem + L.pst + (s ^ 1) * L.pst_stride),
                            reinterpret_cast<T*>(smem + L.xst + (s ^ 1) * L.xst_stride), nullptr, &bars[s ^ 1]);
    const int64_t row0 = tile * a.R;
    const int rows = static_cast<int>(imin64(a.R, a.B - row0));
    if (stage_wait<T, false>(a, tile, cur, xs, nullptr, &bars[s], s ? par1 : par0)) { if (s) par1 ^= 1u; else par0 ^= 1u; }
    __syncthreads();

    // pass A: conditional elements (e = r*Tc + j; parameters at cur + e*Pc)
    const int nA = rows * Tc;
    for (int e = tid; e < nA; e += kThreads) {
      const int r = e / Tc;
      const int j = e - r * Tc;
      const int c = M.cond_col[j];
      const T v = xs[r * D + c];
      RqsElemCfg cfg{K, M.cmode[j], a.inverse, a.outside_zero};
      T out, lad;
      int bin;
      rqs_element<T, KMAX>(cur + static_cast<size_t>(r) * a.prow + j * Pc, v, M.cbound[j], a.wh_scale, cfg, &out, &lad, &bin, &flag);
      ys[r * D + c] = out;
      lads[r * D + c] = lad;
      if (a.bin_out) a.bin_out[(row0 + r) * D + c] = static_cast<int8_t>(bin);
    }
    // pass B: batch-shared and pass-through elements
    const int nB = rows * No;
    for (int e = tid; e < nB; e += kThreads) {
      const int r = e / No;
      const int j = e - r * No;
      const int c = M.other_col[j];
      const int cm = M.other_map[j];
      const T v = xs[r * D + c];
      T out = v, lad = T(0);
      int bin = -2;
      if ((cm >> 16) == VCNF_COL_SHARED) {
        const int idx = cm & 0xffff;
        RqsElemCfg cfg{K, M.smode[idx], a.inverse, a.outside_zero};
        rqs_table_element<T, KMAX>(tab + idx * TK, v, M.sbound[idx], cfg, &out, &lad, &bin, &flag);
      }
      ys[r * D + c] = out;
      lads[r * D + c] = lad;
      if (a.bin_out) a.bin_out[(row0 + r) * D + c] = static_cast<int8_t>(bin);
    }
    __syncthreads();

    // pass C: coalesced row stores + per-sample log|det| (half-warp per sample, shuffle tree)
    const int nY = rows * D;
    if (a.ldy == D) {
      T* dst = a.y + row0 * D;
      for (int e = tid; e < nY; e += kThreads) dst[e] = ys[e];
    } else {
      for (int e = tid; e < nY; e += kThreads) {
        const int r = e / D;
        a.y[(row0 + r) * a.ldy + (e - r * D)] = ys[e];
      }
    }
    const int hl = tid & 15;
    for (int r0 = (tid >> 5) * 2; r0 < rows; r0 += (kThreads >> 5) * 2) {
      const int r = r0 + ((tid >> 4) & 1);
      T acc = T(0);
      if (r < rows)
        for (int c = hl; c < D; c += 16) acc += lads[r * D + c];
#pragma unroll
      for (int o = 8; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
      if (r < rows && hl == 0) {
        if (a.accumulate) acc += a.logdet[row0 + r];
        a.logdet[row0 + r] = acc;
      }
    }
  }
  if (flag && a.err_flag) atomicOr(a.err_flag, flag);
}

// Backward: recomputes the bin record from (x, params) and applies the hand-derived adjoint
// (rqs_math.cuh).  Conditional parameter gradients of a tile are assembled in shared memory and
// leave through one bulk store (TMA engine) per tile.
template <typename T, int KMAX, bool EXACT>
__global__ void __launch_bounds__(kThreads) rqs_layer_bwd_kernel(const RqsLayerArgs<T> a) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ int n_other_s;
  const int K = EXACT ? KMAX : a.K;
  const int TK = 3 * (K + 1);
  const RqsSmem L = rqs_smem_layout<T>(a.R, a.D, a.Tc, a.Pc, a.Ts, K, true, kThreads, a.prow, a.grow);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.bars);
  T* gxs = reinterpret_cast<T*>(smem + L.outs);
  T* gst = reinterpret_cast<T*>(smem + L.gst);
  const T* tab = reinterpret_cast<const T*>(smem + L.tab);
  T* acc = reinterpret_cast<T*>(smem + L.acc);
  T* tabg = reinterpret_cast<T*>(smem + L.tabg);
  const int tid = threadIdx.x;
  if (tid == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    mbar_init_fence();
  }
  const int G = a.Ts > 0 ? kThreads / a.Ts : 0;          // thread groups of pass B2 (one feature per thread)
  for (int i = tid; i < G * a.Ts * TK; i += kThreads) acc[i] = T(0);
  const RqsMeta<T> M = rqs_setup<T, KMAX>(a, K, smem, L, &n_other_s);   // ends with __syncthreads()
  const int D = a.D, Tc = a.Tc, Pc = a.Pc, No = M.n_other, Ts = a.Ts;

  uint32_t par0 = 0u, par1 = 0u;
  int64_t tile = blockIdx.x;
  if (tile < a.num_tiles)
    stage_issue<T, true>(a, tile, reinterpret_cast<T*>(smem + L.pst), reinterpret_cast<T*>(smem + L.xst),
                         reinterpret_cast<T*>(smem + L.gyst), &bars[0]);
  for (int it = 0; tile < a.num_tiles; tile += gridDim.x, ++it) {
    const int s = it & 1;
    T* cur = reinterpret_cast<T*>(smem + L.pst + s * L.pst_stride);
    T* xs = reinterpret_cast<T*>(smem + L.xst + s * L.xst_stride);
    T* gys = reinterpret_cast<T*>(smem + L.gyst + s * L.xst_stride);
    const int64_t next = tile + gridDim.x;
    if (next < a.num_tiles)
      stage_issue<T, true>(a, next, reinterpret_cast<T*>(smem + L.pst + (s ^ 1) * L.pst_stride),
                           reinterpret_cast<T*>(smem + L.xst + (s ^ 1) * L.xst_stride),
                           reinterpret_cast<T*>(smem + L.gyst + (s ^ 1) * L.xst_stride), &bars[s ^ 1]);
    const int64_t row0 = tile * a.R;
    const int rows = static_cast<int>(imin64(a.R, a.B - row0));
    if (stage_wait<T, true>(a, tile, cur, xs, gys, &bars[s], s ? par1 : par0)) { if (s) par1 ^= 1u; else par0 ^= 1u; }
    __syncthreads();

    // pass A: conditional elements
    const int nA = rows * Tc;
    for (int e = tid; e < nA; e += kThreads) {
      const int r = e / Tc;
      const int j = e - r * Tc;
      const int c = M.cond_col[j];
      RqsElemCfg cfg{K, M.cmode[j], a.inverse, a.outside_zero};
      T gv;
      rqs_element_bwd<T, KMAX>(cur + static_cast<size_t>(r) * a.prow + j * Pc, xs[r * D + c], M.cbound[j], a.wh_scale, cfg,
                               gys[r * D + c], a.gld[row0 + r], &gv, gst + static_cast<size_t>(r) * a.grow + j * Pc);
      gxs[r * D + c] = gv;
    }
    // pass B1: pass-through columns;  pass B2: batch-shared columns, thread <-> feature fixed so that
    // the thread-private accumulator table needs no atomics
    if (Ts == 0) {
      const int nB = rows * No;
      for (int e = tid; e < nB; e += kThreads) {
        const int r = e / No;
        const int c = M.other_col[e - r * No];
        gxs[r * D + c] = gys[r * D + c];
      }
    } else {
      for (int j = 0; j < No; ++j) {
        if ((M.other_map[j] >> 16) == VCNF_COL_SHARED) continue;
        const int c = M.other_col[j];
        for (int r = tid; r < rows; r += kThreads) gxs[r * D + c] = gys[r * D + c];
      }
      if (tid < G * Ts) {
        const int g = tid / Ts;
        const int jo = tid - g * Ts;                       // this thread's batch-shared feature
        const int c = M.shared_col[jo];
        if (c >= 0) {
          RqsElemCfg cfg{K, M.smode[jo], a.inverse, a.outside_zero};
          for (int r = g; r < rows; r += G) {
            T gv;
            rqs_table_element_bwd<T, KMAX>(tab + jo * TK, xs[r * D + c], M.sbound[jo], cfg, gys[r * D + c],
                                           a.gld[row0 + r], &gv, acc + static_cast<size_t>(tid) * TK);
            gxs[r * D + c] = gv;
          }
        }
      }
    }
    const int64_t np = static_cast<int64_t>(rows) * a.grow;
    const bool bulk_store = Tc > 0 && span_bulk<T>(a.bulk_gp, np);
    if (bulk_store) fence_async_smem();
    __syncthreads();

    // pass C: coalesced gx rows, parameter-gradient tile
    const int nY = rows * D;
    if (a.ldgx == D) {
      T* dst = a.gx + row0 * D;
      for (int e = tid; e < nY; e += kThreads) dst[e] = gxs[e];
    } else {
      for (int e = tid; e < nY; e += kThreads) {
        const int r = e / D;
        a.gx[(row0 + r) * a.ldgx + (e - r * D)] = gxs[e];
      }
    }
    if (Tc > 0) {
      if (bulk_store) {
        if (tid == 0) {
          bulk_s2g(a.gcond + row0 * a.ld_gcond, gst, static_cast<uint32_t>(np * sizeof(T)));
          bulk_commit();
          bulk_wait_read0();                 // gst is rewritten by the next tile's pass A
        }
      } else {
        const int n = Tc * Pc;
        for (int64_t e = tid; e < static_cast<int64_t>(rows) * n; e += kThreads) {
          const int64_t r = e / n;
          a.gcond[(row0 + r) * a.ld_gcond + (e - r * n)] = gst[r * a.grow + (e - r * n)];
        }
      }
    }
    // (the __syncthreads() after the next stage_wait orders these reads before the next tile's writes)
  }
  if (Ts > 0) {
    __syncthreads();
    for (int q = tid; q < Ts * TK; q += kThreads) {
      const int j = q / TK, i = q - j * TK;
      T sum = T(0);
      for (int g = 0; g < G; ++g) sum += acc[(static_cast<size_t>(g) * Ts + j) * TK + i];
      tabg[q] = sum;
    }
    __syncthreads();
    for (int j = tid; j < Ts; j += kThreads)
      rqs_table_to_raw_grad<T, KMAX>(a.shared + static_cast<size_t>(j) * a.Ps, K, M.smode[j], M.sbound[j],
                                     tabg + j * TK, a.partial + (static_cast<size_t>(blockIdx.x) * Ts + j) * a.Ps);
  }
}

// =====================================================================================================
// fp32 / K = 8 fast path (rqs_fast.cuh) -- the default configuration of every spline wrapper.
//
// Differences from the generic kernels above:
//   * WARP-PRIVATE pipelines: a tile is R = 8 * rpw rows and warp w owns rows [w*rpw, (w+1)*rpw) through
//     all passes (conditional elements, batch-shared elements, row stores + log-det), so the only
//     synchronisation inside the tile loop is __syncwarp plus two mbarriers per stage: `full` (TMA
//     transaction bytes) and `empty` (one arrival per warp when its reads of the stage are done; the
//     producer lane waits on it before re-arming the stage).  No CTA-wide barrier on the fast path.
//   * element math evaluated as packed f32x2 streams (both spline axes per instruction) with MUFU
//     approximations; knot records {X, Y, D} for the batch-shared parameters, searched by bisection;
//   * row / column of an element from a multiply-shift instead of an integer division.
// =====================================================================================================
constexpr int kWarps = kThreads / 32;

// Batch-shared features in the backward: lane <-> (row group g, feature jo) is fixed for the whole launch so
// that every lane owns a private gradient table (no atomics): 32/Ts row groups per warp when Ts <= 32.
__host__ __device__ inline int fast_slots(int Ts) { return Ts <= 0 ? 0 : (Ts <= 32 ? (32 / Ts) * Ts : Ts); }

struct FastSmem {
  size_t bars, pst, pst_stride, xst, xst_stride, gyst, glst, glst_stride, outs, lads, gst, gst_stride, tab, rec, acc, tabg, meta, total;
};
__host__ __device__ inline FastSmem fast_smem_layout(int R, int D, int Tc, int Ts, bool bwd, int prow, int grow,
                                                     int nst) {
  FastSmem m;
  size_t o = 0;
  m.bars = o; o += 128;
  m.pst_stride = round_up(static_cast<size_t>(R) * prow * sizeof(float), 128);
  m.pst = o; o += nst * m.pst_stride;
  m.xst_stride = round_up(static_cast<size_t>(R) * D * sizeof(float), 128);
  m.xst = o; o += nst * m.xst_stride;
  m.gyst = o; if (bwd) o += nst * m.xst_stride;
  m.glst_stride = round_up(static_cast<size_t>(R) * sizeof(float), 128);      // upstream log-det gradients of the tile
  m.glst = o; if (bwd) o += nst * m.glst_stride;
  m.outs = o; o += m.xst_stride;
  m.lads = o; if (!bwd) o += m.xst_stride;
  m.gst_stride = round_up(static_cast<size_t>(R) * grow * sizeof(float), 128);
  m.gst = o; if (bwd) o += m.gst_stride;
  const size_t TK = 3 * 9;
  m.tab = o;
  m.rec = o; o += round_up(static_cast<size_t>(Ts) * 9 * sizeof(fast::Knot4), 128);
  // backward: one private accumulator table per (warp, slot); a slot is a (row group, shared feature) pair
  m.acc = o; if (bwd && Ts > 0) o += round_up(static_cast<size_t>(kWarps) * fast_slots(Ts) * TK * sizeof(float), 128);
  m.tabg = o; if (bwd) o += round_up(Ts * TK * sizeof(float), 128);
  m.meta = o; o += round_up((2 * static_cast<size_t>(Tc) + 2 * D + 2 * Ts) * sizeof(int32_t), 16) +
                   round_up((static_cast<size_t>(Tc) + Ts) * sizeof(float), 16);
  m.total = o;
  return m;
}

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// Consumer-side wait: back off between polls so that a warp waiting for its tile does not take issue slots
// from the warps that are computing (ncu: 10 % of the executed instructions were polls).
__device__ __forceinline__ void mbar_wait_backoff(uint64_t* bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  while (!mbar_try_wait(bar, parity)) __nanosleep(100);
}

// e / n for 0 <= e < 2^20 / n  (magic = 2^20 / n + 1), or a real division when the host passed magic 0
__device__ __forceinline__ int fast_div(int e, int n, uint32_t magic) {
  return magic ? static_cast<int>((static_cast<uint32_t>(e) * magic) >> 20) : e / n;
}

// Column lists, tail modes, bounds and the knot records of the batch-shared features, built by all threads
// in parallel (the generic rqs_setup walks the columns on one thread and runs the libm softmax per feature:
// 7 us of a 38 us launch under ncu).  The outs tile is free before the tile loop and holds the raw column map.
template <int K>
__device__ __forceinline__ RqsMeta<float> fast_setup(const RqsLayerArgs<float>& a, unsigned char* smem,
                                                     const FastSmem& L, fast::Knot4* rec, int* n_pass) {
  RqsMeta<float> m;
  int32_t* ip = reinterpret_cast<int32_t*>(smem + L.meta);
  m.cond_col = ip; ip += a.Tc;
  m.other_col = ip; ip += a.D;
  m.other_map = ip; ip += a.D;
  m.cmode = ip; ip += a.Tc;
  m.smode = ip; ip += a.Ts;
  m.shared_col = ip; ip += a.Ts;
  float* tp = reinterpret_cast<float*>(smem + L.meta + round_up((2 * static_cast<size_t>(a.Tc) + 2 * a.D + 2 * a.Ts) * sizeof(int32_t), 16));
  m.cbound = tp; tp += a.Tc;
  m.sbound = tp;
  int32_t* raw = reinterpret_cast<int32_t*>(smem + L.outs);
  const int tid = threadIdx.x;
  for (int c = tid; c < a.D; c += kThreads) raw[c] = a.col_map[c];
  for (int j = tid; j < a.Tc; j += kThreads) { m.cmode[j] = a.cmodes[j]; m.cbound[j] = a.cbounds[j]; }
  // knot records of the batch-shared features: one thread per feature for the two softmax axes, and one thread
  // per pair of knot derivatives (softplus), so that the longest dependent chain is one axes evaluation
  constexpr int kPairs = (K + 2) / 2;
  for (int w = tid; w < a.Ts * (1 + kPairs); w += kThreads) {
    const int j = w / (1 + kPairs), part = w - j * (1 + kPairs);
    const int mode = a.smodes[j];
    const float bound = a.sbounds[j];
    const float* prm = a.shared + static_cast<size_t>(j) * a.Ps;
    fast::Knot4* rj = rec + j * (K + 1);
    if (part == 0) {
      m.smode[j] = mode; m.sbound[j] = bound; m.shared_col[j] = -1;
      fast::P2 p[K], kn[K + 1];
      fast::axes<K>(prm, 1.f, bound, rqs_lo(bound, mode), p, kn);
#pragma unroll
      for (int k = 0; k <= K; ++k) { rj[k].x = kn[k].x; rj[k].y = kn[k].y; }
    } else {
      const int k = 2 * (part - 1);
      const float edge = RqsConst<float>::edge();
      int sidx;
      const float ra = rqs_raw_deriv(prm + 2 * K, k, K, mode, edge, &sidx);
      const float rb = (k + 1 <= K) ? rqs_raw_deriv(prm + 2 * K, k + 1, K, mode, edge, &sidx) : edge;
      fast::P2 e;
      const fast::P2 sp = fast::softplus2(fast::P2{ra, rb}, &e);
      rj[k].d = RqsConst<float>::min_deriv() + sp.x;
      if (k + 1 <= K) rj[k + 1].d = RqsConst<float>::min_deriv() + sp.y;
    }
  }
  __syncthreads();
  for (int c = tid; c < a.D; c += kThreads) {
    const int cm = raw[c];
    if ((cm >> 16) == VCNF_COL_COND) {
      m.cond_col[cm & 0xffff] = c;
    } else {
      // "other" list: pass-through columns first, batch-shared columns after them (each in column order)
      const bool shared = (cm >> 16) == VCNF_COL_SHARED;
      int n = 0;
      for (int q = 0; q < a.D; ++q) {
        const int k = raw[q] >> 16;
        if (k == VCNF_COL_COND) continue;
        const bool qs = k == VCNF_COL_SHARED;
        n += (qs == shared) ? (q < c ? 1 : 0) : (shared ? 1 : 0);
      }
      m.other_col[n] = c; m.other_map[n] = cm;
      if (shared) m.shared_col[cm & 0xffff] = c;
    }
  }
  int np = 0;
  for (int q = 0; q < a.D; ++q) np += ((raw[q] >> 16) == VCNF_COL_PASS) ? 1 : 0;
  __syncthreads();
  m.n_other = a.D - a.Tc;
  *n_pass = np;
  return m;
}

struct FastExtra { uint32_t magic_tc, magic_no, magic_ts; };

template <bool INV, int NST>
__global__ void __launch_bounds__(kThreads) rqs_layer_fwd_fast_kernel(const RqsLayerArgs<float> a, const FastExtra fx) {
  constexpr int K = 8;
  extern __shared__ __align__(128) unsigned char smem[];
  const FastSmem L = fast_smem_layout(a.R, a.D, a.Tc, a.Ts, false, a.prow, a.grow, NST);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + L.bars);
  uint64_t* empty = full + NST;
  float* ys = reinterpret_cast<float*>(smem + L.outs);
  float* lads = reinterpret_cast<float*>(smem + L.lads);
  fast::Knot4* rec = reinterpret_cast<fast::Knot4*>(smem + L.rec);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  int64_t tile = blockIdx.x;
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < NST; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], kWarps); }
    mbar_init_fence();
    // the first tiles are requested before the setup so that their HBM latency hides behind it
#pragma unroll
    for (int s = 0; s < NST - 1; ++s) {
      const int64_t t = tile + static_cast<int64_t>(s) * gridDim.x;
      if (t < a.num_tiles)
        stage_issue<float, false>(a, t, reinterpret_cast<float*>(smem + L.pst + s * L.pst_stride),
                                  reinterpret_cast<float*>(smem + L.xst + s * L.xst_stride), nullptr, &full[s]);
    }
  }
  int n_pass;
  const RqsMeta<float> M = fast_setup<K>(a, smem, L, rec, &n_pass);   // ends with __syncthreads()
  (void)n_pass;
  const int D = a.D, Tc = a.Tc, Pc = a.Pc, No = M.n_other;
  const int rpw = a.R / kWarps;
  const int r_lo = warp * rpw;
  // a full tile whose spans all go through the TMA engine needs nothing but the mbarrier wait
  const bool full_tiles_bulk = (Tc == 0 || span_bulk<float>(a.bulk_p, static_cast<int64_t>(a.R) * a.prow)) &&
                               span_bulk<float>(a.bulk_x, static_cast<int64_t>(a.R) * D);
  // pass C fast form: at most two output elements and one log-det row pair per lane
  const bool small_c = rpw * D <= 64 && D <= 32 && rpw <= 2 && a.ldy == D;
  float* const y_run = a.y + (static_cast<int64_t>(blockIdx.x) * a.R + r_lo) * D;      // + it * y_step per tile
  const int64_t y_step = static_cast<int64_t>(gridDim.x) * a.R * D;
  float* const ld_run = a.logdet + static_cast<int64_t>(blockIdx.x) * a.R + r_lo;
  const int64_t ld_step = static_cast<int64_t>(gridDim.x) * a.R;

  int flag = 0;
  for (int it = 0; tile < a.num_tiles; tile += gridDim.x, ++it) {
    const int s = it % NST;
    if (tid == 0) {                                    // producer lane: re-arm the stage read in iteration it-1
      const int64_t nt = tile + static_cast<int64_t>(NST - 1) * gridDim.x;
      if (nt < a.num_tiles) {
        const int ps = (it + NST - 1) % NST;
        if (it >= 1) mbar_wait(&empty[ps], static_cast<uint32_t>(((it - 1) / NST) & 1));
        stage_issue<float, false>(a, nt, reinterpret_cast<float*>(smem + L.pst + ps * L.pst_stride),
                                  reinterpret_cast<float*>(smem + L.xst + ps * L.xst_stride), nullptr, &full[ps]);
      }
    }
    __syncwarp();
    float* cur = reinterpret_cast<float*>(smem + L.pst + s * L.pst_stride);
    float* xs = reinterpret_cast<float*>(smem + L.xst + s * L.xst_stride);
    const int64_t row0 = tile * a.R;
    const int rows = static_cast<int>(imin64(a.R, a.B - row0));
    if (rows == a.R && full_tiles_bulk) {
      mbar_wait_backoff(&full[s], static_cast<uint32_t>((it / NST) & 1));
    } else {
      // spans the TMA engine could not take (ragged last tile, unaligned views) are copied cooperatively;
      // this is the only CTA-wide synchronisation and it is off the steady-state path
      __syncthreads();
      stage_wait<float, false>(a, tile, cur, xs, nullptr, &full[s], static_cast<uint32_t>((it / NST) & 1));
      __syncthreads();
    }
    const int nrow = rows - r_lo < rpw ? rows - r_lo : rpw;       // rows of this warp in this tile (may be <= 0)

    // pass A: conditional elements of this warp's rows
    const int nA = nrow * Tc;
    for (int e = lane; e < nA; e += 32) {
      const int rr = fast_div(e, Tc, fx.magic_tc);
      const int j = e - rr * Tc;
      const int r = r_lo + rr;
      const int c = M.cond_col[j];
      float out, lad;
      int bin;
      fast::element<K, INV>(cur + static_cast<size_t>(r) * a.prow + j * Pc, xs[r * D + c], M.cbound[j], a.wh_scale,
                            M.cmode[j], a.outside_zero, &out, &lad, &bin, &flag);
      ys[r * D + c] = out;
      lads[r * D + c] = lad;
      if (a.bin_out) a.bin_out[(row0 + r) * D + c] = static_cast<int8_t>(bin);
    }
    // pass B: batch-shared and pass-through elements
    const int nB = nrow * No;
    for (int e = lane; e < nB; e += 32) {
      const int rr = fast_div(e, No, fx.magic_no);
      const int j = e - rr * No;
      const int r = r_lo + rr;
      const int c = M.other_col[j];
      const int cm = M.other_map[j];
      const float v = xs[r * D + c];
      float out = v, lad = 0.f;
      int bin = -2;
      if ((cm >> 16) == VCNF_COL_SHARED) {
        const int idx = cm & 0xffff;
        fast::table_element<K, INV>(rec + idx * (K + 1), v, M.sbound[idx], a.outside_zero, &out, &lad, &bin, &flag);
      }
      ys[r * D + c] = out;
      lads[r * D + c] = lad;
      if (a.bin_out) a.bin_out[(row0 + r) * D + c] = static_cast<int8_t>(bin);
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[s]);             // this warp is done reading the stage

    // pass C: this warp's rows leave as one contiguous span; per-sample log|det| (half-warp per row)
    if (nrow > 0) {
      const int n = nrow * D;
      const float* src = ys + r_lo * D;
      if (small_c) {
        // common shapes (rpw * D <= 64, D <= 32, contiguous y): two predicated stores per lane, no loops and no
        // per-tile 64-bit index arithmetic beyond the running row pointer
        float* dst = y_run + static_cast<int64_t>(it) * y_step;
        if (lane < n) dst[lane] = src[lane];
        if (lane + 32 < n) dst[lane + 32] = src[lane + 32];
        const int hl = lane & 15, rr = lane >> 4;
        const float* lrow = lads + (r_lo + rr) * D;
        float acc = 0.f;
        if (rr < nrow) acc = (hl < D ? lrow[hl] : 0.f) + (hl + 16 < D ? lrow[hl + 16] : 0.f);
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (rr < nrow && hl == 0) {
          float* ldp = ld_run + static_cast<int64_t>(it) * ld_step + rr;
          *ldp = a.accumulate ? acc + *ldp : acc;
        }
      } else {
        if (a.ldy == D) {
          float* dst = a.y + (row0 + r_lo) * D;
#pragma unroll 1
          for (int e = lane; e < n; e += 32) dst[e] = src[e];
        } else {
          for (int rr = 0; rr < nrow; ++rr)
#pragma unroll 1
            for (int c = lane; c < D; c += 32) a.y[(row0 + r_lo + rr) * a.ldy + c] = src[rr * D + c];
        }
        const int hl = lane & 15;
#pragma unroll 1
        for (int rr0 = 0; rr0 < nrow; rr0 += 2) {
          const int rr = rr0 + (lane >> 4);
          float acc = 0.f;
          if (rr < nrow) {
            const float* lrow = lads + (r_lo + rr) * D;
#pragma unroll 1
            for (int c = hl; c < D; c += 16) acc += lrow[c];
          }
#pragma unroll
          for (int o = 8; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
          if (rr < nrow && hl == 0) {
            const int64_t g = row0 + r_lo + rr;
            a.logdet[g] = a.accumulate ? acc + a.logdet[g] : acc;
          }
        }
      }
    }
    __syncwarp();                                      // ys / lads rows are rewritten by other lanes next tile
  }
  if (flag && a.err_flag) atomicOr(a.err_flag, flag);
}

__device__ __forceinline__ void bulk_wait_read_all() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }

// Backward fast path: same warp-private pipeline as the forward.  Each warp owns its rows of the parameter-
// gradient tile and sends them to HBM with its own bulk store (drains while the warp works on the next tile's
// batch-shared elements); batch-shared gradients go to lane-private tables, summed in a fixed order at the end.
template <bool INV, int NST>
__global__ void __launch_bounds__(kThreads) rqs_layer_bwd_fast_kernel(const RqsLayerArgs<float> a, const FastExtra fx) {
  constexpr int K = 8, TK = 27;
  extern __shared__ __align__(128) unsigned char smem[];
  const FastSmem L = fast_smem_layout(a.R, a.D, a.Tc, a.Ts, true, a.prow, a.grow, NST);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + L.bars);
  uint64_t* empty = full + NST;
  float* gxs = reinterpret_cast<float*>(smem + L.outs);
  float* gst = reinterpret_cast<float*>(smem + L.gst);
  fast::Knot4* rec = reinterpret_cast<fast::Knot4*>(smem + L.rec);
  float* acc = reinterpret_cast<float*>(smem + L.acc);
  float* tabg = reinterpret_cast<float*>(smem + L.tabg);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int D = a.D, Tc = a.Tc, Pc = a.Pc, Ts = a.Ts;
  const bool full_tiles_bulk = (Tc == 0 || span_bulk<float>(a.bulk_p, static_cast<int64_t>(a.R) * a.prow)) &&
                               span_bulk<float>(a.bulk_x, static_cast<int64_t>(a.R) * D) &&
                               span_bulk<float>(a.bulk_gy, static_cast<int64_t>(a.R) * D);
  // the tile's upstream log-det gradients ride along with a fully TMA-staged tile
  const bool gl_bulk = full_tiles_bulk && (reinterpret_cast<uintptr_t>(a.gld) % 16 == 0) && (a.R % 4 == 0);
  auto issue = [&](int64_t t, int ps) {
    float* pd = reinterpret_cast<float*>(smem + L.pst + ps * L.pst_stride);
    float* xd = reinterpret_cast<float*>(smem + L.xst + ps * L.xst_stride);
    float* gd = reinterpret_cast<float*>(smem + L.gyst + ps * L.xst_stride);
    if (full_tiles_bulk && a.B - t * a.R >= a.R) {
      const uint32_t pb = Tc > 0 ? static_cast<uint32_t>(a.R * a.prow * sizeof(float)) : 0u;
      const uint32_t xb = static_cast<uint32_t>(a.R * D * sizeof(float));
      const uint32_t lb = gl_bulk ? static_cast<uint32_t>(a.R * sizeof(float)) : 0u;
      mbar_expect_tx(&full[ps], pb + 2u * xb + lb);
      if (pb) bulk_g2s(pd, a.cond + t * a.R * a.ld_cond, pb, &full[ps]);
      bulk_g2s(xd, a.x + t * a.R * a.ldx, xb, &full[ps]);
      bulk_g2s(gd, a.gy + t * a.R * a.ldgy, xb, &full[ps]);
      if (lb) bulk_g2s(smem + L.glst + ps * L.glst_stride, a.gld + t * a.R, lb, &full[ps]);
    } else {
      stage_issue<float, true>(a, t, pd, xd, gd, &full[ps]);
    }
  };
  int64_t tile = blockIdx.x;
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < NST; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], kWarps); }
    mbar_init_fence();
#pragma unroll
    for (int s = 0; s < NST - 1; ++s) {
      const int64_t t = tile + static_cast<int64_t>(s) * gridDim.x;
      if (t < a.num_tiles) issue(t, s);
    }
  }
  const int slots = fast_slots(Ts);
  for (int i = tid; i < kWarps * slots * TK; i += kThreads) acc[i] = 0.f;
  int n_pass;
  const RqsMeta<float> M = fast_setup<K>(a, smem, L, rec, &n_pass);   // ends with __syncthreads()
  const int No = M.n_other;
  const int rpw = a.R / kWarps;
  const int r_lo = warp * rpw;
  const int Gw = Ts > 0 ? (Ts <= 32 ? 32 / Ts : 1) : 0;      // row groups per warp of the batch-shared pass
  float* my_gst = gst + static_cast<size_t>(r_lo) * a.grow;
  const bool small_c = rpw * D <= 64 && a.ldgx == D;       // pass C: at most two input-gradient elements per lane
  float* const gx_run = a.gx + (static_cast<int64_t>(blockIdx.x) * a.R + r_lo) * D;
  const int64_t gx_step = static_cast<int64_t>(gridDim.x) * a.R * D;
  bool store_pending = false;                                // lane 0: a bulk store of my_gst may still be reading it

  for (int it = 0; tile < a.num_tiles; tile += gridDim.x, ++it) {
    const int s = it % NST;
    if (tid == 0) {
      const int64_t nt = tile + static_cast<int64_t>(NST - 1) * gridDim.x;
      if (nt < a.num_tiles) {
        const int ps = (it + NST - 1) % NST;
        if (it >= 1) mbar_wait(&empty[ps], static_cast<uint32_t>(((it - 1) / NST) & 1));
        issue(nt, ps);
      }
    }
    if (lane == 0 && store_pending) { bulk_wait_read_all(); store_pending = false; }
    __syncwarp();
    float* cur = reinterpret_cast<float*>(smem + L.pst + s * L.pst_stride);
    float* xs = reinterpret_cast<float*>(smem + L.xst + s * L.xst_stride);
    float* gys = reinterpret_cast<float*>(smem + L.gyst + s * L.xst_stride);
    const int64_t row0 = tile * a.R;
    const int rows = static_cast<int>(imin64(a.R, a.B - row0));
    if (rows == a.R && full_tiles_bulk) {
      mbar_wait_backoff(&full[s], static_cast<uint32_t>((it / NST) & 1));
    } else {
      __syncthreads();
      stage_wait<float, true>(a, tile, cur, xs, gys, &full[s], static_cast<uint32_t>((it / NST) & 1));
      __syncthreads();
    }
    const int nrow = rows - r_lo < rpw ? rows - r_lo : rpw;
    // upstream log-det gradients of this tile: staged with the tile, or straight from HBM on the slow path
    const float* gl = (rows == a.R && gl_bulk) ? reinterpret_cast<const float*>(smem + L.glst + s * L.glst_stride)
                                               : a.gld + row0;

    // pass A: conditional elements -> input gradient + this element's parameter-gradient vector
    const int nA = nrow * Tc;
    for (int e = lane; e < nA; e += 32) {
      const int rr = fast_div(e, Tc, fx.magic_tc);
      const int j = e - rr * Tc;
      const int r = r_lo + rr;
      const int c = M.cond_col[j];
      float gv;
      fast::element_bwd<K, INV>(cur + static_cast<size_t>(r) * a.prow + j * Pc, xs[r * D + c], M.cbound[j], a.wh_scale,
                                M.cmode[j], a.outside_zero, gys[r * D + c], gl[r], &gv,
                                gst + static_cast<size_t>(r) * a.grow + j * Pc);
      gxs[r * D + c] = gv;
    }
    // this warp's rows of the parameter-gradient tile leave through the TMA engine
    if (Tc > 0 && nrow > 0) {
      const size_t bytes = static_cast<size_t>(nrow) * a.grow * sizeof(float);
      float* dst = a.gcond + (row0 + r_lo) * a.ld_gcond;
      const bool bulk = a.bulk_gp && bytes % 16 == 0 && (reinterpret_cast<uintptr_t>(dst) % 16 == 0) &&
                        (smem_u32(my_gst) % 16 == 0);
      if (bulk) {
        fence_async_smem();
        __syncwarp();
        if (lane == 0) {
          bulk_s2g(dst, my_gst, static_cast<uint32_t>(bytes));
          bulk_commit();
          store_pending = true;
        }
      } else {
        __syncwarp();
        const int n = Tc * Pc;
        for (int rr = 0; rr < nrow; ++rr)
          for (int q = lane; q < n; q += 32) a.gcond[(row0 + r_lo + rr) * a.ld_gcond + q] = my_gst[rr * a.grow + q];
      }
    }
    // pass B: pass-through columns copy the incoming gradient; batch-shared columns accumulate into the
    // lane-private table of their (row group, feature) slot
    if (Ts == 0) {
      const int nB = nrow * No;
      for (int e = lane; e < nB; e += 32) {
        const int rr = fast_div(e, No, fx.magic_no);
        const int c = M.other_col[e - rr * No];
        gxs[(r_lo + rr) * D + c] = gys[(r_lo + rr) * D + c];
      }
    } else {
      for (int e = lane; e < nrow * n_pass; e += 32) {     // pass-through columns lead the "other" list
        const int rr = e / n_pass;
        const int c = M.other_col[e - rr * n_pass];
        gxs[(r_lo + rr) * D + c] = gys[(r_lo + rr) * D + c];
      }
#pragma unroll 1
      for (int slot = lane; slot < slots; slot += 32) {
        const int g = Ts <= 32 ? fast_div(slot, Ts, fx.magic_ts) : 0;
        const int jo = slot - g * Ts;
        const int c = M.shared_col[jo];
        if (c < 0) continue;
        float* mine = acc + (static_cast<size_t>(warp) * slots + slot) * TK;
        for (int rr = g; rr < nrow; rr += Gw) {
          const int r = r_lo + rr;
          float gv;
          fast::table_element_bwd<K, INV>(rec + jo * (K + 1), xs[r * D + c], M.sbound[jo], a.outside_zero, gys[r * D + c],
                                          gl[r], &gv, mine);
          gxs[r * D + c] = gv;
        }
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[s]);

    // pass C: this warp's input-gradient rows
    if (nrow > 0) {
      const int n = nrow * D;
      const float* src = gxs + r_lo * D;
      if (small_c) {
        float* dst = gx_run + static_cast<int64_t>(it) * gx_step;
        if (lane < n) dst[lane] = src[lane];
        if (lane + 32 < n) dst[lane + 32] = src[lane + 32];
      } else if (a.ldgx == D) {
        float* dst = a.gx + (row0 + r_lo) * D;
#pragma unroll 1
        for (int e = lane; e < n; e += 32) dst[e] = src[e];
      } else {
        for (int rr = 0; rr < nrow; ++rr)
          for (int c = lane; c < D; c += 32) a.gx[(row0 + r_lo + rr) * a.ldgx + c] = src[rr * D + c];
      }
    }
    __syncwarp();
  }
  if (lane == 0 && store_pending) bulk_wait_read_all();
  if (Ts > 0) {
    __syncthreads();
    const int per_feature = kWarps * Gw;                     // tables per feature, summed in a fixed order
    for (int q = tid; q < Ts * TK; q += kThreads) {
      const int j = q / TK, i = q - j * TK;
      float sum = 0.f;
      for (int t = 0; t < per_feature; ++t) {
        const int w = t / Gw, g = t - w * Gw;
        sum += acc[(static_cast<size_t>(w) * slots + g * Ts + j) * TK + i];
      }
      tabg[q] = sum;
    }
    __syncthreads();
    for (int j = tid; j < Ts; j += kThreads)
      rqs_table_to_raw_grad<float, K>(a.shared + static_cast<size_t>(j) * a.Ps, K, M.smode[j], M.sbound[j],
                                      tabg + j * TK, a.partial + (static_cast<size_t>(blockIdx.x) * Ts + j) * a.Ps);
  }
}

// Second stage of the batch-shared parameter gradients: out[j] = sum over CTAs of partial[b, j].  One warp per
// output: lane l adds blocks l, l+32, ... in order, then a fixed xor tree -> bit-reproducible, and 32 loads in
// flight per output instead of one dependent chain (12.7 -> ~3 us for 296 x 345).
template <typename T>
__global__ void rqs_partial_reduce_kernel(const T* partial, int nblocks, int n, T* out) {
  const int j = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (j >= n) return;
  T acc = T(0);
  for (int b = lane; b < nblocks; b += 32) acc += partial[static_cast<size_t>(b) * n + j];
  acc = warp_sum(acc);
  if (lane == 0) out[j] = acc;
}

template <typename T>
__global__ void gather_cols_kernel(const T* src, int64_t lds, T* dst, int64_t ldd, int64_t B, int n,
                                   const int32_t* cols) {
  const int64_t total = B * n;
  if (total < (1ll << 31)) {                               // 32-bit index arithmetic (a 64-bit division per element
    const unsigned tot = static_cast<unsigned>(total), un = static_cast<unsigned>(n);   // was most of the kernel)
    for (unsigned e = blockIdx.x * blockDim.x + threadIdx.x; e < tot; e += gridDim.x * blockDim.x) {
      const unsigned b = e / un, j = e - b * un;
      dst[static_cast<int64_t>(b) * ldd + j] = __ldg(src + static_cast<int64_t>(b) * lds + __ldg(cols + j));
    }
    return;
  }
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
  if (total < (1ll << 31)) {
    const unsigned tot = static_cast<unsigned>(total), ud = static_cast<unsigned>(D_dst);
    for (unsigned e = blockIdx.x * blockDim.x + threadIdx.x; e < tot; e += gridDim.x * blockDim.x) {
      const unsigned b = e / ud, c = e - b * ud;
      const int j = inv[c];
      T* d = dst + static_cast<int64_t>(b) * ldd + c;
      if (j >= 0) {
        const T v = __ldg(src + static_cast<int64_t>(b) * lds + j);
        *d = accumulate ? *d + v : v;
      } else if (!accumulate) {
        *d = T(0);
      }
    }
    return;
  }
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
constexpr int kMaxBwdGrid = 148 * 2;

// Rows per tile: a multiple of 4 (keeps every tile start 16 B aligned for the bulk copies) whose
// parameter block fits the per-stage budget and whose rows*Tc fills whole passes of kThreads best.
static int pick_rows(int Tc, int prow, int D, size_t elt) {
  const size_t per_row = static_cast<size_t>(Tc > 0 ? prow : D) * elt;
  const size_t budget = 24 * 1024;
  int rmax = static_cast<int>(budget / per_row) / 4 * 4;
  if (rmax > 64) rmax = 64;
  if (rmax < 4) rmax = 4;
  const int width = Tc > 0 ? Tc : D;
  int best = 4;
  double best_u = -1.0;
  for (int R = 4; R <= rmax; R += 4) {
    const int n = R * width;
    const double u = static_cast<double>(n) / (static_cast<double>((n + kThreads - 1) / kThreads) * kThreads);
    if (u >= best_u - 1e-9) { best_u = u > best_u ? u : best_u; best = R; }
  }
  return best;
}

// Staged row pitch: the natural Tc*Pc, or the caller's leading dimension when that is a slightly padded
// one (e.g. 324 for 322 = 14 x 23 parameters: the pad makes the pitch a 16 B multiple, which is what
// lets the producing GEMM store and this kernel load whole tiles with the TMA engine).
template <typename T>
static void set_pitches(RqsLayerArgs<T>& a, bool bwd) {
  const int64_t n = static_cast<int64_t>(a.Tc) * a.Pc;
  a.prow = static_cast<int>((a.ld_cond > n && a.ld_cond - n < 16) ? a.ld_cond : n);
  a.grow = static_cast<int>((bwd && a.ld_gcond > n && a.ld_gcond - n < 16) ? a.ld_gcond : n);
}

template <typename T>
static void set_bulk_flags(RqsLayerArgs<T>& a, bool bwd) {
  auto al16 = [](const void* p) { return reinterpret_cast<uintptr_t>(p) % 16 == 0; };
  const size_t prow = static_cast<size_t>(a.R) * a.prow * sizeof(T);
  const size_t grow = static_cast<size_t>(a.R) * a.grow * sizeof(T);
  const size_t xrow = static_cast<size_t>(a.R) * a.D * sizeof(T);
  a.bulk_p = a.Tc > 0 && a.ld_cond == a.prow && prow % 16 == 0 && al16(a.cond);
  a.bulk_x = a.ldx == a.D && xrow % 16 == 0 && al16(a.x);
  a.bulk_gy = bwd && a.ldgy == a.D && xrow % 16 == 0 && al16(a.gy);
  a.bulk_gp = bwd && a.Tc > 0 && a.ld_gcond == a.grow && grow % 16 == 0 && al16(a.gcond);
}

template <typename T, int KMAX, bool EXACT>
static int launch_fwd(RqsLayerArgs<T>& a, cudaStream_t st) {
  const int sms = sm_count();
  if (sms <= 0) return static_cast<int>(cudaErrorNoDevice);
  set_pitches(a, false);
  a.R = pick_rows(a.Tc, a.prow > a.grow ? a.prow : a.grow, a.D, sizeof(T));
  a.num_tiles = (a.B + a.R - 1) / a.R;
  set_bulk_flags(a, false);
  const size_t smem = rqs_smem_layout<T>(a.R, a.D, a.Tc, a.Pc, a.Ts, a.K, false, kThreads, a.prow, a.grow).total;
  if (smem > kSmemMax - 2048) return VCNF_E_UNSUPPORTED;
  auto kern = rqs_layer_fwd_kernel<T, KMAX, EXACT>;
  static bool attr_set = false;     // once per instantiation (not a stream operation: keep it out of graph capture)
  if (!attr_set) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kSmemMax - 2048));   // static __shared__ of the kernel counts too
    attr_set = true;
  }
  int per_sm = static_cast<int>(kSmemMax / (smem + 1024));
  per_sm = per_sm < 1 ? 1 : (per_sm > 4 ? 4 : per_sm);
  const int64_t grid = a.num_tiles < static_cast<int64_t>(sms) * per_sm ? a.num_tiles : static_cast<int64_t>(sms) * per_sm;
  kern<<<static_cast<unsigned>(grid), kThreads, smem, st>>>(a);
  return launch_status();
}

template <typename T, int KMAX, bool EXACT>
static int launch_bwd(RqsLayerArgs<T>& a, T* g_shared, cudaStream_t st) {
  const int sms = sm_count();
  if (sms <= 0) return static_cast<int>(cudaErrorNoDevice);
  if (a.Ts > kThreads) return VCNF_E_UNSUPPORTED;
  set_pitches(a, true);
  a.R = pick_rows(a.Tc, a.prow > a.grow ? a.prow : a.grow, a.D, sizeof(T));
  a.num_tiles = (a.B + a.R - 1) / a.R;
  set_bulk_flags(a, true);
  const size_t smem = rqs_smem_layout<T>(a.R, a.D, a.Tc, a.Pc, a.Ts, a.K, true, kThreads, a.prow, a.grow).total;
  if (smem > kSmemMax - 2048) return VCNF_E_UNSUPPORTED;
  auto kern = rqs_layer_bwd_kernel<T, KMAX, EXACT>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kSmemMax - 2048));   // static __shared__ of the kernel counts too
    attr_set = true;
  }
  int per_sm = static_cast<int>(kSmemMax / (smem + 1024));
  per_sm = per_sm < 1 ? 1 : (per_sm > 2 ? 2 : per_sm);
  int64_t grid = static_cast<int64_t>(sms) * per_sm;
  if (grid > kMaxBwdGrid) grid = kMaxBwdGrid;
  if (grid > a.num_tiles) grid = a.num_tiles;
  kern<<<static_cast<unsigned>(grid), kThreads, smem, st>>>(a);
  int rc = launch_status();
  if (rc == 0 && a.Ts > 0) {
    const int n = a.Ts * a.Ps;
    rqs_partial_reduce_kernel<T><<<(n + 3) / 4, 128, 0, st>>>(a.partial, static_cast<int>(grid), n, g_shared);
    rc = launch_status();
  }
  return rc;
}

// Rows per tile of the fast kernels: 8 warps x rpw rows; the parameter block of a stage stays within the
// budget and rpw*Tc fills whole warp passes best.
static int pick_rows_fast(int Tc, int prow, int D, size_t budget) {
  const size_t per_row = static_cast<size_t>(Tc > 0 ? prow : D) * sizeof(float);
  const int width = Tc > 0 ? Tc : D;
  int best = 1;
  double best_u = -1.0;
  for (int rpw = 1; rpw <= 8; ++rpw) {
    if (rpw > 1 && static_cast<size_t>(rpw) * kWarps * per_row > budget) break;
    if ((rpw * kWarps) % 4 != 0) continue;
    const int n = rpw * width;
    const double u = static_cast<double>(n) / (static_cast<double>((n + 31) / 32) * 32);
    if (u > best_u + 1e-9) { best_u = u; best = rpw; }
  }
  return best * kWarps;
}

static uint32_t div_magic(int n, int64_t max_e) {
  if (n <= 0) return 0;
  return (max_e * n < (1 << 20)) ? static_cast<uint32_t>((1u << 20) / static_cast<uint32_t>(n) + 1u) : 0u;
}

constexpr int kFastStages = 2;
static int fast_stages() {            // VCNF_RQS_STAGES=3: three-stage tile ring (experiments)
  static int v = -1;
  if (v < 0) { const char* e = getenv("VCNF_RQS_STAGES"); v = (e && e[0] == '3') ? 3 : kFastStages; }
  return v;
}

// VCNF_RQS_GENERIC=1 routes fp32 / K = 8 layers through the generic kernels (A/B measurements, tests)
static bool generic_only() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("VCNF_RQS_GENERIC"); v = (e && e[0] == '1') ? 1 : 0; }
  return v == 1;
}

template <bool INV>
static int launch_fwd_fast(RqsLayerArgs<float>& a, cudaStream_t st) {
  const int sms = sm_count();
  if (sms <= 0) return static_cast<int>(cudaErrorNoDevice);
  set_pitches(a, false);
  a.R = pick_rows_fast(a.Tc, a.prow, a.D, 24 * 1024);
  a.num_tiles = (a.B + a.R - 1) / a.R;
  set_bulk_flags(a, false);
  const int nst = fast_stages();
  const size_t smem = fast_smem_layout(a.R, a.D, a.Tc, a.Ts, false, a.prow, a.grow, nst).total;
  if (smem > kSmemMax - 2048) return VCNF_E_UNSUPPORTED;
  auto kern = nst == 3 ? rqs_layer_fwd_fast_kernel<INV, 3> : rqs_layer_fwd_fast_kernel<INV, 2>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kSmemMax - 2048));
    attr_set = true;
  }
  int per_sm = static_cast<int>(kSmemMax / (smem + 1024));
  per_sm = per_sm < 1 ? 1 : (per_sm > 4 ? 4 : per_sm);
  const int64_t grid = a.num_tiles < static_cast<int64_t>(sms) * per_sm ? a.num_tiles : static_cast<int64_t>(sms) * per_sm;
  FastExtra fx;
  fx.magic_tc = div_magic(a.Tc, static_cast<int64_t>(a.R) * a.Tc);
  const int n_other = a.D - a.Tc;          // every column is conditional, batch-shared or pass-through
  fx.magic_no = div_magic(n_other, static_cast<int64_t>(a.R) * n_other);
  fx.magic_ts = 0;
  kern<<<static_cast<unsigned>(grid), kThreads, smem, st>>>(a, fx);
  return launch_status();
}

template <bool INV>
static int launch_bwd_fast(RqsLayerArgs<float>& a, float* g_shared, cudaStream_t st) {
  const int sms = sm_count();
  if (sms <= 0) return static_cast<int>(cudaErrorNoDevice);
  set_pitches(a, true);
  a.R = pick_rows_fast(a.Tc, a.prow > a.grow ? a.prow : a.grow, a.D, 24 * 1024);
  a.num_tiles = (a.B + a.R - 1) / a.R;
  set_bulk_flags(a, true);
  a.bulk_gp = a.Tc > 0 && a.ld_gcond == a.grow && reinterpret_cast<uintptr_t>(a.gcond) % 16 == 0;   // per-warp spans: checked in the kernel
  const int nst = fast_stages();
  const size_t smem = fast_smem_layout(a.R, a.D, a.Tc, a.Ts, true, a.prow, a.grow, nst).total;
  if (smem > kSmemMax - 2048) return VCNF_E_UNSUPPORTED;
  auto kern = nst == 3 ? rqs_layer_bwd_fast_kernel<INV, 3> : rqs_layer_bwd_fast_kernel<INV, 2>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kSmemMax - 2048));
    attr_set = true;
  }
  int per_sm = static_cast<int>(kSmemMax / (smem + 1024));
  per_sm = per_sm < 1 ? 1 : (per_sm > 2 ? 2 : per_sm);
  int64_t grid = static_cast<int64_t>(sms) * per_sm;
  if (grid > kMaxBwdGrid) grid = kMaxBwdGrid;
  if (grid > a.num_tiles) grid = a.num_tiles;
  FastExtra fx;
  const int n_other = a.D - a.Tc;
  fx.magic_tc = div_magic(a.Tc, static_cast<int64_t>(a.R) * a.Tc);
  fx.magic_no = div_magic(n_other, static_cast<int64_t>(a.R) * n_other);
  fx.magic_ts = div_magic(a.Ts, 64);
  kern<<<static_cast<unsigned>(grid), kThreads, smem, st>>>(a, fx);
  int rc = launch_status();
  if (rc == 0 && a.Ts > 0) {
    const int n = a.Ts * a.Ps;
    rqs_partial_reduce_kernel<float><<<(n + 3) / 4, 128, 0, st>>>(a.partial, static_cast<int>(grid), n, g_shared);
    rc = launch_status();
  }
  return rc;
}

// K == 8 (the default of every spline wrapper, neural_spline/wrapper.py:20-31) gets a fully unrolled
// instantiation; other bin counts run the bounded-loop instantiation of the same code.
template <typename T>
static int dispatch_fwd(RqsLayerArgs<T>& a, cudaStream_t st) {
  if constexpr (sizeof(T) == 4) {
    if (a.K == 8 && !generic_only()) {
      const int rc = a.inverse ? launch_fwd_fast<true>(a, st) : launch_fwd_fast<false>(a, st);
      if (rc != VCNF_E_UNSUPPORTED) return rc;
    }
  }
  if (a.K == 8) return launch_fwd<T, 8, true>(a, st);
  if (a.K < 8) return launch_fwd<T, 8, false>(a, st);
  if (a.K <= 16) return launch_fwd<T, 16, false>(a, st);
  if (a.K <= 32) return launch_fwd<T, 32, false>(a, st);
  return VCNF_E_UNSUPPORTED;
}
template <typename T>
static int dispatch_bwd(RqsLayerArgs<T>& a, T* g_shared, cudaStream_t st) {
  if constexpr (sizeof(T) == 4) {
    if (a.K == 8 && !generic_only()) {
      const int rc = a.inverse ? launch_bwd_fast<true>(a, g_shared, st) : launch_bwd_fast<false>(a, g_shared, st);
      if (rc != VCNF_E_UNSUPPORTED) return rc;
    }
  }
  if (a.K == 8) return launch_bwd<T, 8, true>(a, g_shared, st);
  if (a.K < 8) return launch_bwd<T, 8, false>(a, g_shared, st);
  if (a.K <= 16) return launch_bwd<T, 16, false>(a, g_shared, st);
  if (a.K <= 32) return launch_bwd<T, 32, false>(a, g_shared, st);
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
  if (B == 0) return 0;   /* empty batch: pointers may be NULL */
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
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (B == 0) {   /* empty batch: batch pointers may be NULL; shared-parameter gradients are zero */
    if (Ts > 0 && g_shared) cudaMemsetAsync(g_shared, 0, static_cast<size_t>(Ts) * Ps * (dtype == VCNF_F64 ? 8 : 4), st);
    return launch_status();
  }
  if (bad_layer_args(x, B, D, K, col_map, cond_params, Tc, Pc, cond_modes, cond_bounds, shared_params, Ts, Ps,
                     shared_modes, shared_bounds, dtype) || !gy || !glogdet || !gx || (Tc > 0 && !g_cond) ||
      (Ts > 0 && (!g_shared || !workspace)))
    return VCNF_E_BADARG;
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
  if (B == 0) return 0;   /* empty batch: pointers may be NULL */
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
  if (B == 0) return 0;   /* empty batch: pointers may be NULL */
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
