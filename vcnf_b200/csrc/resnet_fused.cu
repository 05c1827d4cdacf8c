// The whole conditioner of a spline coupling layer and the spline itself as ONE kernel:
//
//   h0 = W0 x_id + b0;  h_{i+1} = h_i + Wb_i relu(Wa_i relu(h_i) + ba_i) + bb_i   (normflow/nets/resnet.py:38-106)
//   theta = Wf h_nb + bf                                                           ([B, Tc * Pc], resnet.py:104)
//   y[:, tr] = RQS^{+-1}(x[:, tr]; theta),  y[:, id] = CDF^{+-1}(x[:, id]) or x[:, id],  logdet = sum log|dy/dx|
//                                           (flows/neural_spline/coupling.py:70-125,309-343; utils/splines.py:88-193)
//
// The layer-by-layer path moves the [B, 128] hidden activations through HBM ten times and the [B, 322] spline
// parameters twice (1 524 B per sample-layer); here a 128-row tile stays on the SM from the gathered
// inputs to the transformed outputs (2 * 4 D + 4 B per sample-layer in the no-grad directions; with gradients the
// same launch additionally writes what the backward kernels read: layer outputs and theta):
//
//   * the activation operand of the next layer is written by the epilogue warps straight into TENSOR MEMORY (TF32
//     hi / lo planes, columns [0, 256)) and read from there by tcgen05.mma (3xTF32, see gemm_tf32x3.cu); the
//     accumulators (hi*hi and the cross terms) take the other 256 columns, and are read out whole and released before
//     the epilogue starts converting, so the next layer's MMAs start as soon as its first 32 input columns exist;
//   * the weights (pre-split and pre-tiled by vcnf_resnet_spline_pack: 16-column stages of 16 KB) stream from L2
//     through a 7-stage ring of bulk copies issued by one thread;
//   * the residual stream of the tile lives in shared memory ([column][row], conflict-free both ways); during the head
//     the same buffer stages the parameters of a slice (whole features, each padded to a multiple of 8 columns), the
//     accumulator is released, and the splines are evaluated from there while the tensor core works on the next slice;
//   * x tiles are prefetched one tile ahead with cp.async by two otherwise idle warps, y tiles leave as 16-byte stores.
//
// Warp roles (640 threads, one CTA per SM, persistent over tiles): warp 0 weight copies, warp 1 MMA issue (operands in
// convergent code, one asm block per stage), warps 2-3 x prefetch (warp 2 also allocates TMEM), warps 4-19 epilogue
// (TMEM lane quarter = warp % 4, four warps per quarter split the columns / features).
//
// Measured (B200, cfg 2 layer, 65 536 rows): 122 us against 153 us of GEMM launches + 27 us spline layer by layer;
// one tile takes 28 us of which 16.5 us are MMAs (a 128 x 128 x 8 TF32 MMA with the A operand in tensor memory retires
// every 91 clocks here, 64-column MMAs every 61: splitting a layer into column halves to overlap its epilogue was
// measured and does not pay), the rest are the hand-offs between the MMA stream and the epilogue of the SAME tile.
#include <cstdlib>
#include <cuda.h>
#include "common.cuh"
#include "rqs_fast.cuh"
#include "../../include/vcnf_b200.h"

namespace vcnf {
namespace fusednet {

constexpr int kH = 128;                 // hidden width = UMMA N of the hidden layers (and rows per tile = UMMA M)
constexpr int kRows = 128;
constexpr int kRS = kRows + 1;           // row stride of the [column][row] staging buffer: conflict-free both ways
constexpr int kMaxBlocks = 4;
constexpr int kMaxUnits = 1 + 2 * kMaxBlocks + 4;      // layers + head slices
constexpr int kMaxSlices = 4;
constexpr int kWK = 16;                 // columns of the contraction per weight stage
constexpr int kWStages = 7;
constexpr uint32_t kWPlane = kH * kWK * 4;             // 8 KB: hi (or lo) plane of a weight stage
constexpr uint32_t kWStage = 2 * kWPlane;
constexpr uint32_t kWSbo = kWK * 32;                   // byte stride between 8-row groups of a weight stage
constexpr int kAChunk = 32;             // activation columns per chunk (readiness granularity)
constexpr uint32_t kAPlane = kRows * kAChunk * 4;      // 16 KB
constexpr uint32_t kAChunkBytes = 2 * kAPlane;         // hi | lo
constexpr uint32_t kASbo = kAChunk * 32;
constexpr int kEpiWarp0 = 4, kEpiWarps = 16, kEpiThreads = kEpiWarps * 32;
constexpr int kThreads = (kEpiWarp0 + kEpiWarps) * 32;
constexpr int kG = kEpiWarps / 4;       // column groups: epilogue threads per row
constexpr int kCW = 32 / kG;            // columns of a 32-column chunk per thread
constexpr int kMaxD = 32;
constexpr uint32_t kTmemCols = 512;     // 2 buffers x (main 128 | cross terms 128)

struct Args {
  const float* x; int64_t ldx;
  float* y; int64_t ldy;
  float* logdet; int accumulate;
  int64_t B; int D;
  const unsigned char* packed;          // weight stages, then biases (see pack kernel)
  int n_in, nb, Tc, Pc, PcP, fs, n_slices;
  int id_cols[kMaxD], tr_cols[kMaxD];
  unsigned char modes[kMaxD];
  float bounds[kMaxD];
  float wh_scale;
  int outside_zero;
  int* err_flag;
  const float* shared; int Ts, Ps;      // optional: batch-shared spline on the id columns (unconditional transform)
  unsigned char smodes[kMaxD];
  float sbounds[kMaxD];
  float* params; int64_t ldp;           // optional: theta [B, Tc * Pc]
  float* acts; int64_t act_stride;      // optional: pre-activation outputs of every layer, [1 + 2 nb][B, 128]
  float* xid; int64_t ldxid;            // optional: the gathered net inputs x[:, id_cols], [B, n_in]
};

// ---- raw PTX wrappers (same conventions as gemm_tf32x3.cu) ------------------------------------------------------
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(cols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void mma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                         uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
      "}\n" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void mma_tf32_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t"
      "}\n" ::"r"(d_tmem),
      "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ bool elect_one() {
  uint32_t leader;
  asm volatile("{.reg .b32 rx; .reg .pred px; elect.sync rx|px, 0xffffffff; selp.b32 %0, 1, 0, px;}" : "=r"(leader));
  return leader != 0;
}
// One weight stage (16 columns of the contraction = two K = 8 steps) of the split product, activations from tensor
// memory: A_lo*W_hi and A_hi*W_lo into d_small, A_hi*W_hi into d_main.  acc = 0 starts both accumulators.
#define VCNF_MMA_TS "tcgen05.mma.cta_group::1.kind::tf32 "
__device__ __forceinline__ void mma_stage2(uint32_t d_main, uint32_t d_small, uint32_t a_hi, uint32_t a_lo, uint64_t b_hi,
                                           uint64_t b_lo, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t"
      ".reg .pred pa, pt;\n\t"
      ".reg .b32 ah1, al1;\n\t"
      ".reg .b64 bh1, bl1;\n\t"
      "setp.ne.b32 pa, %7, 0;\n\t"
      "setp.eq.b32 pt, %6, %6;\n\t"
      "add.u32 ah1, %2, 8;\n\t add.u32 al1, %3, 8;\n\t add.u64 bh1, %4, 16;\n\t add.u64 bl1, %5, 16;\n\t"
      VCNF_MMA_TS "[%1], [%3], %4, %6, pa;\n\t"
      VCNF_MMA_TS "[%1], [%2], %5, %6, pt;\n\t"
      VCNF_MMA_TS "[%0], [%2], %4, %6, pa;\n\t"
      VCNF_MMA_TS "[%1], [al1], bh1, %6, pt;\n\t"
      VCNF_MMA_TS "[%1], [ah1], bl1, %6, pt;\n\t"
      VCNF_MMA_TS "[%0], [ah1], bh1, %6, pt;\n\t"
      "}\n" ::"r"(d_main), "r"(d_small), "r"(a_hi), "r"(a_lo), "l"(b_hi), "l"(b_lo), "r"(idesc), "r"(acc)
      : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t* r) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr)
               : "memory");
}
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t* r) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr), "r"(r[0]),
               "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
               : "memory");
}
template <int N> __device__ __forceinline__ void tmem_ld(uint32_t taddr, uint32_t* r) {
  if (N == 16) tmem_ld16(taddr, r); else tmem_ld8(taddr, r);
}
template <int N> __device__ __forceinline__ void tmem_st(uint32_t taddr, const uint32_t* r) {
  if (N == 16) tmem_st16(taddr, r); else tmem_st8(taddr, r);
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads) : "memory"); }

// K-major, no swizzle: 8 rows x 16 B core matrices, LBO = distance between the two 16 B k-chunks of one MMA (K = 8
// tf32), SBO = distance between 8-row groups (cute/arch/mma_sm100_desc.hpp bit layout, version 1).
__device__ __forceinline__ uint64_t smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((saddr >> 4) & 0x3fffu);
  d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3fffu) << 16;
  d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3fffu) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  return d;
}
__host__ __device__ constexpr uint32_t instr_desc_tf32(int M, int N) {
  return (1u << 4) | (2u << 7) | (2u << 10) | (static_cast<uint32_t>(N >> 3) << 17) |
         (static_cast<uint32_t>(M >> 4) << 24);
}
__host__ __device__ __forceinline__ float rn_tf32(float a) {
#if defined(__CUDA_ARCH__)
  return __uint_as_float((__float_as_uint(a) + 0x1000u) & 0xffffe000u);
#else
  return a;
#endif
}
__device__ __forceinline__ void split_tf32(float a, float& hi, float& lo) {
  hi = rn_tf32(a);
  lo = rn_tf32(a - hi);
}

// ---- packed parameter buffer -------------------------------------------------------------------------------------
// Units u = 0 .. L + S - 1: the L = 1 + 2 nb linear layers, then the S head slices.  Unit 0 has ceil(n_in / 16)
// weight stages, every other unit 8 (K = 128).  A stage is [hi plane | lo plane], each 128 rows x 16 k in the
// canonical layout (row = output feature of the unit).  Head slice s, row fi * PcP + i = parameter i of feature
// s * fs + fi (zero rows pad).  After the stages: (L + S) x 128 biases in the same row order.
__host__ __device__ inline int n_layers(int nb) { return 1 + 2 * nb; }
__host__ __device__ inline int stages0(int n_in) { return (n_in + kWK - 1) / kWK; }
__host__ __device__ inline int total_stages(int n_in, int nb, int S) { return stages0(n_in) + 8 * (2 * nb + S); }
__host__ __device__ inline size_t packed_bytes(int n_in, int nb, int S) {
  return static_cast<size_t>(total_stages(n_in, nb, S)) * kWStage + static_cast<size_t>(n_layers(nb) + S) * kH * 4;
}

struct PackArgs {
  const float* W[kMaxUnits];     // [0]: initial [128, n_in]; [1 .. 2 nb]: hidden [128, 128]; [2 nb + 1]: head [Tc*Pc, 128]
  const float* b[kMaxUnits];
  int n_in, nb, Tc, Pc, PcP, fs, S;
  unsigned char* out;
};

__global__ void __launch_bounds__(256) pack_kernel(const PackArgs a) {
  const int L = n_layers(a.nb), S = a.S;
  const int st0 = stages0(a.n_in);
  const int n_st = total_stages(a.n_in, a.nb, S);
  const int64_t n_w = static_cast<int64_t>(n_st) * kH * kWK;
  const int64_t n_b = static_cast<int64_t>(L + S) * kH;
  for (int64_t e = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; e < n_w + n_b;
       e += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    if (e < n_w) {
      const int st = static_cast<int>(e / (kH * kWK));
      const int rem = static_cast<int>(e - static_cast<int64_t>(st) * (kH * kWK));
      const int n = rem / kWK, kk = rem - n * kWK;
      int u, t;
      if (st < st0) { u = 0; t = st; } else { u = 1 + (st - st0) / 8; t = (st - st0) % 8; }
      const int k = t * kWK + kk;
      float v = 0.f;
      if (u == 0) {
        if (k < a.n_in) v = a.W[0][n * a.n_in + k];
      } else if (u < L) {
        v = a.W[u][n * kH + k];
      } else {
        const int s = u - L, fi = n / a.PcP, i = n - fi * a.PcP, f = s * a.fs + fi;
        if (fi < a.fs && i < a.Pc && f < a.Tc) v = a.W[L][static_cast<int64_t>(f * a.Pc + i) * kH + k];
      }
      float hi, lo;
      split_tf32(v, hi, lo);
      const uint32_t off = static_cast<uint32_t>(n >> 3) * kWSbo + (kk >> 2) * 128 + (n & 7) * 16 + (kk & 3) * 4;
      unsigned char* base = a.out + static_cast<size_t>(st) * kWStage;
      *reinterpret_cast<float*>(base + off) = hi;
      *reinterpret_cast<float*>(base + kWPlane + off) = lo;
    } else {
      const int eb = static_cast<int>(e - n_w);
      const int u = eb / kH, n = eb - u * kH;
      float v = 0.f;
      if (u < L) {
        v = a.b[u] ? a.b[u][n] : 0.f;
      } else {
        const int s = u - L, fi = n / a.PcP, i = n - fi * a.PcP, f = s * a.fs + fi;
        if (fi < a.fs && i < a.Pc && f < a.Tc && a.b[L]) v = a.b[L][f * a.Pc + i];
      }
      reinterpret_cast<float*>(a.out + static_cast<size_t>(n_st) * kWStage)[eb] = v;
    }
  }
}

// One spline element whose parameter vector sits in REGISTERS: fast::locate with the two derivative parameters picked
// by selects instead of a dynamically indexed load (same values, same arithmetic).
template <int K, bool INV, int PCP>
__device__ __forceinline__ void element_regs(const float (&prm)[PCP], float v, float bound, float scale, int mode,
                                             int outside_zero, float* out, float* lad, int* flag) {
  using namespace fast;
  P2 p[K], kn[K + 1];
  const float left = rqs_lo(bound, mode);
  axes<K>(prm, scale, bound, left, p, kn);
  P2 lo = kn[0], hi = kn[1];
  int idx = 0;
#pragma unroll
  for (int k = 1; k < K; ++k) {
    const bool ge = v >= (INV ? kn[k].y : kn[k].x);
    lo.x = ge ? kn[k].x : lo.x;  lo.y = ge ? kn[k].y : lo.y;
    hi.x = ge ? kn[k + 1].x : hi.x;  hi.y = ge ? kn[k + 1].y : hi.y;
    idx = ge ? k : idx;
  }
  RqsBin<float> b;
  b.idx = idx;
  const P2 wh = p2add(hi, P2{-lo.x, -lo.y});
  b.x0 = lo.x; b.w = wh.x; b.y0 = lo.y; b.h = wh.y;
  deriv_slots(idx, K, mode, &b.s0, &b.s1);
  const float edge = RqsConst<float>::edge();
  float r0 = edge, r1 = edge;
#pragma unroll
  for (int k = 0; k < PCP - 2 * K; ++k) {
    r0 = (b.s0 == k) ? prm[2 * K + k] : r0;
    r1 = (b.s1 == k) ? prm[2 * K + k] : r1;
  }
  b.r0 = r0; b.r1 = r1;
  P2 e01;
  const P2 sp = softplus2(P2{r0, r1}, &e01);
  b.d0 = RqsConst<float>::min_deriv() + sp.x;
  b.d1 = RqsConst<float>::min_deriv() + sp.y;
  float o, l;
  int ok = 1;
  if (INV) rq_inverse(v, b, &o, &l, &ok); else rq_forward(v, b, &o, &l);
  const bool inside = (v >= left) && (v <= left + (bound - (-bound)));
  *out = inside ? o : (outside_zero ? 0.f : v);
  *lad = inside ? l : 0.f;
  if (inside && !ok) *flag |= 1;
}

// Timeline of CTA 0 (clock64 at the hand-offs of the MMA warp and of the first epilogue warp), read back with
// vcnf_debug_fused_trace by tools/fused_trace.py; compiled in with -DVCNF_FUSED_TRACE only.
#ifdef VCNF_FUSED_TRACE
__device__ long long g_ftrace[2][128];
#define FTRACE(role, idx) do { if (blockIdx.x == 0 && lane == 0 && (idx) < 128) g_ftrace[role][idx] = clock64(); } while (0)
#else
#define FTRACE(role, idx) do { } while (0)
#endif

template <bool INV, int PCP>
__global__ void __launch_bounds__(kThreads, 1) resnet_spline_kernel(const __grid_constant__ Args a) {
  extern __shared__ __align__(1024) unsigned char smem[];
  uint64_t* w_full = reinterpret_cast<uint64_t*>(smem);           // [kWStages] copy engine -> MMA
  uint64_t* w_empty = w_full + kWStages;                          // [kWStages] MMA -> copy issuer
  uint64_t* a_ready = reinterpret_cast<uint64_t*>(smem + 128);    // [4] epilogue -> MMA: activation chunk j written
  uint64_t* acc_full = a_ready + 4;                               // MMA -> epilogue
  uint64_t* acc_empty = acc_full + 1;                             // epilogue -> MMA: accumulator read out
  uint64_t* x_full = acc_empty + 1;                               // [2] prefetch warps -> epilogue
  uint64_t* x_empty = x_full + 2;                                 // [2] epilogue -> prefetch warps
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(x_empty + 2);
  int* s_id = reinterpret_cast<int*>(smem + 256);                 // [32]
  int* s_tr = s_id + kMaxD;                                       // [32]
  int* s_mode = s_tr + kMaxD;                                     // [32]
  float* s_bound = reinterpret_cast<float*>(s_mode + kMaxD);      // [32]
  float* s_sbound = reinterpret_cast<float*>(smem + 768);         // [32]
  unsigned char* Wst = smem + 1024;                               // kWStages x [hi 8 KB | lo 8 KB]
  float* res = reinterpret_cast<float*>(Wst + kWStages * kWStage);    // [128 columns][128 rows]: residual stream, then
                                                                      // the staged spline parameters of a head slice
  float* bias = res + kH * kRS + 32;                              // (+32 floats keeps 16-byte alignment)                                 // [(L + S)][128]
  const int L = n_layers(a.nb), S = a.n_slices, D = a.D;
  float* lds = bias + (L + S) * kH;                               // [128]
  fast::Knot4* rec = reinterpret_cast<fast::Knot4*>(lds + (kG - 1) * kRows); // [Ts][9] knot records of the batch-shared splines
  float* xs0 = reinterpret_cast<float*>(rec + a.Ts * 9);          // 2 x [128][D]: x tile in, y tile out

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int st0 = stages0(a.n_in);
  const int n_st = total_stages(a.n_in, a.nb, S);
  const int64_t n_tiles = (a.B + kRows - 1) / kRows;

  if (tid == 0) {
    for (int s = 0; s < kWStages; ++s) { mbar_init(&w_full[s], 1); mbar_init(&w_empty[s], 1); }
    for (int j = 0; j < 4; ++j) mbar_init(&a_ready[j], kEpiWarps);
    mbar_init(acc_full, 1);
    mbar_init(acc_empty, kEpiWarps);
    for (int b = 0; b < 2; ++b) { mbar_init(&x_full[b], 2); mbar_init(&x_empty[b], kEpiWarps); }
    mbar_init_fence();
  }
  if (warp == 2) tmem_alloc(tmem_slot, kTmemCols);
  if (tid < kMaxD) {
    s_id[tid] = a.id_cols[tid];
    s_tr[tid] = a.tr_cols[tid];
    s_mode[tid] = a.modes[tid];
    s_bound[tid] = a.bounds[tid];
    s_sbound[tid] = a.sbounds[tid];
    if (tid < a.Ts) fast::build_records<8>(a.shared + tid * a.Ps, a.smodes[tid], a.sbounds[tid], rec + tid * 9);
  }
  {
    const float* bsrc = reinterpret_cast<const float*>(a.packed + static_cast<size_t>(n_st) * kWStage);
    for (int i = tid; i < (L + S) * kH; i += kThreads) bias[i] = __ldg(bsrc + i);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t col_alo = 128, col_main = 256, col_cross = 384;   // A hi at column 0

  if (warp == 0) {
    // ===================================== weight stages: bulk copies from L2 =====================================
    if (lane == 0) {
      // Every CTA walks the same 0.9 MB of weights: they stay in L2 (replicating the block per CTA group changed
      // nothing measurable, so one copy it is).
      const unsigned char* wsrc = a.packed;
      uint32_t s = 0, ph = 0;
      for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        for (int st = 0; st < n_st; ++st) {
          mbar_wait(&w_empty[s], ph ^ 1u);
          mbar_expect_tx(&w_full[s], kWStage);
          bulk_g2s(Wst + s * kWStage, wsrc + static_cast<size_t>(st) * kWStage, kWStage, &w_full[s]);
          if (++s == kWStages) { s = 0; ph ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================================== MMA issue (activations from tensor memory) =====================================
    const uint32_t idesc = instr_desc_tf32(kRows, kH);
    uint32_t s = 0, ph = 0, uc = 0, aph = 0;
    int ti = 0;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
      for (int u = 0; u < L + S; ++u, ++uc) {
        FTRACE(0, ti); ++ti;
        mbar_wait(acc_empty, (uc & 1u) ^ 1u);
        tc_fence_after();
        FTRACE(0, ti); ++ti;
        const int nst = u == 0 ? st0 : 8;
        for (int t = 0; t < nst; ++t) {
          const int j = t >> 1;
          if (u <= L && (t & 1) == 0) {               // chunk j of this unit's activation operand (head slices share one)
            mbar_wait(&a_ready[j], (aph >> j) & 1u);
            aph ^= 1u << j;
          }
          mbar_wait(&w_full[s], ph);
          tc_fence_after();
          // Operands are computed by the WHOLE warp in convergent code and the six MMAs of a stage leave from one asm
          // block: inside a divergent `if (lane == 0)` every tcgen05.mma is wrapped in an ELECT / broadcast loop
          // (measured here: 117 clk per MMA issued that way against the 64 clk the shape needs).
          const uint32_t a_hi = tmem_base + static_cast<uint32_t>(t) * kWK;
          const uint32_t b_hi = smem_u32(Wst) + s * kWStage;
          const uint64_t bh = smem_desc(b_hi, 128, kWSbo);
          const uint64_t bl = smem_desc(b_hi + kWPlane, 128, kWSbo);
          const uint32_t acc0 = t > 0 ? 1u : 0u;
          const bool last = t == nst - 1;
          if (elect_one()) {
            mma_stage2(tmem_base + col_main, tmem_base + col_cross, a_hi, a_hi + col_alo, bh, bl, idesc, acc0);
            tc_commit(&w_empty[s]);
            if (last) tc_commit(acc_full);
          }
          __syncwarp();
          if (t == 0 || t == nst - 1) { FTRACE(0, ti); ++ti; }
          if (++s == kWStages) { s = 0; ph ^= 1u; }
        }
      }
    }
  } else if (warp < kEpiWarp0) {
    // ===================================== x tiles: prefetched one tile ahead =====================================
    const int pt = tid - 64;   // warps 2 and 3
    const bool fast_copy = a.ldx == D && reinterpret_cast<uintptr_t>(a.x) % 16 == 0;
    uint32_t k = 0;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++k) {
      const uint32_t b = k & 1u;
      float* xs = xs0 + b * (kRows * D);
      const int64_t m0 = tile * kRows;
      mbar_wait(&x_empty[b], ((k >> 1) & 1u) ^ 1u);
      if (fast_copy) {
        const int64_t valid = (a.B - m0 < kRows ? a.B - m0 : kRows) * D * 4;       // bytes of the tile that exist
        const unsigned char* src = reinterpret_cast<const unsigned char*>(a.x + m0 * D);
        for (int c = pt; c < kRows * D / 4; c += 64) {
          const int64_t left = valid - static_cast<int64_t>(c) * 16;
          const uint32_t nbytes = left >= 16 ? 16u : (left > 0 ? static_cast<uint32_t>(left) : 0u);
          cp_async16(smem_u32(xs) + c * 16, nbytes ? src + c * 16 : src, nbytes);
        }
        asm volatile("cp.async.wait_all;" ::: "memory");
      } else {
        for (int e0 = pt; e0 < kRows * D; e0 += 64 * 8) {
          float v[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int e = e0 + i * 64, r = e / D, c = e - r * D;
            v[i] = (e < kRows * D && m0 + r < a.B) ? __ldg(a.x + (m0 + r) * a.ldx + c) : 0.f;
          }
#pragma unroll
          for (int i = 0; i < 8; ++i) if (e0 + i * 64 < kRows * D) xs[e0 + i * 64] = v[i];
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&x_full[b]);
    }
  } else {
    // ===================================== epilogue / spline =====================================
    // thread = (row, g): TMEM lane quarter q = warp % 4, column group g = kCW columns of every 32-column chunk
    const int ew = warp - kEpiWarp0, q = warp & 3, g = ew >> 2;
    const int et = ew * 32 + lane, row = q * 32 + lane;
    const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    uint32_t uc = 0, k = 0;
    int flag = 0;
    int ei = warp == kEpiWarp0 ? 0 : 1000;
    // Prologue of a tile: wait for its prefetched x, (sampling direction) unconditional splines of the id columns,
    // first-layer operand -> tensor memory, (density direction) unconditional splines.  Returns the log-det collected
    // so far.  It runs for tile t + 1 in the middle of tile t's LAST head slice (right after that slice's accumulator is
    // staged: every MMA of tile t has retired, the operand columns are free), so the first layer of the next tile
    // computes while the last splines of this one are evaluated and its outputs stored.
    auto prologue = [&](int64_t tile_p, uint32_t kp) -> float {
      const int64_t m0 = tile_p * kRows;
      float* xs = xs0 + (kp & 1u) * (kRows * D);
      mbar_wait(&x_full[kp & 1u], (kp >> 1) & 1u);
      float ldsum = 0.f;
      if (INV && a.Ts > 0) {            // sampling direction: the id columns go through their own spline FIRST
        for (int j = g; j < a.Ts; j += kG) {
          float o, l;
          int bin;
          fast::table_element<8, true>(rec + j * 9, xs[row * D + s_id[j]], s_sbound[j], a.outside_zero, &o, &l, &bin, &flag);
          xs[row * D + s_id[j]] = o;
          ldsum += l;
        }
        epi_bar();
      }
      // gathered inputs of the net -> columns [kCW g, kCW g + kCW) of the activation operand (tensor memory, hi | lo)
      if (g * kCW < st0 * kWK) {
        uint32_t hh[kCW], ll[kCW];
#pragma unroll
        for (int c = 0; c < kCW; ++c) {
          const int kcol = g * kCW + c;
          const float v = (kcol < a.n_in) ? xs[row * D + s_id[kcol < a.n_in ? kcol : 0]] : 0.f;
          if (a.xid && kcol < a.n_in && m0 + row < a.B) a.xid[(m0 + row) * a.ldxid + kcol] = v;
          float hi, lo;
          split_tf32(v, hi, lo);
          hh[c] = __float_as_uint(hi); ll[c] = __float_as_uint(lo);
        }
        tmem_st<kCW>(t_lane + g * kCW, hh);
        tmem_st<kCW>(t_lane + col_alo + g * kCW, ll);
        tmem_st_wait();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&a_ready[0]);
      if (!INV && a.Ts > 0) {           // density direction: the net saw the raw id columns, the output gets their CDF
        epi_bar();
        for (int j = g; j < a.Ts; j += kG) {
          float o, l;
          int bin;
          fast::table_element<8, false>(rec + j * 9, xs[row * D + s_id[j]], s_sbound[j], a.outside_zero, &o, &l, &bin, &flag);
          xs[row * D + s_id[j]] = o;
          ldsum += l;
        }
      }
      return ldsum;
    };
    float ld_next = 0.f;
    if (static_cast<int64_t>(blockIdx.x) < n_tiles) ld_next = prologue(blockIdx.x, 0);
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++k) {
      const int64_t m0 = tile * kRows;
      const uint32_t xb = k & 1u;
      float* xs = xs0 + xb * (kRows * D);
      const bool has_next = tile + gridDim.x < n_tiles;
      float ldsum = ld_next;

      FTRACE(1, ei); ++ei;
      for (int u = 0; u < L; ++u, ++uc) {
        mbar_wait(acc_full, uc & 1u);
        tc_fence_after();
        FTRACE(1, ei); ++ei;
        float v[4 * kCW];                              // this thread's 4 x kCW accumulator columns, read out at once
        {
          uint32_t r[4 * kCW], r2[4 * kCW];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            tmem_ld<kCW>(t_lane + col_main + j * 32 + g * kCW, r + j * kCW);
            tmem_ld<kCW>(t_lane + col_cross + j * 32 + g * kCW, r2 + j * kCW);
          }
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 4 * kCW; ++i) v[i] = __uint_as_float(r[i]) + __uint_as_float(r2[i]);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(acc_empty);         // the next unit may start accumulating
        FTRACE(1, ei); ++ei;
        const bool add_res = u >= 2 && (u & 1) == 0, write_res = (u & 1) == 0, relu_next = u != L - 1;
        const float* bu = bias + u * kH;
        float* act_row = a.acts ? a.acts + static_cast<int64_t>(u) * a.act_stride + (m0 + row) * kH : nullptr;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int c0 = j * 32 + g * kCW;
          uint32_t hh[kCW], ll[kCW];
          float bb[kCW], rr[kCW];                        // all loads of the chunk first: few warps per scheduler
#pragma unroll                                           // cannot hide a load -> add -> store chain per element
          for (int i = 0; i < kCW; i += 4) {
            const float4 b4 = *reinterpret_cast<const float4*>(bu + c0 + i);
            bb[i] = b4.x; bb[i + 1] = b4.y; bb[i + 2] = b4.z; bb[i + 3] = b4.w;
          }
#pragma unroll
          for (int i = 0; i < kCW; ++i) rr[i] = add_res ? res[(c0 + i) * kRS + row] : 0.f;
#pragma unroll
          for (int i = 0; i < kCW; ++i) {
            const float t = (v[j * kCW + i] + bb[i]) + rr[i];
            if (write_res) res[(c0 + i) * kRS + row] = t;
            v[j * kCW + i] = t;
            float hi, lo;
            split_tf32(relu_next ? fmaxf(t, 0.f) : t, hi, lo);
            hh[i] = __float_as_uint(hi); ll[i] = __float_as_uint(lo);
          }
          tmem_st<kCW>(t_lane + c0, hh);
          tmem_st<kCW>(t_lane + col_alo + c0, ll);
          tmem_st_wait();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&a_ready[j]);
          if (act_row && m0 + row < a.B) {
#pragma unroll
            for (int i = 0; i < kCW; i += 4)
              *reinterpret_cast<float4*>(act_row + c0 + i) =
                  make_float4(v[j * kCW + i], v[j * kCW + i + 1], v[j * kCW + i + 2], v[j * kCW + i + 3]);
          }
        }
        FTRACE(1, ei); ++ei;
      }

      for (int s = 0; s < S; ++s, ++uc) {
        mbar_wait(acc_full, uc & 1u);
        tc_fence_after();
        FTRACE(1, ei); ++ei;
        const float* bu = bias + (L + s) * kH;
        if (s > 0) epi_bar();                          // every thread has read the previous slice's staged parameters
        {                                              // accumulator -> [column][row] staging (the residual buffer is free)
          uint32_t r[4 * kCW], r2[4 * kCW];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            tmem_ld<kCW>(t_lane + col_main + j * 32 + g * kCW, r + j * kCW);
            tmem_ld<kCW>(t_lane + col_cross + j * 32 + g * kCW, r2 + j * kCW);
          }
          float bb[4 * kCW];
#pragma unroll
          for (int i = 0; i < 4 * kCW; i += 4) {
            const float4 b4 = *reinterpret_cast<const float4*>(bu + (i / kCW) * 32 + g * kCW + (i % kCW));
            bb[i] = b4.x; bb[i + 1] = b4.y; bb[i + 2] = b4.z; bb[i + 3] = b4.w;
          }
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 4 * kCW; ++i) {
            const int c = (i / kCW) * 32 + g * kCW + (i % kCW);
            res[c * kRS + row] = __uint_as_float(r[i]) + __uint_as_float(r2[i]) + bb[i];
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(acc_empty);
        FTRACE(1, ei); ++ei;
        if (s == S - 1 && has_next) ld_next = prologue(tile + gridDim.x, k + 1);
        epi_bar();
        const int nf = (a.Tc - s * a.fs) < a.fs ? (a.Tc - s * a.fs) : a.fs;
        if (a.params) {                                // the slice's parameters, one row per warp pass: coalesced
          for (int rr = ew; rr < kRows; rr += kEpiWarps) {
            if (m0 + rr < a.B && lane < a.Pc) {
              float* pr = a.params + (m0 + rr) * a.ldp + (s * a.fs) * a.Pc + lane;
              for (int fi = 0; fi < nf; ++fi) pr[fi * a.Pc] = res[(fi * PCP + lane) * kRS + rr];
            }
          }
        }
        for (int fi = g; fi < nf; fi += kG) {
          const int f = s * a.fs + fi;
          float prm[PCP];
#pragma unroll
          for (int i = 0; i < PCP; ++i) prm[i] = res[(fi * PCP + i) * kRS + row];
          const int col = s_tr[f];
          const float xv = xs[row * D + col];
          float o, l;
          element_regs<8, INV, PCP>(prm, xv, s_bound[f], a.wh_scale, s_mode[f], a.outside_zero, &o, &l, &flag);
          xs[row * D + col] = o;
          ldsum += l;
        }
        FTRACE(1, ei); ++ei;
      }
      if (g > 0) lds[(g - 1) * kRows + row] = ldsum;
      epi_bar();
      if (g == 0 && m0 + row < a.B) {
        float t = ldsum;
#pragma unroll
        for (int gg = 1; gg < kG; ++gg) t += lds[(gg - 1) * kRows + row];
        a.logdet[m0 + row] = a.accumulate ? a.logdet[m0 + row] + t : t;
      }
      if (a.ldy == D && reinterpret_cast<uintptr_t>(a.y) % 16 == 0) {      // the tile is one contiguous block of y
        const int64_t n_valid = (a.B - m0 < kRows ? a.B - m0 : kRows) * D;   // floats
        float* yt = a.y + m0 * D;
        for (int e = et * 4; e < kRows * D; e += kEpiThreads * 4) {
          const float4 v4 = *reinterpret_cast<const float4*>(xs + e);
          if (e + 3 < n_valid) {
            *reinterpret_cast<float4*>(yt + e) = v4;
          } else {
            if (e < n_valid) yt[e] = v4.x;
            if (e + 1 < n_valid) yt[e + 1] = v4.y;
            if (e + 2 < n_valid) yt[e + 2] = v4.z;
          }
        }
      } else {
        for (int e = et; e < kRows * D; e += kEpiThreads) {
          const int r = e / D, c = e - r * D;
          if (m0 + r < a.B) a.y[(m0 + r) * a.ldy + c] = xs[e];
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&x_empty[xb]);
      FTRACE(1, ei); ++ei;
    }
    if (flag && a.err_flag) atomicOr(a.err_flag, 1);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, kTmemCols);
}

inline size_t smem_bytes(int nb, int S, int D) {
  return 1024 + kWStages * static_cast<size_t>(kWStage) + (static_cast<size_t>(kH) * kRS + 32) * 4 +
         (static_cast<size_t>(n_layers(nb) + S) * kH + (kG - 1) * kRows + 2 * static_cast<size_t>(kRows) * D) * 4 + kMaxD * 9 * 16;
}

}  // namespace fusednet
}  // namespace vcnf

using namespace vcnf;
using namespace vcnf::fusednet;

static bool shape_ok(int n_in, int hidden, int nb, int Tc, int Pc, int* PcP, int* fs, int* S) {
  if (hidden != kH || n_in < 1 || n_in > kMaxD || nb < 0 || nb > kMaxBlocks || Tc < 1 || Tc > kMaxD || Pc < 1 || Pc > 32)
    return false;
  *PcP = Pc <= 24 ? 24 : 32;
  *fs = kH / *PcP;
  *S = (Tc + *fs - 1) / *fs;
  return *S <= kMaxSlices;
}

extern "C" int64_t vcnf_resnet_spline_pack_bytes(int n_in, int hidden, int num_blocks, int Tc, int Pc) {
  int PcP, fs, S;
  if (!shape_ok(n_in, hidden, num_blocks, Tc, Pc, &PcP, &fs, &S)) return VCNF_E_UNSUPPORTED;
  return static_cast<int64_t>(packed_bytes(n_in, num_blocks, S));
}

extern "C" int vcnf_resnet_spline_pack(const void* const* W, const void* const* b, int n_in, int hidden, int num_blocks,
                                       int Tc, int Pc, void* packed, void* stream) {
  if (!W || !b || !packed) return VCNF_E_BADARG;
  PackArgs p = {};
  if (!shape_ok(n_in, hidden, num_blocks, Tc, Pc, &p.PcP, &p.fs, &p.S)) return VCNF_E_UNSUPPORTED;
  const int L = n_layers(num_blocks);
  for (int u = 0; u <= L; ++u) {
    if (!W[u]) return VCNF_E_BADARG;
    p.W[u] = static_cast<const float*>(W[u]);
    p.b[u] = static_cast<const float*>(b[u]);
  }
  p.n_in = n_in; p.nb = num_blocks; p.Tc = Tc; p.Pc = Pc;
  p.out = static_cast<unsigned char*>(packed);
  const int64_t n = static_cast<int64_t>(total_stages(n_in, num_blocks, p.S)) * kH * kWK + static_cast<int64_t>(L + p.S) * kH;
  const int64_t blocks = (n + 255) / 256;
  pack_kernel<<<static_cast<unsigned>(blocks < 1184 ? blocks : 1184), 256, 0, static_cast<cudaStream_t>(stream)>>>(p);
  return launch_status();
}

extern "C" int vcnf_resnet_spline_fwd(const void* x, int64_t ldx, void* y, int64_t ldy, void* logdet, int accumulate,
                                      int64_t B, int D, int K, const int32_t* id_cols, int n_in, const int32_t* tr_cols,
                                      int Tc, int Pc, int hidden, int num_blocks, const void* packed,
                                      const uint8_t* modes, const float* bounds, double wh_scale, int inverse,
                                      int outside_zero, int32_t* err_flag, const void* shared_params, int Ts, int Ps,
                                      const uint8_t* shared_modes, const float* shared_bounds, void* params,
                                      int64_t ldp, void* acts, void* xid, int64_t ldxid, void* stream) {
  if (Ts < 0 || (Ts > 0 && (!shared_params || !shared_modes || !shared_bounds))) return VCNF_E_BADARG;
  if (Ts != 0 && Ts != n_in) return VCNF_E_UNSUPPORTED;   // the shared splines act on the net's input columns
  if (!x || !y || !logdet || !id_cols || !tr_cols || !packed || !modes || !bounds || B < 0 || D < 1) return VCNF_E_BADARG;
  Args a = {};
  if (K != 8 || D > kMaxD || !shape_ok(n_in, hidden, num_blocks, Tc, Pc, &a.PcP, &a.fs, &a.n_slices)) return VCNF_E_UNSUPPORTED;
  if (B == 0) return 0;
  for (int i = 0; i < n_in; ++i) { if (id_cols[i] < 0 || id_cols[i] >= D) return VCNF_E_BADARG; a.id_cols[i] = id_cols[i]; }
  for (int i = 0; i < Tc; ++i) {
    if (tr_cols[i] < 0 || tr_cols[i] >= D) return VCNF_E_BADARG;
    a.tr_cols[i] = tr_cols[i]; a.modes[i] = modes[i]; a.bounds[i] = bounds[i];
  }
  a.x = static_cast<const float*>(x); a.ldx = ldx;
  a.y = static_cast<float*>(y); a.ldy = ldy;
  a.logdet = static_cast<float*>(logdet); a.accumulate = accumulate;
  a.B = B; a.D = D;
  a.packed = static_cast<const unsigned char*>(packed);
  a.n_in = n_in; a.nb = num_blocks; a.Tc = Tc; a.Pc = Pc;
  a.wh_scale = static_cast<float>(wh_scale);
  a.outside_zero = outside_zero;
  a.err_flag = err_flag;
  a.shared = static_cast<const float*>(shared_params); a.Ts = Ts; a.Ps = Ps;
  for (int i = 0; i < Ts; ++i) { a.smodes[i] = shared_modes[i]; a.sbounds[i] = shared_bounds[i]; }
  a.params = static_cast<float*>(params); a.ldp = ldp;
  a.acts = static_cast<float*>(acts); a.act_stride = B * kH;
  a.xid = static_cast<float*>(xid); a.ldxid = ldxid;
  const size_t smem = smem_bytes(num_blocks, a.n_slices, D);
  const int sms = sm_count() > 0 ? sm_count() : VCNF_SM_COUNT_DEFAULT;
  const int64_t tiles = (B + kRows - 1) / kRows;
  const unsigned grid = static_cast<unsigned>(tiles < sms ? tiles : sms);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
#define VCNF_FUSED_LAUNCH(INV, PCP)                                                                         \
  do {                                                                                                      \
    auto k = resnet_spline_kernel<INV, PCP>;                                                                \
    static bool attr = false;                                                                               \
    if (!attr) { cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024); attr = true; } \
    k<<<grid, kThreads, smem, st>>>(a);                                                                     \
  } while (0)
  if (a.PcP == 24) { if (inverse) VCNF_FUSED_LAUNCH(true, 24); else VCNF_FUSED_LAUNCH(false, 24); }
  else { if (inverse) VCNF_FUSED_LAUNCH(true, 32); else VCNF_FUSED_LAUNCH(false, 32); }
#undef VCNF_FUSED_LAUNCH
  return launch_status();
}

#ifdef VCNF_FUSED_TRACE
extern "C" int vcnf_debug_fused_trace(long long* out) {
  return static_cast<int>(cudaMemcpyFromSymbol(out, vcnf::fusednet::g_ftrace, sizeof(long long) * 256));
}
#endif
