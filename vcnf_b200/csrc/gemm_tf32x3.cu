// Conditioner dense layers on the 5th-generation tensor cores (tcgen05 + TMEM), fp32-faithful.
//
//   C[m, n] = [mask[m, n] > 0] * ( sum_k act(A[m, k]) * W[n, k] + bias[n] ) + residual[m, n]
//
// replaces the cuBLAS SIMT sgemm + elementwise chain behind nn.Linear in the reference's
// conditioners (normflow/nets/resnet.py:38-106 ResidualNet / ResidualBlock, nets/mlp.py:7-58,
// nets/made.py:79-80 after pre-masking) in BOTH directions: forward (A = activations, W = weight)
// and input-gradient (A = grad-output, W = weight^T through strides, mask = relu'(saved input)).
//
// Why 3xTF32: the stated parity tolerance (1e-4 rel on spline outputs) rules out plain TF32/BF16
// products (2^-11).  Every fp32 operand is split a = hi + lo with hi = rn_tf32(a) and lo =
// rn_tf32(a - hi) (|lo| <= 2^-11 |a|), and three tcgen05.mma.kind::tf32 chains lo*hi + hi*lo + hi*hi
// accumulate in fp32 TMEM accumulators (hi*hi and the cross terms separately, summed in the
// epilogue): relative error ~2^-22 per product, the level of an fp32 FMA chain.
//
// Structure (one CTA per SM, persistent over 128-row tiles of a fixed <=128-column slice of W):
//   warps 0-7  loaders : ld.global (128-bit, 8 rows x 64 B per warp request, three k-chunks in flight
//                        per thread) -> optional ReLU -> hi/lo split -> st.shared in the canonical
//                        no-swizzle K-major UMMA layout (8 x 16 B core matrices), fence.proxy.async,
//                        mbarrier arrive                                                [2-stage ring]
//   warp  12   MMA     : one elected lane issues 12 tcgen05.mma (M=128, N<=128, K=8) per 32-wide
//                        k-chunk; tcgen05.commit releases the stage / publishes the accumulator
//   warps 8-11 epilogue: tcgen05.ld 32x32b.x32 (TMEM -> registers), bias / mask / residual, 128-bit
//                        global stores; double-buffered accumulators (2 x 128 TMEM columns) let the
//                        epilogue of tile i overlap the MMAs of tile i+1
// W (hi and lo, <= 2 x 64 KB) stays resident in shared memory for the CTA's lifetime.
#include <cstdlib>
#include "common.cuh"
#include "../../include/vcnf_b200.h"

namespace vcnf {
namespace gemm {

constexpr int kLoaderWarps = 8, kEpiWarps = 4;
constexpr int kThreads = (kLoaderWarps + kEpiWarps + 1) * 32;     // 416
constexpr int kMmaWarp = kLoaderWarps + kEpiWarps;                // 12
constexpr int kBM = 128;          // rows per tile = UMMA M
constexpr int kBN = 128;          // max columns per CTA slice = max UMMA N used
constexpr int kBK = 16;           // fp32 elements per k-chunk (one stage)
constexpr int kStages = 4;        // handshake latency (arrive -> MMA -> commit -> refill) is hidden by depth, not size
constexpr int kLoadsPerChunk = (kBM * kBK / 4) / (kLoaderWarps * 32);   // float4 per loader thread per k-chunk
constexpr int kBlkPerGroup = kBK / 16;                                  // warp-loads (8 rows x 4 chunks) per 8-row group
constexpr uint32_t kASbo = (kBK / 4) * 128;                             // byte stride between 8-row groups of an A stage
constexpr int kMaxK = 128;        // W slice resident in smem
constexpr uint32_t kTmemCols = 512;   // 2 buffers x (hi*hi accumulator + cross-term accumulator) x 128 columns

// ---- raw PTX wrappers ---------------------------------------------------------------------------
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
// arrive on an mbarrier once every tcgen05 op issued so far by this thread has completed
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc]^T, kind::tf32, issued by ONE thread
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
// 32 lanes x 32 consecutive fp32 columns of the accumulator -> 32 registers per thread
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ---- descriptors (bit layouts: cute/arch/mma_sm100_desc.hpp, UMMA::SmemDescriptor / InstrDescriptor)
// Shared-memory matrix descriptor, K-major, SWIZZLE_NONE ("interleave"): 8 rows x 16 B core matrices,
// LBO = byte distance between the two 16 B k-chunks of one MMA (K = 8 tf32), SBO = byte distance
// between consecutive 8-row groups.  version = 1 (Blackwell), base_offset = 0, layout_type = 0.
__device__ __forceinline__ uint64_t smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((saddr >> 4) & 0x3fffu);
  d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3fffu) << 16;
  d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3fffu) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  return d;
}
// Instruction descriptor: D = F32 (bits 4-5 = 1), A = B = TF32 (bits 7-9 / 10-12 = 2), both K-major
// (bits 15, 16 = 0), N >> 3 at bits 17-22, M >> 4 at bits 24-28, dense, no negate, no saturate.
__host__ __device__ constexpr uint32_t instr_desc_tf32(int M, int N) {
  return (1u << 4) | (2u << 7) | (2u << 10) | (static_cast<uint32_t>(N >> 3) << 17) |
         (static_cast<uint32_t>(M >> 4) << 24);
}

struct Args {
  const float* A; int64_t lda;
  const float* W; int64_t sw_n, sw_k;
  const float* bias;
  const float* residual; int64_t ldr;
  const float* mask; int64_t ldm;
  float* C; int64_t ldc;
  int64_t M; int N; int K;
  int relu_in;
  int dbg;            // timing experiments only (VCNF_GEMM_DEBUG): 1 no global loads, 2 no MMA, 4 no stores
  int n_tiles;        // ceil(N / kBN)
  int ctas_per_ntile; // persistent CTAs sharing one column slice
  int KP;             // K rounded up to a multiple of kBK
};

// byte offset of element (row, k) inside a [rows x KP] operand in the canonical layout
__device__ __forceinline__ uint32_t canon_off(int row, int k, int KP) {
  return static_cast<uint32_t>((row >> 3) * (KP * 32) + (k >> 2) * 128 + (row & 7) * 16 + (k & 3) * 4);
}

// a = hi + lo with BOTH parts rounded to nearest TF32 (10 explicit mantissa bits).  The tensor core
// truncates whatever low bits it is given; with a truncating split the dropped part of lo is up to
// 2^-20 |a| (measured: max error / sum|terms| 5e-7 even for K = 15), with round-to-nearest it is
// <= 2^-23 |a|.  Integer rounding (add half an ulp of TF32, clear the low 13 bits) = cvt.rna.tf32.f32
// without its lower issue rate.
__device__ __forceinline__ float rn_tf32(float a) {
  return __uint_as_float((__float_as_uint(a) + 0x1000u) & 0xffffe000u);
}
__device__ __forceinline__ void split_tf32(float a, float& hi, float& lo) {
  hi = rn_tf32(a);
  lo = rn_tf32(a - hi);
}

__global__ void __launch_bounds__(kThreads, 1) linear_tf32x3_kernel(const Args a) {
  extern __shared__ __align__(1024) unsigned char smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem);            // [kStages] loaders -> MMA
  uint64_t* empty = full + kStages;                              // [kStages] MMA -> loaders
  uint64_t* acc_full = empty + kStages;                          // [2] MMA -> epilogue
  uint64_t* acc_empty = acc_full + 2;                            // [2] epilogue -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);
  const int KP = a.KP;
  const uint32_t w_bytes = static_cast<uint32_t>(kBN) * KP * 4;
  unsigned char* w_hi = smem + 1024;
  unsigned char* w_lo = w_hi + w_bytes;
  unsigned char* a_st = w_lo + w_bytes;                          // stage s: [hi 16 KB | lo 16 KB]
  constexpr uint32_t kAHalf = kBM * kBK * 4;
  unsigned char* epi_st = a_st + kStages * 2 * kAHalf;           // 4 x [32][36] fp32 transpose tiles

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int nt = blockIdx.x % a.n_tiles;
  const int n0 = nt * kBN;
  const int NT = (a.N - n0) < kBN ? (a.N - n0) : kBN;            // valid columns of this slice
  const int UN = (NT + 15) & ~15;                                // UMMA N
  const int64_t m_tiles = (a.M + kBM - 1) / kBM;
  const int64_t mt0 = blockIdx.x / a.n_tiles;
  const int nk = KP / kBK;                                       // k-chunks per tile
  const int nk8 = (a.K + 7) >> 3;                                // 8-wide MMA steps that carry data

  if (tid == 0) {
    for (int s = 0; s < kStages; ++s) { mbar_init(&full[s], kLoaderWarps); mbar_init(&empty[s], 1); }   // one arrival per WARP:
    for (int b = 0; b < 2; ++b) { mbar_init(&acc_full[b], 1); mbar_init(&acc_empty[b], kEpiWarps); }      // 256 same-word arrivals serialise
    mbar_init_fence();
  }
  if (warp == kMmaWarp) tmem_alloc(tmem_slot, kTmemCols);

  // resident W slice: hi / lo planes, zero padded to [kBN x KP]
  {
    const bool k_fast = a.sw_k == 1;
    const bool w_vec = reinterpret_cast<uintptr_t>(a.W) % 16 == 0 &&
                       (k_fast ? (a.sw_n % 4 == 0 && a.K % 4 == 0) : (a.sw_n == 1 && a.sw_k % 4 == 0 && n0 % 4 == 0 && NT % 4 == 0));
    if (w_vec) {
      // 128-bit loads, all of a thread's loads issued before the first use
      const int total4 = kBN * KP / 4;
      constexpr int kU = 4;
      for (int base = tid; base < total4; base += kThreads * kU) {
        float4 v[kU];
        int nn[kU], kk[kU];
#pragma unroll
        for (int u = 0; u < kU; ++u) {
          const int idx = base + u * kThreads;
          int n, k;
          if (k_fast) { n = idx / (KP / 4); k = (idx - n * (KP / 4)) * 4; } else { k = idx / (kBN / 4); n = (idx - k * (kBN / 4)) * 4; }
          nn[u] = n; kk[u] = k;
          v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (idx < total4 && n < NT && k < a.K)
            v[u] = __ldg(reinterpret_cast<const float4*>(a.W + static_cast<int64_t>(n0 + n) * a.sw_n + static_cast<int64_t>(k) * a.sw_k));
        }
#pragma unroll
        for (int u = 0; u < kU; ++u) {
          if (base + u * kThreads >= total4) continue;
          const float e[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
          float hi[4], lo[4];
#pragma unroll
          for (int c = 0; c < 4; ++c) split_tf32(e[c], hi[c], lo[c]);
          if (k_fast) {                                   // 4 consecutive k of one row = one 16 B chunk
            const uint32_t off = canon_off(nn[u], kk[u], KP);
            *reinterpret_cast<float4*>(w_hi + off) = make_float4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<float4*>(w_lo + off) = make_float4(lo[0], lo[1], lo[2], lo[3]);
          } else {                                        // 4 consecutive rows at one k
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              const uint32_t off = canon_off(nn[u] + c, kk[u], KP);
              *reinterpret_cast<float*>(w_hi + off) = hi[c];
              *reinterpret_cast<float*>(w_lo + off) = lo[c];
            }
          }
        }
      }
    } else {
    const int total = kBN * KP;
    constexpr int kU = 8;                                  // independent loads in flight per thread
    for (int base = tid; base < total; base += kThreads * kU) {
      float v[kU];
      int nn[kU], kk[kU];
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const int idx = base + u * kThreads;
        int n, k;
        if (k_fast) { n = idx / KP; k = idx - n * KP; } else { k = idx / kBN; n = idx - k * kBN; }
        nn[u] = n; kk[u] = k;
        v[u] = 0.f;
        if (idx < total && n < NT && k < a.K)
          v[u] = __ldg(a.W + static_cast<int64_t>(n0 + n) * a.sw_n + static_cast<int64_t>(k) * a.sw_k);
      }
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        if (base + u * kThreads >= total) continue;
        float hi, lo;
        split_tf32(v[u], hi, lo);
        const uint32_t off = canon_off(nn[u], kk[u], KP);
        *reinterpret_cast<float*>(w_hi + off) = hi;
        *reinterpret_cast<float*>(w_lo + off) = lo;
      }
    }
    }
    fence_async_smem();          // generic-proxy writes -> visible to the tensor-core (async) proxy
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < kLoaderWarps) {
    // ===================================== loaders =====================================
    // The (tile, k-chunk) pairs of this CTA form one linear stream g = 0 .. G-1.  Each thread keeps the
    // global loads of THREE chunks in flight (register ring b0/b1/b2, 4 x 128 bit each): with 256 loader
    // threads that is 48 KB per SM outstanding -- what HBM latency x bandwidth needs (Little's law).
    const bool vec_ok = (a.lda % 4 == 0) && (reinterpret_cast<uintptr_t>(a.A) % 16 == 0);
    const int64_t my_tiles = mt0 < m_tiles ? (m_tiles - mt0 + a.ctas_per_ntile - 1) / a.ctas_per_ntile : 0;
    const int64_t G = my_tiles * nk;
    auto load_chunk = [&](float4 (&b)[kLoadsPerChunk], int64_t g) {
      if (g >= G) return;
      const int64_t m0 = (mt0 + (g / nk) * a.ctas_per_ntile) * kBM;
      const int k0 = static_cast<int>(g % nk) * kBK;
#pragma unroll
      for (int i = 0; i < kLoadsPerChunk; ++i) {
        const int q = i * (kLoaderWarps * 32) + tid;
        const int blk = q >> 5, l = q & 31;
        const int row = (blk / kBlkPerGroup) * 8 + (l & 7);
        const int chunk = (blk % kBlkPerGroup) * 4 + (l >> 3);
        const int64_t gr = m0 + row;
        const int gk = k0 + chunk * 4;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (gr < a.M && !(a.dbg & 1)) {
          const float* src = a.A + gr * a.lda + gk;
          if (vec_ok && gk + 3 < a.K) {
            v = __ldg(reinterpret_cast<const float4*>(src));
          } else {
            if (gk + 0 < a.K) v.x = __ldg(src + 0);
            if (gk + 1 < a.K) v.y = __ldg(src + 1);
            if (gk + 2 < a.K) v.z = __ldg(src + 2);
            if (gk + 3 < a.K) v.w = __ldg(src + 3);
          }
        }
        b[i] = v;
      }
    };
    auto store_chunk = [&](const float4 (&b)[kLoadsPerChunk], int64_t g) {
      if (g >= G) return;
      const uint32_t it = static_cast<uint32_t>(g);
      const int s = it % kStages;
      mbar_wait(&empty[s], ((it / kStages) & 1u) ^ 1u);
      unsigned char* hi_base = a_st + static_cast<uint32_t>(s) * 2 * kAHalf;
      unsigned char* lo_base = hi_base + kAHalf;
#pragma unroll
      for (int i = 0; i < kLoadsPerChunk; ++i) {
        const int q = i * (kLoaderWarps * 32) + tid;
        const int blk = q >> 5, l = q & 31;
        const int row = (blk / kBlkPerGroup) * 8 + (l & 7);
        const int chunk = (blk % kBlkPerGroup) * 4 + (l >> 3);
        float4 v = b[i];
        if (a.relu_in) { v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f); }
        float4 h, lo4;
        split_tf32(v.x, h.x, lo4.x); split_tf32(v.y, h.y, lo4.y);
        split_tf32(v.z, h.z, lo4.z); split_tf32(v.w, h.w, lo4.w);
        const uint32_t off = static_cast<uint32_t>(row >> 3) * kASbo + chunk * 128 + (row & 7) * 16;
        *reinterpret_cast<float4*>(hi_base + off) = h;
        *reinterpret_cast<float4*>(lo_base + off) = lo4;
      }
      fence_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(&full[s]);
    };
    float4 b0[kLoadsPerChunk], b1[kLoadsPerChunk], b2[kLoadsPerChunk];
    load_chunk(b0, 0);
    load_chunk(b1, 1);
    for (int64_t g = 0; g < G; g += 3) {
      load_chunk(b2, g + 2);
      store_chunk(b0, g);
      load_chunk(b0, g + 3);
      store_chunk(b1, g + 1);
      load_chunk(b1, g + 4);
      store_chunk(b2, g + 2);
    }
  } else if (warp == kMmaWarp) {
    // ===================================== MMA issuer =====================================
    const uint32_t idesc = instr_desc_tf32(kBM, UN);
    const uint32_t w_sbo = static_cast<uint32_t>(KP) * 32;
    uint32_t it = 0, tc = 0;
    for (int64_t mt = mt0; mt < m_tiles; mt += a.ctas_per_ntile, ++tc) {
      const uint32_t buf = tc & 1u;
      mbar_wait(&acc_empty[buf], ((tc >> 1) & 1u) ^ 1u);
      tc_fence_after();
      // Two accumulators per tile: the tensor core adds into the fp32 accumulator with truncation, so the
      // error grows with the number of accumulate steps on the LARGE sum.  hi*hi (16 steps for K = 128)
      // goes to d_main; the 2^-11 smaller cross terms lo*hi + hi*lo go to d_small and are added once in
      // the epilogue (measured: max error / sum|terms| 1.0e-6 with one accumulator).
      const uint32_t d_main = tmem_base + buf * (2 * kBN);
      const uint32_t d_small = d_main + kBN;
      for (int kc = 0; kc < nk; ++kc, ++it) {
        const int s = it % kStages;
        mbar_wait(&full[s], (it / kStages) & 1u);
        tc_fence_after();
        if (lane == 0) {
          const uint32_t a_hi = smem_u32(a_st + static_cast<uint32_t>(s) * 2 * kAHalf);
          const uint32_t a_lo = a_hi + kAHalf;
          const uint32_t b_hi = smem_u32(w_hi) + static_cast<uint32_t>(kc) * (kBK / 4) * 128;
          const uint32_t b_lo = smem_u32(w_lo) + static_cast<uint32_t>(kc) * (kBK / 4) * 128;
          const int steps = (nk8 - kc * (kBK / 8)) < (kBK / 8) ? (nk8 - kc * (kBK / 8)) : (kBK / 8);
#pragma unroll
          for (int combo = 0; combo < 3; ++combo) {          // small terms first: lo*hi, hi*lo, hi*hi
            const uint32_t ab = combo == 0 ? a_lo : a_hi;
            const uint32_t bb = combo == 1 ? b_lo : b_hi;
            for (int kk = 0; kk < steps; ++kk) {
              const uint64_t ad = smem_desc(ab + kk * 256, 128, kASbo);
              const uint64_t bd = smem_desc(bb + kk * 256, 128, w_sbo);
              if (a.dbg & 2) continue;
              if (combo == 2) mma_tf32(d_main, ad, bd, idesc, (kc > 0 || kk > 0) ? 1u : 0u);
              else mma_tf32(d_small, ad, bd, idesc, (kc > 0 || combo > 0 || kk > 0) ? 1u : 0u);
            }
          }
          tc_commit(&empty[s]);                              // stage reusable once these MMAs retire
          if (kc == nk - 1) tc_commit(&acc_full[buf]);       // accumulator complete
        }
        __syncwarp();
      }
    }
  } else {
    // ===================================== epilogue =====================================
    // TMEM gives every lane one ROW (32 consecutive columns in registers).  Written out directly that
    // is 32 different 128 B lines per store instruction (measured: 15 sectors per request, 7 % of DRAM
    // peak).  Each warp therefore transposes its 32 x 32 block through a padded shared-memory tile so
    // that one instruction reads / writes one full 128 B line of a row (lane = column).
    const int q = warp & 3;                                   // TMEM lane quarter this warp may read
    constexpr int kTS = 36;                                   // tile row stride (floats): 16 B aligned, conflict-free
    float* tbuf = reinterpret_cast<float*>(epi_st) + (warp - kLoaderWarps) * (32 * kTS);
    const bool vec_out = (a.ldc % 4 == 0) && (a.N % 4 == 0) && (reinterpret_cast<uintptr_t>(a.C) % 16 == 0) &&
                         (!a.bias || reinterpret_cast<uintptr_t>(a.bias) % 16 == 0) &&
                         (!a.residual || (a.ldr % 4 == 0 && reinterpret_cast<uintptr_t>(a.residual) % 16 == 0)) &&
                         (!a.mask || (a.ldm % 4 == 0 && reinterpret_cast<uintptr_t>(a.mask) % 16 == 0));
    uint32_t tc = 0;
    for (int64_t mt = mt0; mt < m_tiles; mt += a.ctas_per_ntile, ++tc) {
      const uint32_t buf = tc & 1u;
      mbar_wait(&acc_full[buf], (tc >> 1) & 1u);
      tc_fence_after();
      const int64_t row0 = mt * kBM + q * 32;
      int rows = static_cast<int>(a.M - row0 < 32 ? a.M - row0 : 32);
      if (rows < 0) rows = 0;
      for (int cb = 0; cb * 32 < NT; ++cb) {
        uint32_t r[32], r2[32];
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + buf * (2 * kBN) + cb * 32;
        tmem_ld32(taddr, r);
        tmem_ld32(taddr + kBN, r2);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; j += 4)                        // lane = row: 8 x 16 B, conflict-free
          *reinterpret_cast<float4*>(tbuf + lane * kTS + j) =
              make_float4(__uint_as_float(r[j]) + __uint_as_float(r2[j]), __uint_as_float(r[j + 1]) + __uint_as_float(r2[j + 1]),
                          __uint_as_float(r[j + 2]) + __uint_as_float(r2[j + 2]), __uint_as_float(r[j + 3]) + __uint_as_float(r2[j + 3]));
        __syncwarp();
        if (a.dbg & 4) {
        } else if (vec_out) {
          // read back with lane = (row group of 4, 16 B column chunk): one instruction moves 4 rows x 128 B.
          // All residual / mask loads of the block are issued before the first store (the compiler cannot
          // move loads across stores through possibly-aliasing pointers itself).
          const int cc = (lane & 7) * 4;                        // column offset inside the 32-column block
          const int col = n0 + cb * 32 + cc;
          const bool col_ok = col + 3 < a.N;                    // N % 4 == 0 on this path
          float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
          if (a.bias && col_ok) bv = __ldg(reinterpret_cast<const float4*>(a.bias + col));
          float4 rv[8], mv[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int rr = i * 4 + (lane >> 3);
            const bool ok = rr < rows && col_ok;
            rv[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            mv[i] = make_float4(1.f, 1.f, 1.f, 1.f);
            if (a.residual && ok) rv[i] = __ldg(reinterpret_cast<const float4*>(a.residual + (row0 + rr) * a.ldr + col));
            if (a.mask && ok) mv[i] = __ldg(reinterpret_cast<const float4*>(a.mask + (row0 + rr) * a.ldm + col));
          }
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int rr = i * 4 + (lane >> 3);
            if (!(rr < rows && col_ok)) continue;
            float4 t = *reinterpret_cast<const float4*>(tbuf + rr * kTS + cc);
            t.x = (mv[i].x > 0.f ? t.x + bv.x : 0.f) + rv[i].x;
            t.y = (mv[i].y > 0.f ? t.y + bv.y : 0.f) + rv[i].y;
            t.z = (mv[i].z > 0.f ? t.z + bv.z : 0.f) + rv[i].z;
            t.w = (mv[i].w > 0.f ? t.w + bv.w : 0.f) + rv[i].w;
            *reinterpret_cast<float4*>(a.C + (row0 + rr) * a.ldc + col) = t;
          }
        } else {
          // unaligned leading dimensions (e.g. the 322-wide head): lane = column, one 128 B row segment
          // per instruction
          const int col = n0 + cb * 32 + lane;
          if (col < a.N) {
            const float bv = a.bias ? __ldg(a.bias + col) : 0.f;
            float* cp = a.C + row0 * a.ldc + col;
            const float* rp = a.residual ? a.residual + row0 * a.ldr + col : nullptr;
            const float* mp = a.mask ? a.mask + row0 * a.ldm + col : nullptr;
#pragma unroll 8
            for (int rr = 0; rr < rows; ++rr) {
              float t = tbuf[rr * kTS + lane] + bv;
              if (mp && !(__ldg(mp + rr * a.ldm) > 0.f)) t = 0.f;
              if (rp) t += __ldg(rp + rr * a.ldr);
              cp[rr * a.ldc] = t;
            }
          }
        }
        __syncwarp();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[buf]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == kMmaWarp) tmem_dealloc(tmem_base, kTmemCols);
}


// =================================================================================================
// Weight gradient:  gW[o, i] = sum_b gY[b, o] * act(X[b, i]),   gb[o] = sum_b gY[b, o]
// (autograd of F.linear in the conditioners: the cuBLAS "nt" sgemm + the bias-gradient reduce_kernel,
// 25 ms of the first measured 60 ms training step of cfg 2).
//
// The reduction runs over the BATCH, which is the slow index of both operands in HBM, while the
// tensor core wants K(= batch)-major operands (tf32 MN-major no-swizzle operands returned zeros on
// B200, tools/gemm_test.cu probe).  So every loader thread transposes in registers: it loads a 4
// samples x 4 features block with four coalesced 128-bit loads (a warp reads 512 contiguous bytes of
// one sample row per instruction) and stores four 16 B chunks "feature row x 4 consecutive samples" of
// the canonical K-major layout.  The 8-row-group stride (SBO) is skewed by 16 B (1040 instead of 1024)
// so that the 8 lanes of a store phase hit 8 different 16 B bank groups.
// UMMA M = output features (<= 128 per CTA), N = input features + a ones column (-> gb for free), K =
// samples.  Split-K over the batch: every CTA accumulates its share in TMEM and writes one partial;
// wgrad_reduce_kernel sums the partials in a fixed order (deterministic).
// =================================================================================================
constexpr int kWStages = 3;
constexpr int kWBK = 32;             // samples per stage (4 MMA k-steps of 8)
constexpr uint32_t kWSbo = 1024 + 16;

struct WArgs {
  const float* gY; int64_t ldg;      // [B, N]
  const float* X; int64_t ldx;       // [B, K]
  float* partial;                    // [splits, o_tiles * 128, K + 1]
  int64_t B; int N; int K;
  int relu_in;
  int o_tiles, splits;
  int NX;                            // K + 1 rounded up to 16 = UMMA N
  int64_t chunks;                    // ceil(B / kWBK)
};

// Thread item: feature chunk c = tid & 31 (features 4c .. 4c+3), sample group sg = tid >> 5 (samples
// 4sg .. 4sg+3 of the stage).  v[j] = the four features of sample 4sg + j.
__device__ __forceinline__ void wload(float4 (&v)[4], const float* src, int64_t ld, int f0, int n_valid, int n_chunks,
                                      int64_t b0, int64_t B, bool vec_ok, int tid) {
  const int c = tid & 31, sg = tid >> 5;
  const int f = c * 4;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
    const int64_t b = b0 + sg * 4 + j;
    if (c < n_chunks && b < B) {
      const float* p = src + b * ld + f0 + f;
      if (vec_ok && f + 3 < n_valid) {
        t = __ldg(reinterpret_cast<const float4*>(p));
      } else {
        if (f + 0 < n_valid) t.x = __ldg(p + 0);
        if (f + 1 < n_valid) t.y = __ldg(p + 1);
        if (f + 2 < n_valid) t.z = __ldg(p + 2);
        if (f + 3 < n_valid) t.w = __ldg(p + 3);
      }
    }
    v[j] = t;
  }
}

// Transpose the 4x4 block and store the hi / lo planes.  ones_at: feature index that must read 1.0
// (bias-gradient column) when it falls inside this thread's chunk, -1 otherwise.
__device__ __forceinline__ void wstore(const float4 (&v)[4], unsigned char* hi, unsigned char* lo, int n_chunks, bool relu,
                                       int ones_at, int tid) {
  const int c = tid & 31, sg = tid >> 5;
  if (c >= n_chunks) return;
  const float in[4][4] = {{v[0].x, v[0].y, v[0].z, v[0].w}, {v[1].x, v[1].y, v[1].z, v[1].w},
                          {v[2].x, v[2].y, v[2].z, v[2].w}, {v[3].x, v[3].y, v[3].z, v[3].w}};
#pragma unroll
  for (int r = 0; r < 4; ++r) {                           // feature row 4c + r: samples 4sg .. 4sg+3
    const int row = c * 4 + r;
    float e[4], h[4], l[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      e[j] = in[j][r];
      if (relu) e[j] = fmaxf(e[j], 0.f);
      if (row == ones_at) e[j] = 1.f;
      split_tf32(e[j], h[j], l[j]);
    }
    const uint32_t off = static_cast<uint32_t>(row >> 3) * kWSbo + static_cast<uint32_t>(sg) * 128 + (row & 7) * 16;
    *reinterpret_cast<float4*>(hi + off) = make_float4(h[0], h[1], h[2], h[3]);
    *reinterpret_cast<float4*>(lo + off) = make_float4(l[0], l[1], l[2], l[3]);
  }
}

__global__ void __launch_bounds__(kThreads, 1) linear_wgrad_tf32x3_kernel(const WArgs a) {
  extern __shared__ __align__(1024) unsigned char smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem);            // [kWStages]
  uint64_t* empty = full + kWStages;                             // [kWStages]
  uint64_t* acc_full = empty + kWStages;                         // [1]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 1);
  const int NX = a.NX;
  const uint32_t a_half = (128 / 8) * kWSbo;                     // gY operand plane (hi or lo)
  const uint32_t b_half = static_cast<uint32_t>(NX / 8) * kWSbo; // X operand plane
  const uint32_t stage_bytes = 2 * a_half + 2 * b_half;
  unsigned char* st0 = smem + 1024;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int ot = blockIdx.x % a.o_tiles;
  const int split = blockIdx.x / a.o_tiles;
  const int o0 = ot * 128;
  const int o_valid = (a.N - o0) < 128 ? (a.N - o0) : 128;
  const int64_t per = (a.chunks + a.splits - 1) / a.splits;
  const int64_t c_begin = split * per;
  const int64_t c_end = (c_begin + per) < a.chunks ? (c_begin + per) : a.chunks;
  const int64_t G = c_end > c_begin ? c_end - c_begin : 0;
  const int x_chunks = (a.K + 3) / 4;                            // feature chunks of X that hold data
  const int ones_in_data = (a.K % 4 != 0) ? a.K : -1;            // ones column shares a chunk with data

  if (tid == 0) {
    for (int s = 0; s < kWStages; ++s) { mbar_init(&full[s], kLoaderWarps); mbar_init(&empty[s], 1); }
    mbar_init(acc_full, 1);
    mbar_init_fence();
  }
  if (warp == kMmaWarp) tmem_alloc(tmem_slot, kTmemCols);
  // constant rows of the X operand (features >= 4 * x_chunks): ones column (if not inside a data chunk)
  // and zero padding up to NX, written once for every stage
  for (int idx = tid; idx < kWStages * (NX - 4 * x_chunks) * (kWBK / 4); idx += kThreads) {
    const int per_stage = (NX - 4 * x_chunks) * (kWBK / 4);
    const int s = idx / per_stage, rem = idx - s * per_stage;
    const int row = 4 * x_chunks + rem / (kWBK / 4), sg = rem % (kWBK / 4);
    const float val = (row == a.K) ? 1.f : 0.f;
    unsigned char* bh = st0 + static_cast<uint32_t>(s) * stage_bytes + 2 * a_half;
    const uint32_t off = static_cast<uint32_t>(row >> 3) * kWSbo + static_cast<uint32_t>(sg) * 128 + (row & 7) * 16;
    *reinterpret_cast<float4*>(bh + off) = make_float4(val, val, val, val);
    *reinterpret_cast<float4*>(bh + b_half + off) = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  fence_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < kLoaderWarps) {
    // ------------------------------ loaders: chunk i+1 in registers while chunk i is stored
    const bool vec_g = (a.ldg % 4 == 0) && (reinterpret_cast<uintptr_t>(a.gY) % 16 == 0) && (o0 % 4 == 0);
    const bool vec_x = (a.ldx % 4 == 0) && (reinterpret_cast<uintptr_t>(a.X) % 16 == 0);
    float4 ga[4], xa[4], gb[4], xb[4];
    auto load = [&](float4 (&g)[4], float4 (&x)[4], int64_t i) {
      if (i >= G) return;
      const int64_t b0 = (c_begin + i) * kWBK;
      wload(g, a.gY, a.ldg, o0, o_valid, 32, b0, a.B, vec_g, tid);
      wload(x, a.X, a.ldx, 0, a.K, x_chunks, b0, a.B, vec_x, tid);
    };
    auto store = [&](const float4 (&g)[4], const float4 (&x)[4], int64_t i) {
      if (i >= G) return;
      const uint32_t it = static_cast<uint32_t>(i);
      const int s = it % kWStages;
      mbar_wait(&empty[s], ((it / kWStages) & 1u) ^ 1u);
      unsigned char* st = st0 + static_cast<uint32_t>(s) * stage_bytes;
      wstore(g, st, st + a_half, 32, false, -1, tid);
      wstore(x, st + 2 * a_half, st + 2 * a_half + b_half, x_chunks, a.relu_in != 0, ones_in_data, tid);
      fence_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(&full[s]);
    };
    load(ga, xa, 0);
    for (int64_t i = 0; i < G; i += 2) {
      load(gb, xb, i + 1);
      store(ga, xa, i);
      load(ga, xa, i + 2);
      store(gb, xb, i + 1);
    }
  } else if (warp == kMmaWarp) {
    // ------------------------------ MMA issuer
    const uint32_t idesc = instr_desc_tf32(128, NX);
    const uint32_t d_main = tmem_base, d_small = tmem_base + 256;
    for (int64_t i = 0; i < G; ++i) {
      const uint32_t it = static_cast<uint32_t>(i);
      const int s = it % kWStages;
      mbar_wait(&full[s], (it / kWStages) & 1u);
      tc_fence_after();
      if (lane == 0) {
        const uint32_t a_hi = smem_u32(st0 + static_cast<uint32_t>(s) * stage_bytes);
        const uint32_t a_lo = a_hi + a_half;
        const uint32_t b_hi = a_hi + 2 * a_half;
        const uint32_t b_lo = b_hi + b_half;
#pragma unroll
        for (int combo = 0; combo < 3; ++combo) {
          const uint32_t ab = combo == 0 ? a_lo : a_hi;
          const uint32_t bb = combo == 1 ? b_lo : b_hi;
#pragma unroll
          for (int kk = 0; kk < kWBK / 8; ++kk) {
            const uint64_t ad = smem_desc(ab + kk * 256, 128, kWSbo);
            const uint64_t bd = smem_desc(bb + kk * 256, 128, kWSbo);
            if (combo == 2) mma_tf32(d_main, ad, bd, idesc, (i > 0 || kk > 0) ? 1u : 0u);
            else mma_tf32(d_small, ad, bd, idesc, (i > 0 || combo > 0 || kk > 0) ? 1u : 0u);
          }
        }
        tc_commit(&empty[s]);
        if (i == G - 1) tc_commit(acc_full);
      }
      __syncwarp();
    }
  } else {
    // ------------------------------ epilogue: one partial [128, K + 1] per CTA
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const int W1 = a.K + 1;
    float* dst = a.partial + (static_cast<int64_t>(split) * a.o_tiles * 128 + o0 + row) * W1;
    if (G > 0) {
      mbar_wait(acc_full, 0);
      tc_fence_after();
    }
    for (int cb = 0; cb * 32 < W1; ++cb) {
      uint32_t r[32], r2[32];
      if (G > 0) {
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + cb * 32;
        tmem_ld32(taddr, r);
        tmem_ld32(taddr + 256, r2);
        tmem_ld_wait();
      }
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const int col = cb * 32 + j;
        if (col < W1) dst[col] = G > 0 ? __uint_as_float(r[j]) + __uint_as_float(r2[j]) : 0.f;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kMmaWarp) tmem_dealloc(tmem_base, kTmemCols);
}

// gW[o, i] = sum_s partial[s][o][i], gb[o] = sum_s partial[s][o][K]   (fixed order)
__global__ void wgrad_reduce_kernel(const float* partial, int splits, int rows_pad, int N, int K, float* gW, float* gb) {
  const int W1 = K + 1;
  const int64_t total = static_cast<int64_t>(N) * W1;
  for (int64_t e = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; e < total;
       e += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int o = static_cast<int>(e / W1), i = static_cast<int>(e - static_cast<int64_t>(o) * W1);
    float acc = 0.f;
    for (int s = 0; s < splits; ++s) acc += partial[(static_cast<int64_t>(s) * rows_pad + o) * W1 + i];
    if (i < K) gW[static_cast<int64_t>(o) * K + i] = acc;
    else if (gb) gb[o] = acc;
  }
}

}  // namespace gemm
}  // namespace vcnf

using namespace vcnf;

extern "C" int vcnf_linear_tf32x3(const void* A, int64_t lda, const void* W, int64_t sw_n, int64_t sw_k,
                                  const void* bias, const void* residual, int64_t ldr, const void* mask,
                                  int64_t ldm, void* C, int64_t ldc, int64_t M, int N, int K, int relu_in,
                                  void* stream) {
  if (M == 0) return 0;   /* empty batch: pointers may be NULL */
  if (!A || !W || !C || M < 0 || N <= 0 || K <= 0 || lda < K || ldc < N) return VCNF_E_BADARG;
  if (K > gemm::kMaxK) return VCNF_E_UNSUPPORTED;
  const int sms = sm_count();
  if (sms <= 0) return static_cast<int>(cudaErrorNoDevice);
  gemm::Args a{};
  a.A = static_cast<const float*>(A); a.lda = lda;
  a.W = static_cast<const float*>(W); a.sw_n = sw_n; a.sw_k = sw_k;
  a.bias = static_cast<const float*>(bias);
  a.residual = static_cast<const float*>(residual); a.ldr = ldr;
  a.mask = static_cast<const float*>(mask); a.ldm = ldm;
  a.C = static_cast<float*>(C); a.ldc = ldc;
  a.M = M; a.N = N; a.K = K; a.relu_in = relu_in;
  { const char* e = getenv("VCNF_GEMM_DEBUG"); a.dbg = e ? atoi(e) : 0; }
  a.n_tiles = (N + gemm::kBN - 1) / gemm::kBN;
  a.KP = (K + gemm::kBK - 1) / gemm::kBK * gemm::kBK;
  const int64_t m_tiles = (M + gemm::kBM - 1) / gemm::kBM;
  int cpt = sms / a.n_tiles;
  if (cpt < 1) cpt = 1;
  if (cpt > m_tiles) cpt = static_cast<int>(m_tiles);
  a.ctas_per_ntile = cpt;
  const size_t smem = 1024 + 2 * static_cast<size_t>(gemm::kBN) * a.KP * 4 +
                      static_cast<size_t>(gemm::kStages) * 2 * gemm::kBM * gemm::kBK * 4 +
                      static_cast<size_t>(gemm::kEpiWarps) * 32 * 36 * 4;
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(gemm::linear_tf32x3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    attr_set = true;
  }
  gemm::linear_tf32x3_kernel<<<static_cast<unsigned>(a.n_tiles * cpt), gemm::kThreads, smem,
                               static_cast<cudaStream_t>(stream)>>>(a);
  return launch_status();
}


static int wgrad_splits(int64_t B, int N) {
  const int sms = sm_count() > 0 ? sm_count() : VCNF_SM_COUNT_DEFAULT;
  const int o_tiles = (N + 127) / 128;
  const int64_t chunks = (B + gemm::kWBK - 1) / gemm::kWBK;
  int64_t s = sms / o_tiles;
  if (s < 1) s = 1;
  if (s > chunks) s = chunks;
  return static_cast<int>(s);
}

extern "C" int64_t vcnf_linear_wgrad_workspace(int64_t B, int N, int K) {
  if (B <= 0 || N <= 0 || K <= 0) return 0;
  const int o_tiles = (N + 127) / 128;
  return static_cast<int64_t>(wgrad_splits(B, N)) * o_tiles * 128 * (K + 1) * 4;
}

extern "C" int vcnf_linear_wgrad_tf32x3(const void* gY, int64_t ldg, const void* X, int64_t ldx, void* gW, void* gb,
                                        void* workspace, int64_t B, int N, int K, int relu_in, void* stream) {
  if (!gW || N <= 0 || K <= 0 || B < 0) return VCNF_E_BADARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (B == 0) {   /* empty batch: gradients are zero */
    cudaMemsetAsync(gW, 0, static_cast<size_t>(N) * K * 4, st);
    if (gb) cudaMemsetAsync(gb, 0, static_cast<size_t>(N) * 4, st);
    return launch_status();
  }
  if (!gY || !X || !workspace || ldg < N || ldx < K) return VCNF_E_BADARG;
  if (K > gemm::kMaxK) return VCNF_E_UNSUPPORTED;
  gemm::WArgs a{};
  a.gY = static_cast<const float*>(gY); a.ldg = ldg;
  a.X = static_cast<const float*>(X); a.ldx = ldx;
  a.partial = static_cast<float*>(workspace);
  a.B = B; a.N = N; a.K = K; a.relu_in = relu_in;
  a.o_tiles = (N + 127) / 128;
  a.splits = wgrad_splits(B, N);
  a.NX = (K + 1 + 15) / 16 * 16;
  a.chunks = (B + gemm::kWBK - 1) / gemm::kWBK;
  const size_t smem = 1024 + static_cast<size_t>(gemm::kWStages) * 2 * (128 / 8 + static_cast<size_t>(a.NX) / 8) * gemm::kWSbo;
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(gemm::linear_wgrad_tf32x3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    attr_set = true;
  }
  gemm::linear_wgrad_tf32x3_kernel<<<static_cast<unsigned>(a.o_tiles * a.splits), gemm::kThreads, smem, st>>>(a);
  int rc = launch_status();
  if (rc) return rc;
  const int64_t total = static_cast<int64_t>(N) * (K + 1);
  int grid = static_cast<int>((total + 255) / 256);
  if (grid > 1184) grid = 1184;
  gemm::wgrad_reduce_kernel<<<grid, 256, 0, st>>>(a.partial, a.splits, a.o_tiles * 128, N, K, static_cast<float*>(gW),
                                                   static_cast<float*>(gb));
  return launch_status();
}
