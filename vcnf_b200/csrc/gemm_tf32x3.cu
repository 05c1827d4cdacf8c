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
#include <cuda.h>
#include "common.cuh"
#include "../../include/vcnf_b200.h"

// ---- programmatic dependent launch (PDL) ----------------------------------------------------------------------
// A training step is ~340 back-to-back GEMM launches of 13-50 us, each with a serial prologue (barrier init, TMEM
// allocation, W -> TMEM: 4.4 us) behind ~2.5 us of launch latency.  The single-wave GEMM kernels signal
// `launch_dependents` as soon as they are set up, and are themselves launched with the programmatic-serialization
// attribute: the next GEMM's CTAs are placed on an SM the moment the previous kernel's CTA leaves it and run their
// prologue (which only touches parameters) while the rest of the previous grid drains; `griddepcontrol.wait` stands
// before the first access to anything a preceding kernel may have produced (activations, gradients) or may still
// read (outputs).  Kernels that do not signal (everything else in the stream) keep the ordinary full dependency.
// VCNF_PDL=0 turns the attribute off.
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

static bool pdl_enabled() {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("VCNF_PDL");
    on = (e && e[0] == '0') ? 0 : 1;
  }
  return on == 1;
}

template <typename... KArgs, typename... Args>
static void launch_pdl(void (*kernel)(KArgs...), unsigned grid, unsigned block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid, 1, 1);
  cfg.blockDim = dim3(block, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

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
// v2: weights resident in TENSOR MEMORY, activations streamed by cp.async.
//
// The contraction is evaluated transposed,  D^T[o, b] = sum_k W[o, k] * act(A[b, k]):
//   UMMA "A" operand = W  (M = 128 output features), read from TMEM (tcgen05.mma [d], [a_tmem], b_desc):
//                      hi plane in columns [0, 128), lo plane in [128, 256), written once per CTA with
//                      tcgen05.st -- no shared memory is spent on the weights;
//   UMMA "B" operand = activations, N = 64 batch rows per tile, K-major in shared memory;
//   accumulators     = 2 buffers x (64 columns hi*hi + 64 columns cross terms) in columns [256, 512).
// What this buys over v1 (W in smem, register-staged A):
//   * all 221 KB of shared memory is the activation ring (12 stages of 64 rows x 32 k, hi + lo planes);
//     producers issue 16 B cp.async copies 8 stages ahead (64 KB in flight per SM, no registers held),
//     then convert their OWN chunks in place (ReLU, round-to-nearest TF32 hi, lo) -- per-thread
//     cp.async.wait_group is the only synchronisation between copy and convert;
//   * each MMA reads only the 2 KB activation operand from shared memory (v1: 8 KB per MMA);
//   * the accumulator comes out feature-major: lane = output feature, register j = batch row j, so one
//     warp instruction stores 32 consecutive floats of one output row -- coalesced without the
//     shared-memory transpose, and the residual row segment is prefetched before the accumulator is
//     even complete.
// =================================================================================================
namespace v2 {
#ifdef VCNF_GEMM_TRACE
__device__ long long g_trace[160][16];
#define VTRACE(i) g_trace[blockIdx.x][i] = clock64()
#else
#define VTRACE(i)
#endif
constexpr int kProdWarps = 4, kEpiWarps = 8;   // (8 producer warps = 544 threads: 96 registers, the epilogue spills -- measured slower)
constexpr int kThreads = (kProdWarps + kEpiWarps + 1) * 32;          // 416
constexpr int kMmaWarp = kProdWarps + kEpiWarps;                     // 12
constexpr int kProd = kProdWarps * 32;
constexpr int kTN = 64;                    // batch rows per tile = UMMA N
constexpr int kCK = 32;                    // fp32 per k-chunk = one ring stage
// Stage plane = 64 rows x 128 B in the SWIZZLE_128B K-major UMMA layout: row r at r * 128, its 16 B chunk c
// at ((c ^ (r & 7)) * 16); 8-row atoms of 1024 B (SBO).  (The no-swizzle "interleave" layout of v1 is
// fetched at ~16 B/clk by the tensor core: measured 155 clk per M128 N64 K8 MMA against a 32 clk floor.)
constexpr uint32_t kSbo = 1024;
constexpr uint32_t kPlane = kTN * 128;                    // one hi (or lo) plane of a stage (8 KB)
constexpr int kStages = 8;
constexpr int kAhead = 6;                  // stages requested ahead of the one being converted
constexpr int kChunksPerThread = (kTN * kCK / 4) / kProd;            // 16 B chunks per producer thread per stage (4)
constexpr uint32_t kColWlo = 128, kColAcc = 256;
constexpr int kEBufs = 3;
constexpr uint32_t kETile = kTN * 128 * 4;                // one 64 x 128 fp32 epilogue tile (32 KB)
constexpr size_t kSmemBytes = 1024 + static_cast<size_t>(kStages) * 2 * kPlane + kEBufs * kETile;

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
// 32 registers of this thread -> 32 consecutive TMEM columns of its lane
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]),
      "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]),
      "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// K-major SWIZZLE_128B operand: LBO unused (1), SBO = 8-row atom stride, layout_type 2 at bits 61-63; the
// k-step inside the 128 B swizzle row is taken by advancing the start address (32 B per K = 8 tf32).
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t saddr, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((saddr >> 4) & 0x3fffu);
  d |= static_cast<uint64_t>(1) << 16;
  d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3fffu) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(2) << 61;
  return d;
}

// One 64-row x 32-float box of A -> shared memory in the 128 B swizzled layout (rows / k beyond the tensor
// read as zeros), completion in bytes on `bar`.
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, int c0, int c1, uint32_t src) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.tile.bulk_group [%0, {%1, %2}], [%3];" ::"l"(
                   reinterpret_cast<uint64_t>(map)),
               "r"(c0), "r"(c1), "r"(src)
               : "memory");
}
__device__ __forceinline__ bool elect_one() {
  uint32_t leader;
  asm volatile("{.reg .b32 rx; .reg .pred px; elect.sync rx|px, 0xffffffff; selp.b32 %0, 1, 0, px;}" : "=r"(leader));
  return leader != 0;
}
// One k-step (K = 8) of the split product: Wlo*Ahi and Whi*Alo into d_small, Whi*Ahi into d_main.
// `acc` = 0 starts both accumulators (first k-step of a tile).
#define VCNF_MMA_TS "tcgen05.mma.cta_group::1.kind::tf32 "
__device__ __forceinline__ void mma_step(uint32_t d_main, uint32_t d_small, uint32_t w_hi, uint32_t w_lo, uint64_t b_hi,
                                         uint64_t b_lo, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t"
      ".reg .pred pa, pt;\n\t"
      "setp.ne.b32 pa, %7, 0;\n\t"
      "setp.eq.b32 pt, %6, %6;\n\t"
      VCNF_MMA_TS "[%1], [%3], %4, %6, pa;\n\t"
      VCNF_MMA_TS "[%1], [%2], %5, %6, pt;\n\t"
      VCNF_MMA_TS "[%0], [%2], %4, %6, pa;\n\t"
      "}\n" ::"r"(d_main), "r"(d_small), "r"(w_hi), "r"(w_lo), "l"(b_hi), "l"(b_lo), "r"(idesc), "r"(acc)
      : "memory");
}
// A full 32-wide k-chunk: 4 k-steps x 3 products = 12 MMAs issued back to back (descriptor start address
// advances 32 B = 2 units, the W columns by 8 per k-step).
__device__ __forceinline__ void mma_stage4(uint32_t d_main, uint32_t d_small, uint32_t w_hi, uint32_t w_lo, uint64_t b_hi,
                                           uint64_t b_lo, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t"
      ".reg .pred pa, pt;\n\t"
      ".reg .b32 wh1, wl1, wh2, wl2, wh3, wl3;\n\t"
      ".reg .b64 bh1, bl1, bh2, bl2, bh3, bl3;\n\t"
      "setp.ne.b32 pa, %7, 0;\n\t"
      "setp.eq.b32 pt, %6, %6;\n\t"
      "add.u32 wh1, %2, 8;\n\t add.u32 wl1, %3, 8;\n\t add.u64 bh1, %4, 2;\n\t add.u64 bl1, %5, 2;\n\t"
      "add.u32 wh2, %2, 16;\n\t add.u32 wl2, %3, 16;\n\t add.u64 bh2, %4, 4;\n\t add.u64 bl2, %5, 4;\n\t"
      "add.u32 wh3, %2, 24;\n\t add.u32 wl3, %3, 24;\n\t add.u64 bh3, %4, 6;\n\t add.u64 bl3, %5, 6;\n\t"
      VCNF_MMA_TS "[%1], [%3], %4, %6, pa;\n\t"
      VCNF_MMA_TS "[%1], [%2], %5, %6, pt;\n\t"
      VCNF_MMA_TS "[%0], [%2], %4, %6, pa;\n\t"
      VCNF_MMA_TS "[%1], [wl1], bh1, %6, pt;\n\t"
      VCNF_MMA_TS "[%1], [wh1], bl1, %6, pt;\n\t"
      VCNF_MMA_TS "[%0], [wh1], bh1, %6, pt;\n\t"
      VCNF_MMA_TS "[%1], [wl2], bh2, %6, pt;\n\t"
      VCNF_MMA_TS "[%1], [wh2], bl2, %6, pt;\n\t"
      VCNF_MMA_TS "[%0], [wh2], bh2, %6, pt;\n\t"
      VCNF_MMA_TS "[%1], [wl3], bh3, %6, pt;\n\t"
      VCNF_MMA_TS "[%1], [wh3], bl3, %6, pt;\n\t"
      VCNF_MMA_TS "[%0], [wh3], bh3, %6, pt;\n\t"
      "}\n" ::"r"(d_main), "r"(d_small), "r"(w_hi), "r"(w_lo), "l"(b_hi), "l"(b_lo), "r"(idesc), "r"(acc)
      : "memory");
}

struct Args2 {
  const float* A; int64_t lda;
  const float* W; int64_t sw_n, sw_k;
  const float* bias;
  const float* residual; int64_t ldr;
  const float* mask; int64_t ldm;
  float* C; int64_t ldc;
  int64_t M; int N; int K;
  int relu_in;
  int w_dynamic;      // W was written by a preceding kernel of this stream (MADE's W * mask): wait before reading it
  int n_tiles;        // ceil(N / 128) column slices
  int ctas_per_ntile; // persistent CTAs sharing one column slice
  int use_tma;        // A is 16 B aligned with a 16 B multiple row pitch: stages arrive by tensor-map TMA
  int e_mode;         // elementwise operand staged by TMA: 0 none, 1 residual, 2 mask
  int mask_last;      // C = [mask > 0] (A W^T + bias + residual) instead of [mask > 0] (A W^T + bias) + residual
  int c_tma;          // output leaves by TMA tensor stores
};

__global__ void __launch_bounds__(kThreads, 1) linear_ts_kernel(const Args2 a, const __grid_constant__ CUtensorMap tm_a,
                                                                 const __grid_constant__ CUtensorMap tm_e,
                                                                 const __grid_constant__ CUtensorMap tm_c) {
  extern __shared__ __align__(1024) unsigned char smem[];
  uint64_t* ready = reinterpret_cast<uint64_t*>(smem);           // [kStages] producers -> MMA (converted)
  uint64_t* empty = ready + kStages;                             // [kStages] MMA -> producers
  uint64_t* acc_full = empty + kStages;                          // [2] MMA -> epilogue
  uint64_t* acc_empty = acc_full + 2;                            // [2] epilogue -> MMA
  uint64_t* w_ready = acc_empty + 2;                             // [1] W planes written to TMEM
  uint64_t* full = w_ready + 1;                                  // [kStages] TMA -> producers (raw fp32 landed)
  uint64_t* e_full = full + kStages;                             // [kEBufs] epilogue tile ready (E landed / buffer free)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(e_full + kEBufs);
  unsigned char* ring = smem + 1024;                             // stage s: [hi plane | lo plane]
  unsigned char* ebuf = ring + kStages * 2 * kPlane;             // kEBufs epilogue tiles

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int nt = blockIdx.x % a.n_tiles;
  const int n0 = nt * 128;
  const int NT = (a.N - n0) < 128 ? (a.N - n0) : 128;            // valid output features of this slice
  const int64_t m_tiles = (a.M + kTN - 1) / kTN;
  const int64_t mt0 = blockIdx.x / a.n_tiles;
  const int nk = (a.K + kCK - 1) / kCK;                          // ring stages per tile
  const int nk8 = (a.K + 7) >> 3;                                // 8-wide MMA steps that carry data
  const int64_t my_tiles = mt0 < m_tiles ? (m_tiles - mt0 + a.ctas_per_ntile - 1) / a.ctas_per_ntile : 0;

  if (tid == 0) {
    VTRACE(0);
    for (int s = 0; s < kStages; ++s) { mbar_init(&ready[s], kProdWarps); mbar_init(&empty[s], 1); mbar_init(&full[s], 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(&acc_full[b], 1); mbar_init(&acc_empty[b], kEpiWarps); }
    mbar_init(w_ready, kEpiWarps);
    for (int b = 0; b < kEBufs; ++b) mbar_init(&e_full[b], 1);
    mbar_init_fence();
  }
  if (warp == kMmaWarp) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (tid == 0) VTRACE(1);
  pdl_trigger();                                                 // the next launch may start its prologue

  if (warp < kProdWarps) {
    // ===================================== producers =====================================
    pdl_wait();                                                  // activations come from preceding kernels
    // Stage g of this CTA's stream = (tile g / nk, k-chunk g % nk): 64 rows x 32 fp32.  Thread item i:
    // 16 B chunk c = tid & 7 of row r = (tid >> 3) + 16 i -- 8 consecutive lanes read one 128 B line,
    // and land in 8 different 16 B bank groups (128 B swizzle).
    const int64_t G = my_tiles * nk;
    if (a.use_tma) {
      // The raw fp32 box lands by TMA in the swizzled layout (= the hi operand once rounded); the 128
      // producer threads then make one elementwise pass over the 8 KB plane: ReLU, round-to-nearest TF32
      // hi in place, lo into the second plane.  Layout-agnostic, so a linear 16 B / thread sweep.
      auto issue_tma = [&](int64_t g) {                          // thread 0 only
        if (g >= G) return;
        const uint32_t it = static_cast<uint32_t>(g);
        const int s = it % kStages;
        mbar_wait(&empty[s], ((it / kStages) & 1u) ^ 1u);
        mbar_expect_tx(&full[s], kPlane);
        const int64_t m0 = (mt0 + (g / nk) * a.ctas_per_ntile) * kTN;
        tma_load_2d(smem_u32(ring) + static_cast<uint32_t>(s) * 2 * kPlane, &tm_a, static_cast<int>(g % nk) * kCK,
                    static_cast<int>(m0), &full[s]);
      };
      if (tid == 0) {
        for (int g = 0; g < kAhead; ++g) issue_tma(g);
        VTRACE(11);
      }
      // Two stages per pass: the fixed cost of a pass (barrier polls, the async-proxy fence, the arrives and the
      // look-ahead TMA issue) is paid once per 16 KB instead of once per 8 KB, and the 8 independent 16 B chunks of
      // a thread cover the shared-memory latency that 4 did not (one producer warp per scheduler: nothing else
      // hides it).  tools/gemm_exp.cu: the head layer lost 10 us of 53 to this loop.
      for (int64_t g = 0; g < G; g += 2) {
        const uint32_t it0 = static_cast<uint32_t>(g), it1 = it0 + 1;
        const int s0 = it0 % kStages, s1 = it1 % kStages;
        const bool two = g + 1 < G;
        mbar_wait(&full[s0], (it0 / kStages) & 1u);
        if (two) mbar_wait(&full[s1], (it1 / kStages) & 1u);
        unsigned char* p0 = ring + static_cast<uint32_t>(s0) * 2 * kPlane + tid * 16;
        unsigned char* p1 = ring + static_cast<uint32_t>(s1) * 2 * kPlane + tid * 16;
#ifndef VCNF_EXP_SKIP_CONVERT                                    // tools/gemm_exp.cu: timing experiments only
        float4 v[2 * kChunksPerThread];
#pragma unroll
        for (int i = 0; i < kChunksPerThread; ++i) v[i] = *reinterpret_cast<const float4*>(p0 + i * (kProd * 16));
        if (two) {
#pragma unroll
          for (int i = 0; i < kChunksPerThread; ++i)
            v[kChunksPerThread + i] = *reinterpret_cast<const float4*>(p1 + i * (kProd * 16));
        }
#pragma unroll
        for (int i = 0; i < 2 * kChunksPerThread; ++i) {
          if (i >= kChunksPerThread && !two) break;
          unsigned char* dst = (i < kChunksPerThread ? p0 : p1) + (i % kChunksPerThread) * (kProd * 16);
          float4 x = v[i];
          if (a.relu_in) { x.x = fmaxf(x.x, 0.f); x.y = fmaxf(x.y, 0.f); x.z = fmaxf(x.z, 0.f); x.w = fmaxf(x.w, 0.f); }
          float4 h, l;
          split_tf32(x.x, h.x, l.x); split_tf32(x.y, h.y, l.y);
          split_tf32(x.z, h.z, l.z); split_tf32(x.w, h.w, l.w);
          *reinterpret_cast<float4*>(dst) = h;
          *reinterpret_cast<float4*>(dst + kPlane) = l;
        }
#endif
        fence_async_smem();
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(&ready[s0]);
          if (two) mbar_arrive(&ready[s1]);
        }
        if (tid == 0) {
          if (g == 0) { VTRACE(3); }
          issue_tma(g + kAhead);
          issue_tma(g + 1 + kAhead);
        }
      }
      if (tid == 0) { VTRACE(7); }
    } else {
    const bool vec_ok = (a.lda % 4 == 0) && (a.K % 4 == 0) && (reinterpret_cast<uintptr_t>(a.A) % 16 == 0);
    const int c = tid & 7, r0 = tid >> 3;
    const uint32_t ring_u32 = smem_u32(ring);
    uint32_t off[kChunksPerThread];
#pragma unroll
    for (int i = 0; i < kChunksPerThread; ++i) {
      const int r = r0 + (kProd / 8) * i;
      off[i] = static_cast<uint32_t>(r) * 128 + static_cast<uint32_t>((c ^ (r & 7)) << 4);
    }
    auto issue = [&](int64_t g) {
      if (g < G) {
        const uint32_t it = static_cast<uint32_t>(g);
        const int s = it % kStages;
        mbar_wait(&empty[s], ((it / kStages) & 1u) ^ 1u);
        const int64_t m0 = (mt0 + (g / nk) * a.ctas_per_ntile) * kTN;
        const int gk = static_cast<int>(g % nk) * kCK + c * 4;
        const uint32_t base = static_cast<uint32_t>(s) * 2 * kPlane;
#pragma unroll
        for (int i = 0; i < kChunksPerThread; ++i) {
          const int64_t gr = m0 + r0 + (kProd / 8) * i;
          const bool ok = gr < a.M && gk < a.K;
          if (vec_ok) {
            cp_async16(ring_u32 + base + off[i], ok ? a.A + gr * a.lda + gk : a.A, ok ? 16u : 0u);
          } else {
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (ok) {
              const float* src = a.A + gr * a.lda + gk;
              v.x = __ldg(src);
              if (gk + 1 < a.K) v.y = __ldg(src + 1);
              if (gk + 2 < a.K) v.z = __ldg(src + 2);
              if (gk + 3 < a.K) v.w = __ldg(src + 3);
            }
            *reinterpret_cast<float4*>(ring + base + off[i]) = v;
          }
        }
      }
      cp_async_commit();                                         // one group per stage, empty past the end
    };
    for (int g = 0; g < kAhead; ++g) issue(g);
    for (int64_t g = 0; g < G; ++g) {
      cp_async_wait<kAhead - 1>();                               // this thread's copies of stage g have landed
      const int s = static_cast<uint32_t>(g) % kStages;
      unsigned char* hi_p = ring + static_cast<uint32_t>(s) * 2 * kPlane;
#pragma unroll
      for (int i = 0; i < kChunksPerThread; ++i) {
        float4 v = *reinterpret_cast<const float4*>(hi_p + off[i]);
        if (a.relu_in) { v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f); }
        float4 h, l;
        split_tf32(v.x, h.x, l.x); split_tf32(v.y, h.y, l.y);
        split_tf32(v.z, h.z, l.z); split_tf32(v.w, h.w, l.w);
        *reinterpret_cast<float4*>(hi_p + off[i]) = h;
        *reinterpret_cast<float4*>(hi_p + kPlane + off[i]) = l;
      }
      fence_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(&ready[s]);
      if (tid == 0 && g == 0) VTRACE(3);
      issue(g + kAhead);
    }
    cp_async_wait<0>();
    if (tid == 0) { VTRACE(7); }
    }
  } else if (warp == kMmaWarp) {
    // ===================================== MMA issuer =====================================
    const uint32_t idesc = instr_desc_tf32(128, kTN);
    mbar_wait(w_ready, 0);
    tc_fence_after();
    const uint32_t ring_u32 = smem_u32(ring);
    uint32_t it = 0;
    for (int64_t tc = 0; tc < my_tiles; ++tc) {
      const uint32_t buf = static_cast<uint32_t>(tc) & 1u;
      mbar_wait(&acc_empty[buf], ((static_cast<uint32_t>(tc) >> 1) & 1u) ^ 1u);
      tc_fence_after();
      const uint32_t d_main = tmem_base + kColAcc + buf * (2 * kTN);
      const uint32_t d_small = d_main + kTN;
      for (int kc = 0; kc < nk; ++kc, ++it) {
        const int s = it % kStages;
        mbar_wait(&ready[s], (it / kStages) & 1u);
        tc_fence_after();
        if (lane == 0 && it == 0) { VTRACE(4); }
        // Everything up to the issue is computed by the WHOLE warp in convergent code, and the MMAs of a
        // k-step go out back to back from one asm block: operands computed inside a divergent `if (lane == 0)`
        // make the compiler wrap every UTCHMMA in an ELECT / R2UR.BROADCAST waterfall loop (measured 230
        // clk per MMA against the 32 clk this shape needs, tools/mma_rate.cu).
        const uint32_t w_hi_col = tmem_base + static_cast<uint32_t>(kc) * kCK;
        const uint32_t w_lo_col = w_hi_col + kColWlo;
        const uint64_t bd_hi = smem_desc_sw128(ring_u32 + static_cast<uint32_t>(s) * 2 * kPlane, kSbo);
        const uint64_t bd_lo = smem_desc_sw128(ring_u32 + static_cast<uint32_t>(s) * 2 * kPlane + kPlane, kSbo);
        const int steps = (nk8 - kc * 4) < 4 ? (nk8 - kc * 4) : 4;
        const uint32_t acc0 = kc > 0 ? 1u : 0u;
        const bool last = kc == nk - 1;
        if (elect_one()) {
#ifndef VCNF_EXP_SKIP_MMA
          if (steps == 4) {
            mma_stage4(d_main, d_small, w_hi_col, w_lo_col, bd_hi, bd_lo, idesc, acc0);
          } else {
            for (int kk = 0; kk < steps; ++kk)
              mma_step(d_main, d_small, w_hi_col + kk * 8, w_lo_col + kk * 8, bd_hi + kk * 2, bd_lo + kk * 2, idesc,
                       (kc > 0 || kk > 0) ? 1u : 0u);
          }
#endif
          tc_commit(&empty[s]);
          if (last) tc_commit(&acc_full[buf]);
        }
        __syncwarp();
      }
    }
    if (lane == 0) VTRACE(8);
  } else {
    // ===================================== epilogue (8 warps) =====================================
    // warp = (TMEM lane quarter q = warp % 4 -> output features n0 + 32 q + lane, half h -> batch rows
    // 32 h .. 32 h + 31 of the tile).  First job: put this warp's share of W into tensor memory.
    const int q = warp & 3, h = (warp - kProdWarps) >> 2;
    const int o_loc = q * 32 + lane;
    const bool o_ok = o_loc < NT;
    const int o = n0 + o_loc;
    const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    // W and bias are normally parameters (no kernel of this chain writes them), so they are fetched BEFORE the
    // dependency wait and overlap the previous launch's tail.  A weight produced by the preceding kernel (the masked
    // weight W * mask of nets/made.py) is only guaranteed visible after griddepcontrol.wait.
    if (a.w_dynamic) pdl_wait();
    {
      const float* wrow = a.W + static_cast<int64_t>(o_ok ? o : n0) * a.sw_n;
      // 32 x 32 transpose tile per warp, xor-swizzled; aliases epilogue tile 2, which is first requested
      // after every epilogue warp has passed the named barrier of tile 0
      float* tb = reinterpret_cast<float*>(ebuf + 2 * kETile) + (warp - kProdWarps) * (32 * 32);
      // Both 32-column chunks of this warp are requested before the first one is used (one memory latency
      // instead of two on the critical path to the first MMA).
      float wv[2][32];
#pragma unroll
      for (int cc = 0; cc < 2; ++cc) {
        const int k0 = h * 64 + cc * 32;
        if (k0 >= nk8 * 8) continue;                              // warp-uniform: columns no MMA reads
        if (a.sw_k == 1) {
          // row-major W: a warp instruction reads 128 contiguous bytes of ONE row (lane = k); the 32 x 32
          // block is transposed through a smem tile below so that lane = output feature afterwards.  (Per-lane
          // row reads are 32 sectors per request; with all 148 CTAs hitting the same 64 KB at once they took
          // 10 us of every launch.)
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const int oi = q * 32 + i, k = k0 + lane;
            wv[cc][i] = (oi < NT && k < a.K) ? __ldg(a.W + static_cast<int64_t>(n0 + oi) * a.sw_n + k) : 0.f;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const int k = k0 + j;
            wv[cc][j] = (o_ok && k < a.K) ? __ldg(wrow + static_cast<int64_t>(k) * a.sw_k) : 0.f;
          }
        }
      }
#pragma unroll
      for (int cc = 0; cc < 2; ++cc) {
        const int k0 = h * 64 + cc * 32;
        if (k0 >= nk8 * 8) continue;
        uint32_t whi[32], wlo[32];
        if (a.sw_k == 1) {
          __syncwarp();
#pragma unroll
          for (int i = 0; i < 32; ++i) tb[i * 32 + (lane ^ i)] = wv[cc][i];
          __syncwarp();
#pragma unroll
          for (int j = 0; j < 32; ++j) wv[cc][j] = tb[lane * 32 + (j ^ lane)];
        }
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          float hi, lo;
          split_tf32(wv[cc][j], hi, lo);
          whi[j] = __float_as_uint(hi);
          wlo[j] = __float_as_uint(lo);
        }
        tmem_st32(lane_base + k0, whi);
        tmem_st32(lane_base + kColWlo + k0, wlo);
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(w_ready);
      if (warp == kProdWarps && lane == 0) VTRACE(2);
    }
    pdl_wait();                                                  // W, bias: parameters; everything below: activations
    const float bv = (a.bias && o_ok) ? __ldg(a.bias + o) : 0.f;
    // Elementwise operand E (the residual, or the ReLU mask of the input-gradient call) and the output
    // tile travel by TMA through three 64 x 128 fp32 tiles in shared memory: E for tile t+2 is requested
    // while tile t is finished, the result overwrites E in place and leaves with one bulk tensor store.
    // In-flight bytes live in shared memory, not in registers of 8 warps (the register version topped
    // out at ~20 KB/us per SM: one memory latency per tile).  Operands whose pitch / alignment rules out
    // a tensor map fall back to direct, still row-coalesced, accesses.
    const bool staged = a.e_mode != 0 || a.c_tma != 0;
    const bool elead = warp == kProdWarps && lane == 0;
    const float* res_reg = (a.residual && a.e_mode != 1) ? a.residual : nullptr;
    const float* mask_reg = (a.mask && a.e_mode != 2) ? a.mask : nullptr;
    auto request_e = [&](int64_t t) {                              // elead only: arm e_full[t % 3]
      if (t >= my_tiles) return;
      uint64_t* bar = &e_full[t % kEBufs];
      if (a.e_mode != 0) {
        mbar_expect_tx(bar, kETile);
        tma_load_2d(smem_u32(ebuf) + static_cast<uint32_t>(t % kEBufs) * kETile, &tm_e, n0,
                    static_cast<int>((mt0 + t * a.ctas_per_ntile) * kTN), bar);
      } else {
        mbar_arrive(bar);                                          // buffer free, nothing to load
      }
    };
    if (staged && elead) { request_e(0); request_e(1); }
    for (int64_t tc = 0; tc < my_tiles; ++tc) {
      const uint32_t buf = static_cast<uint32_t>(tc) & 1u;
      const int64_t row0 = (mt0 + tc * a.ctas_per_ntile) * kTN + h * 32;
      int rows = static_cast<int>(a.M - row0 < 32 ? a.M - row0 : 32);
      if (rows < 0 || !o_ok) rows = 0;
      float* et = reinterpret_cast<float*>(ebuf + static_cast<uint32_t>(tc % kEBufs) * kETile) + (h * 32) * 128 + o_loc;
      float res[32];
      if (res_reg) {                                              // independent of the MMAs: fetch early
#pragma unroll
        for (int j = 0; j < 32; ++j) res[j] = (j < rows) ? __ldg(res_reg + (row0 + j) * a.ldr + o) : 0.f;
      }
      mbar_wait(&acc_full[buf], (static_cast<uint32_t>(tc) >> 1) & 1u);
      tc_fence_after();
      if (warp == kProdWarps && lane == 0 && tc == 0) VTRACE(5);
      uint32_t r[32], r2[32];
      const uint32_t taddr = lane_base + kColAcc + buf * (2 * kTN) + h * 32;
#ifndef VCNF_EXP_SKIP_EPI
      tmem_ld32(taddr, r);
      tmem_ld32(taddr + kTN, r2);
      tmem_ld_wait();
#else
#pragma unroll
      for (int j = 0; j < 32; ++j) { r[j] = 0u; r2[j] = 0u; }
#endif
      tc_fence_before();                                          // values are in registers: hand the
      __syncwarp();                                               // accumulator back before the stores
      if (lane == 0) mbar_arrive(&acc_empty[buf]);
      float acc[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) acc[j] = __uint_as_float(r[j]) + __uint_as_float(r2[j]) + bv;
      if (staged) mbar_wait(&e_full[tc % kEBufs], static_cast<uint32_t>(tc / kEBufs) & 1u);
      if (a.mask_last) {
        // last slice of a sliced contraction: the running sum (residual == C of the earlier slices) comes first, the
        // ReLU mask applies to the total
        if (a.e_mode == 1) {
#pragma unroll
          for (int j = 0; j < 32; ++j) acc[j] += et[j * 128];
        }
        if (res_reg) {
#pragma unroll
          for (int j = 0; j < 32; ++j) acc[j] += res[j];
        }
      }
      if (a.e_mode == 2) {
#pragma unroll
        for (int j = 0; j < 32; ++j) acc[j] = et[j * 128] > 0.f ? acc[j] : 0.f;
      }
      if (mask_reg) {
        float mk[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) mk[j] = (j < rows) ? __ldg(mask_reg + (row0 + j) * a.ldm + o) : 1.f;
#pragma unroll
        for (int j = 0; j < 32; ++j) acc[j] = mk[j] > 0.f ? acc[j] : 0.f;
      }
      if (!a.mask_last) {
        if (a.e_mode == 1) {
#pragma unroll
          for (int j = 0; j < 32; ++j) acc[j] += et[j * 128];
        }
        if (res_reg) {
#pragma unroll
          for (int j = 0; j < 32; ++j) acc[j] += res[j];
        }
      }
      if (a.c_tma) {
#ifndef VCNF_EXP_SKIP_EPI
#pragma unroll
        for (int j = 0; j < 32; ++j) et[j * 128] = acc[j];        // lane = column: conflict-free rows
#endif
        fence_async_smem();
      } else {
        float* cp = a.C + row0 * a.ldc + o;
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (j < rows) cp[j * a.ldc] = acc[j];                   // one 128 B row segment per warp instruction
      }
      if (staged) {
        asm volatile("bar.sync 1, %0;" ::"n"(kEpiWarps * 32) : "memory");   // tile complete / E consumed by all 8 warps
        if (elead) {
          if (a.c_tma) {
            tma_store_2d(&tm_c, n0, static_cast<int>((mt0 + tc * a.ctas_per_ntile) * kTN),
                         smem_u32(ebuf) + static_cast<uint32_t>(tc % kEBufs) * kETile);
            bulk_commit();
            asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");   // store of tile tc-1 has left its buffer
          }
          request_e(tc + 2);                                      // (tc + 2) % 3 == (tc - 1) % 3
        }
      }
      if (warp == kProdWarps && lane == 0 && tc == 0) VTRACE(6);
    }
    if (elead && a.c_tma) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    if (warp == kProdWarps && lane == 0) VTRACE(9);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == kMmaWarp) tmem_dealloc(tmem_base, 512);
  if (tid == 0) VTRACE(10);
}
}  // namespace v2

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

// One 32-sample stage of the split product with BOTH operands in shared memory (no-swizzle K-major, 256 B per
// k-step): 4 k-steps x {lo*hi, hi*lo -> d_small; hi*hi -> d_main}, issued back to back by the elected lane of a
// CONVERGENT warp (see v2::mma_stage4 for why).
__device__ __forceinline__ void mma_stage4_ss(uint32_t d_main, uint32_t d_small, uint64_t a_hi, uint64_t a_lo, uint64_t b_hi,
                                              uint64_t b_lo, uint32_t idesc, uint32_t acc) {
#define VCNF_MMA_SS "tcgen05.mma.cta_group::1.kind::tf32 "
  asm volatile(
      "{\n\t"
      ".reg .pred pa, pt;\n\t"
      ".reg .b64 ah1, al1, bh1, bl1, ah2, al2, bh2, bl2, ah3, al3, bh3, bl3;\n\t"
      "setp.ne.b32 pa, %7, 0;\n\t"
      "setp.eq.b32 pt, %6, %6;\n\t"
      "add.u64 ah1, %2, 16;\n\t add.u64 al1, %3, 16;\n\t add.u64 bh1, %4, 16;\n\t add.u64 bl1, %5, 16;\n\t"
      "add.u64 ah2, %2, 32;\n\t add.u64 al2, %3, 32;\n\t add.u64 bh2, %4, 32;\n\t add.u64 bl2, %5, 32;\n\t"
      "add.u64 ah3, %2, 48;\n\t add.u64 al3, %3, 48;\n\t add.u64 bh3, %4, 48;\n\t add.u64 bl3, %5, 48;\n\t"
      VCNF_MMA_SS "[%1], %3, %4, %6, pa;\n\t"
      VCNF_MMA_SS "[%1], %2, %5, %6, pt;\n\t"
      VCNF_MMA_SS "[%0], %2, %4, %6, pa;\n\t"
      VCNF_MMA_SS "[%1], al1, bh1, %6, pt;\n\t"
      VCNF_MMA_SS "[%1], ah1, bl1, %6, pt;\n\t"
      VCNF_MMA_SS "[%0], ah1, bh1, %6, pt;\n\t"
      VCNF_MMA_SS "[%1], al2, bh2, %6, pt;\n\t"
      VCNF_MMA_SS "[%1], ah2, bl2, %6, pt;\n\t"
      VCNF_MMA_SS "[%0], ah2, bh2, %6, pt;\n\t"
      VCNF_MMA_SS "[%1], al3, bh3, %6, pt;\n\t"
      VCNF_MMA_SS "[%1], ah3, bl3, %6, pt;\n\t"
      VCNF_MMA_SS "[%0], ah3, bh3, %6, pt;\n\t"
      "}\n" ::"r"(d_main), "r"(d_small), "l"(a_hi), "l"(a_lo), "l"(b_hi), "l"(b_lo), "r"(idesc), "r"(acc)
      : "memory");
}

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
      const uint32_t a_hi = smem_u32(st0 + static_cast<uint32_t>(s) * stage_bytes);
      const uint32_t a_lo = a_hi + a_half;
      const uint32_t b_hi = a_hi + 2 * a_half;
      const uint32_t b_lo = b_hi + b_half;
      const uint64_t ad_hi = smem_desc(a_hi, 128, kWSbo), ad_lo = smem_desc(a_lo, 128, kWSbo);
      const uint64_t bd_hi = smem_desc(b_hi, 128, kWSbo), bd_lo = smem_desc(b_lo, 128, kWSbo);
      const uint32_t acc0 = i > 0 ? 1u : 0u;
      const bool last = i == G - 1;
      if (v2::elect_one()) {
        mma_stage4_ss(d_main, d_small, ad_hi, ad_lo, bd_hi, bd_lo, idesc, acc0);
        tc_commit(&empty[s]);
        if (last) tc_commit(acc_full);
      }
      __syncwarp();
    }
  } else {
    // ------------------------------ epilogue: one partial [128, K + 1] per CTA
    // TMEM hands every lane one ROW of the partial; each 32 x 32 block is transposed through a padded tile
    // in the (now idle) stage memory so that a warp instruction writes 128 contiguous bytes of one row
    // (lane = row stores were 32 sectors per request: 78 MB of L2 write traffic for a 9.8 MB partial).
    const int q = warp & 3;
    const int W1 = a.K + 1;
    float* dst = a.partial + (static_cast<int64_t>(split) * a.o_tiles * 128 + o0 + q * 32) * W1;
    float* tb = reinterpret_cast<float*>(st0) + (warp - kLoaderWarps) * (32 * 33);
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
      __syncwarp();
#pragma unroll
      for (int j = 0; j < 32; ++j) tb[lane * 33 + j] = G > 0 ? __uint_as_float(r[j]) + __uint_as_float(r2[j]) : 0.f;
      __syncwarp();
      const int col = cb * 32 + lane;
      if (col < W1) {
#pragma unroll 8
        for (int i = 0; i < 32; ++i) dst[static_cast<int64_t>(i) * W1 + col] = tb[i * 33 + lane];
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kMmaWarp) tmem_dealloc(tmem_base, kTmemCols);
}


// -------------------------------------------------------------------------------------------------
// Weight gradient, TMA-fed (used whenever both operands have 16-byte-multiple row pitches).
// The contraction runs over samples, and both operands are stored sample-major with the FEATURE index
// contiguous -- "MN-major" for the tensor core.  For 32-bit operands UMMA reads MN-major tiles in the
// SWIZZLE_128B_BASE32B layout (32 features = 128 B per sample row, 32 B chunks xor-ed with (row & 3),
// 4-row atoms of 512 B), which is exactly what a TMA box {32 features, 32 samples} with
// CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B writes (probed in tools/mn_probe.cu; plain SWIZZLE_128B and the
// no-swizzle layout return zeros for tf32).  So nothing is transposed any more: raw tiles land by TMA,
// 8 warps make one elementwise pass (ReLU, round-to-nearest hi in place, lo plane, ones column), and the
// MMAs read them as they lie.
// -------------------------------------------------------------------------------------------------
namespace wg2 {
#ifdef VCNF_GEMM_TRACE
__device__ long long g_wtrace[160][16];
#define WTRACE(i) g_wtrace[blockIdx.x][i] = clock64()
#else
#define WTRACE(i)
#endif
constexpr int kR = 32;                         // samples per stage
constexpr uint32_t kGroup = kR * 128;          // one 32-feature group of a stage (4 KB)
constexpr int kConvWarps = 8;                  // X converters (warps 0-7); warps 8-11 move gY into tensor memory
constexpr int kStages = 4;                     // raw tiles in flight: [gY raw 16 KB | X raw -> hi, gx groups]
constexpr int kLoBufs = 2;                     // X lo planes live only from the conversion to the MMAs that read them
constexpr uint32_t kColSmall = 128, kColA = 256;   // TMEM: d_main [0,128) d_small [128,256) gY stage s at 256 + 64 s: hi | lo

struct Args {
  float* partial;                              // [splits, o_tiles * 128, K + 1]
  int64_t B; int N; int K;
  int relu_in;
  int o_tiles, splits;
  int gx;                                      // 32-feature groups of X: ceil(K / 32)  (UMMA N = 32 gx)
  int64_t chunks;                              // ceil(B / kR)
};

__device__ __forceinline__ uint64_t desc_mn(uint32_t saddr) {     // MN-major, 128B_BASE32B: LBO = group, SBO = 4-row atom
  uint64_t d = 0;
  d |= static_cast<uint64_t>((saddr >> 4) & 0x3fffu);
  d |= static_cast<uint64_t>((kGroup >> 4) & 0x3fffu) << 16;
  d |= static_cast<uint64_t>((512u >> 4) & 0x3fffu) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(1) << 61;
  return d;
}
// 4 k-steps (8 samples: 8 TMEM columns of the A operand, 1024 B of the B operand) x
// {Alo*Bhi, Ahi*Blo -> d_small; Ahi*Bhi -> d_main}; A = gY^T from tensor memory, B = X MN-major from smem.
__device__ __forceinline__ void mma_stage4_ts_mn(uint32_t d_main, uint32_t d_small, uint32_t a_hi, uint32_t a_lo, uint64_t b_hi,
                                                 uint64_t b_lo, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t"
      ".reg .pred pa, pt;\n\t"
      ".reg .b32 ah1, al1, ah2, al2, ah3, al3;\n\t"
      ".reg .b64 bh1, bl1, bh2, bl2, bh3, bl3;\n\t"
      "setp.ne.b32 pa, %7, 0;\n\t"
      "setp.eq.b32 pt, %6, %6;\n\t"
      "add.u32 ah1, %2, 8;\n\t add.u32 al1, %3, 8;\n\t add.u64 bh1, %4, 64;\n\t add.u64 bl1, %5, 64;\n\t"
      "add.u32 ah2, %2, 16;\n\t add.u32 al2, %3, 16;\n\t add.u64 bh2, %4, 128;\n\t add.u64 bl2, %5, 128;\n\t"
      "add.u32 ah3, %2, 24;\n\t add.u32 al3, %3, 24;\n\t add.u64 bh3, %4, 192;\n\t add.u64 bl3, %5, 192;\n\t"
      VCNF_MMA_SS "[%1], [%3], %4, %6, pa;\n\t"
      VCNF_MMA_SS "[%1], [%2], %5, %6, pt;\n\t"
      VCNF_MMA_SS "[%0], [%2], %4, %6, pa;\n\t"
      VCNF_MMA_SS "[%1], [al1], bh1, %6, pt;\n\t"
      VCNF_MMA_SS "[%1], [ah1], bl1, %6, pt;\n\t"
      VCNF_MMA_SS "[%0], [ah1], bh1, %6, pt;\n\t"
      VCNF_MMA_SS "[%1], [al2], bh2, %6, pt;\n\t"
      VCNF_MMA_SS "[%1], [ah2], bl2, %6, pt;\n\t"
      VCNF_MMA_SS "[%0], [ah2], bh2, %6, pt;\n\t"
      VCNF_MMA_SS "[%1], [al3], bh3, %6, pt;\n\t"
      VCNF_MMA_SS "[%1], [ah3], bl3, %6, pt;\n\t"
      VCNF_MMA_SS "[%0], [ah3], bh3, %6, pt;\n\t"
      "}\n" ::"r"(d_main), "r"(d_small), "r"(a_hi), "r"(a_lo), "l"(b_hi), "l"(b_lo), "r"(idesc), "r"(acc)
      : "memory");
}

// Shared-memory bandwidth, not HBM, bounded the first TMA-fed version (236 KB per 32-sample stage: both
// operands written as hi + lo planes and read by 12 MMAs).  Here gY goes to TENSOR MEMORY: the 128 threads
// of warps 8-11 (thread = output feature = TMEM lane) read the raw tile column-wise (conflict-free in the
// 32 B-atom swizzle), split in registers and tcgen05.st the hi / lo rows, summing the bias gradient on the
// way -- so there is no ones column, UMMA N is 32 ceil(K/32), and per stage shared memory sees 144 KB.
__global__ void __launch_bounds__(kThreads, 1) linear_wgrad_tma_kernel(const Args a, const __grid_constant__ CUtensorMap tm_g,
                                                                       const __grid_constant__ CUtensorMap tm_x) {
  extern __shared__ __align__(1024) unsigned char smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem);            // [kStages] TMA -> converters
  uint64_t* ready = full + kStages;                              // [kStages] converters (12 warps) -> MMA
  uint64_t* empty = ready + kStages;                             // [kStages] MMA -> TMA
  uint64_t* acc_full = empty + kStages;                          // [1]
  uint64_t* lo_free = acc_full + 1;                              // [kLoBufs] MMA -> X converters
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(lo_free + kLoBufs);
  const uint32_t a_plane = 4 * kGroup;                           // raw gY tile: 128 features x 32 samples
  const uint32_t b_plane = static_cast<uint32_t>(a.gx) * kGroup; // X tile (raw -> hi in place; lo in a lo buffer)
  const uint32_t stage_bytes = a_plane + b_plane;
  unsigned char* st0 = smem + 1024;                              // kStages slots [gY raw | X hi]
  unsigned char* lo0 = st0 + kStages * stage_bytes;              // kLoBufs X lo planes

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int ot = blockIdx.x % a.o_tiles;
  const int split = blockIdx.x / a.o_tiles;
  const int o0 = ot * 128;
  const int64_t per = (a.chunks + a.splits - 1) / a.splits;
  const int64_t c_begin = split * per;
  const int64_t c_end = (c_begin + per) < a.chunks ? (c_begin + per) : a.chunks;
  const int64_t G = c_end > c_begin ? c_end - c_begin : 0;

  if (tid == 0) {
    WTRACE(0);
    for (int s = 0; s < kStages; ++s) { mbar_init(&full[s], 1); mbar_init(&ready[s], kConvWarps + 4); mbar_init(&empty[s], 1); }
    for (int j = 0; j < kLoBufs; ++j) mbar_init(&lo_free[j], 1);
    mbar_init(acc_full, 1);
    mbar_init_fence();
  }
  if (warp == kMmaWarp) tmem_alloc(tmem_slot, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (tid == 0) { WTRACE(1); }
  pdl_trigger();
  pdl_wait();                                                    // both operands and the partial buffer belong to
                                                                 // preceding kernels until they have completed
  if (warp < kConvWarps) {
    // ------------------------------ TMA issue (thread 0) + X conversion (256 threads)
    const uint32_t tx_bytes = static_cast<uint32_t>(4 + a.gx) * kGroup;
    auto issue = [&](int64_t i) {
      if (i >= G) return;
      const uint32_t it = static_cast<uint32_t>(i);
      const int s = it % kStages;
      mbar_wait(&empty[s], ((it / kStages) & 1u) ^ 1u);
      mbar_expect_tx(&full[s], tx_bytes);
      const int b0 = static_cast<int>((c_begin + i) * kR);
      const uint32_t base = smem_u32(st0) + static_cast<uint32_t>(s) * stage_bytes;
      for (int g = 0; g < 4; ++g) v2::tma_load_2d(base + g * kGroup, &tm_g, o0 + g * 32, b0, &full[s]);
      for (int g = 0; g < a.gx; ++g) v2::tma_load_2d(base + a_plane + g * kGroup, &tm_x, g * 32, b0, &full[s]);
    };
    if (tid == 0) { issue(0); issue(1); issue(2); }
    for (int64_t i = 0; i < G; ++i) {
      const uint32_t it = static_cast<uint32_t>(i);
      const int s = it % kStages;
      if (tid == 0 && i == 8) { WTRACE(9); }
      mbar_wait(&full[s], (it / kStages) & 1u);
      if (tid == 0 && i == 0) { WTRACE(2); }
      if (tid == 0 && i == 8) { WTRACE(10); }
      const int j = it % kLoBufs;
      mbar_wait(&lo_free[j], ((it / kLoBufs) & 1u) ^ 1u);          // MMAs of stage i - 2 have read this lo buffer
      unsigned char* st = st0 + static_cast<uint32_t>(s) * stage_bytes + a_plane + tid * 16;
      unsigned char* lo = lo0 + static_cast<uint32_t>(j) * b_plane + tid * 16;
      for (int g = 0; g < a.gx; ++g) {                             // elementwise: layout-agnostic 16 B sweep
        float4 v = *reinterpret_cast<const float4*>(st + g * kGroup);
        if (a.relu_in) { v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f); }
        float4 h, l;
        split_tf32(v.x, h.x, l.x); split_tf32(v.y, h.y, l.y); split_tf32(v.z, h.z, l.z); split_tf32(v.w, h.w, l.w);
        *reinterpret_cast<float4*>(st + g * kGroup) = h;
        *reinterpret_cast<float4*>(lo + g * kGroup) = l;
      }
      fence_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(&ready[s]);
      if (tid == 0 && i == 0) { WTRACE(3); }
      if (tid == 0 && i == 8) { WTRACE(11); }
      if (tid == 0) issue(i + 3);
      if (tid == 0 && i == 8) { WTRACE(12); }
    }
    if (tid == 0) { WTRACE(4); }
  } else if (warp == kMmaWarp) {
    // ------------------------------ MMA issuer (whole warp convergent, one elected lane issues)
    const uint32_t idesc = instr_desc_tf32(128, a.gx * 32) | (1u << 16);     // A: TMEM (K-major), B: MN-major smem
    const uint32_t d_main = tmem_base, d_small = tmem_base + kColSmall;
    for (int64_t i = 0; i < G; ++i) {
      const uint32_t it = static_cast<uint32_t>(i);
      const int s = it % kStages;
      mbar_wait(&ready[s], (it / kStages) & 1u);
      tc_fence_after();
      if (lane == 0 && i == 8) { WTRACE(13); }
      const int j = it % kLoBufs;
      const uint32_t a_hi = tmem_base + kColA + static_cast<uint32_t>(s) * 64;
      const uint64_t bd_hi = desc_mn(smem_u32(st0) + static_cast<uint32_t>(s) * stage_bytes + a_plane);
      const uint64_t bd_lo = desc_mn(smem_u32(lo0) + static_cast<uint32_t>(j) * b_plane);
      const uint32_t acc0 = i > 0 ? 1u : 0u;
      const bool last = i == G - 1;
      if (v2::elect_one()) {
        mma_stage4_ts_mn(d_main, d_small, a_hi, a_hi + 32, bd_hi, bd_lo, idesc, acc0);
        tc_commit(&empty[s]);
        tc_commit(&lo_free[j]);
        if (last) tc_commit(acc_full);
      }
      if (lane == 0 && i == 8) { WTRACE(14); }
      __syncwarp();
    }
    if (lane == 0) { WTRACE(5); }
  } else {
    // ------------------------------ warps 8-11: gY^T -> tensor memory, bias gradient, then the epilogue
    const int q = warp & 3;
    const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    // element (feature f, sample k) of the raw tile: group f / 32, row k, 32 B chunk ((f % 32) / 8) ^ (k & 3)
    const uint32_t f_off = static_cast<uint32_t>(q) * kGroup + static_cast<uint32_t>(lane & 7) * 4;
    const int c32 = lane >> 3;
    float bsum = 0.f;
    for (int64_t i = 0; i < G; ++i) {
      const uint32_t it = static_cast<uint32_t>(i);
      const int s = it % kStages;
      mbar_wait(&full[s], (it / kStages) & 1u);
      if (warp == kConvWarps && lane == 0 && i == 8) { WTRACE(15); }
      const unsigned char* raw = st0 + static_cast<uint32_t>(s) * stage_bytes + f_off;
      uint32_t hi[32], lo[32];
#pragma unroll
      for (int k = 0; k < 32; ++k) {
        const float v = *reinterpret_cast<const float*>(raw + k * 128 + ((c32 ^ (k & 3)) << 5));
        bsum += v;
        float h, l;
        split_tf32(v, h, l);
        hi[k] = __float_as_uint(h);
        lo[k] = __float_as_uint(l);
      }
      const uint32_t ta = lane_base + kColA + static_cast<uint32_t>(s) * 64;
      v2::tmem_st32(ta, hi);
      v2::tmem_st32(ta + 32, lo);
      v2::tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&ready[s]);
    }
    // epilogue: one partial [128, K + 1] per CTA, coalesced through a smem tile (the stage memory is idle now)
    const int W1 = a.K + 1;
    float* dst = a.partial + (static_cast<int64_t>(split) * a.o_tiles * 128 + o0 + q * 32) * W1;
    float* tb = reinterpret_cast<float*>(st0) + (warp - kConvWarps) * (32 * 33);
    if (G > 0) {
      mbar_wait(acc_full, 0);
      tc_fence_after();
    }
    if (warp == kConvWarps && lane == 0) { WTRACE(6); }
    for (int cb = 0; cb * 32 < a.K; ++cb) {
      uint32_t r[32], r2[32];
      if (G > 0) {
        const uint32_t taddr = lane_base + cb * 32;
        tmem_ld32(taddr, r);
        tmem_ld32(taddr + kColSmall, r2);
        tmem_ld_wait();
      }
      __syncwarp();
#pragma unroll
      for (int jj = 0; jj < 32; ++jj) tb[lane * 33 + jj] = G > 0 ? __uint_as_float(r[jj]) + __uint_as_float(r2[jj]) : 0.f;
      __syncwarp();
      const int col = cb * 32 + lane;
      if (col < a.K) {
#pragma unroll 8
        for (int i = 0; i < 32; ++i) dst[static_cast<int64_t>(i) * W1 + col] = tb[i * 33 + lane];
      }
    }
    dst[static_cast<int64_t>(lane) * W1 + a.K] = bsum;             // bias-gradient column of this thread's feature
    if (warp == kConvWarps && lane == 0) { WTRACE(7); }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kMmaWarp) tmem_dealloc(tmem_base, kTmemCols);
  if (tid == 0) { WTRACE(8); }
}
}  // namespace wg2

// Sum of the split-K partials in a FIXED order (bit-reproducible): 8 warps of a block share 32 output
// elements, warp g adds splits g, g + 8, ... (independent loads, unrolled), then the 8 sub-sums are added
// in order 0..7.  (One thread per element walked 148 dependent L2 round trips: 10 us per call.)
__global__ void __launch_bounds__(256) wgrad_reduce_kernel(const float* partial, int splits, int rows_pad, int N, int K,
                                                           float* gW, float* gb) {
  __shared__ float sub[8][33];
  pdl_wait();                                                    // the partials of the weight-gradient launch
  pdl_trigger();                                                 // writes gW / gb only: no GEMM prologue reads them
  const int W1 = K + 1;
  const int64_t total = static_cast<int64_t>(N) * W1;
  const int ex = threadIdx.x & 31, g = threadIdx.x >> 5;
  for (int64_t base = static_cast<int64_t>(blockIdx.x) * 32; base < total; base += static_cast<int64_t>(gridDim.x) * 32) {
    const int64_t e = base + ex;
    float acc = 0.f;
    int o = 0, i = 0;
    if (e < total) {
      o = static_cast<int>(e / W1);
      i = static_cast<int>(e - static_cast<int64_t>(o) * W1);
      const float* p = partial + static_cast<int64_t>(o) * W1 + i;
      const int64_t stride = static_cast<int64_t>(rows_pad) * W1;
#pragma unroll 8
      for (int s = g; s < splits; s += 8) acc += p[s * stride];
    }
    sub[g][ex] = acc;
    __syncthreads();
    if (g == 0 && e < total) {
      float t = sub[0][ex];
#pragma unroll
      for (int k = 1; k < 8; ++k) t += sub[k][ex];
      if (i < K) gW[static_cast<int64_t>(o) * K + i] = t;
      else if (gb) gb[o] = t;
    }
    __syncthreads();
  }
}

}  // namespace gemm
}  // namespace vcnf

using namespace vcnf;

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
// cuTensorMapEncodeTiled through the runtime's driver entry point (the library does not link libcuda).
static EncodeTiledFn tensor_map_encoder() {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess ||
      q != cudaDriverEntryPointSuccess)
    return nullptr;
  return reinterpret_cast<EncodeTiledFn>(fn);
}

extern "C" int vcnf_linear_tf32x3(const void* A, int64_t lda, const void* W, int64_t sw_n, int64_t sw_k,
                                  const void* bias, const void* residual, int64_t ldr, const void* mask,
                                  int64_t ldm, void* C, int64_t ldc, int64_t M, int N, int K, int relu_in,
                                  void* stream) {
  if (M == 0) return 0;   /* empty batch: pointers may be NULL */
  if (!A || !W || !C || M < 0 || N <= 0 || K <= 0 || lda < K || ldc < N) return VCNF_E_BADARG;
  if (K > gemm::kMaxK) return VCNF_E_UNSUPPORTED;
  const int sms = sm_count();
  if (sms <= 0) return static_cast<int>(cudaErrorNoDevice);
  static const bool use_v1 = [] { const char* e = getenv("VCNF_GEMM_V1"); return e && atoi(e) != 0; }();
  if (use_v1 && (relu_in & 4)) return VCNF_E_UNSUPPORTED;
  if (!use_v1) {
    gemm::v2::Args2 b{};
    b.A = static_cast<const float*>(A); b.lda = lda;
    b.W = static_cast<const float*>(W); b.sw_n = sw_n; b.sw_k = sw_k;
    b.bias = static_cast<const float*>(bias);
    b.residual = static_cast<const float*>(residual); b.ldr = ldr;
    b.mask = static_cast<const float*>(mask); b.ldm = ldm;
    b.C = static_cast<float*>(C); b.ldc = ldc;
    b.M = M; b.N = N; b.K = K; b.relu_in = relu_in & 1; b.w_dynamic = (relu_in >> 1) & 1; b.mask_last = (relu_in >> 2) & 1;
    b.n_tiles = (N + 127) / 128;
    const int64_t m_tiles2 = (M + gemm::v2::kTN - 1) / gemm::v2::kTN;
    int cpt2 = sms / b.n_tiles;
    if (cpt2 < 1) cpt2 = 1;
    if (cpt2 > m_tiles2) cpt2 = static_cast<int>(m_tiles2);
    b.ctas_per_ntile = cpt2;
    CUtensorMap tm{}, tm_e{}, tm_c{};
    b.use_tma = b.e_mode = b.c_tma = 0;
    static EncodeTiledFn encode = tensor_map_encoder();
    // 2-D fp32 view [rows, cols] of pitch ld; eligible when 16 B aligned with a 16 B multiple pitch
    auto make_map = [&](CUtensorMap* out, const void* ptr, int64_t cols, int64_t ld, uint32_t box_c, uint32_t box_r,
                        CUtensorMapSwizzle swz) {
      if (!encode || !ptr || ld % 4 != 0 || reinterpret_cast<uintptr_t>(ptr) % 16 != 0 || M >= (1ll << 31)) return false;
      const cuuint64_t gdim[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(M)};
      const cuuint64_t gstride[1] = {static_cast<cuuint64_t>(ld) * 4};
      const cuuint32_t box[2] = {box_c, box_r};
      const cuuint32_t estr[2] = {1, 1};
      return encode(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(ptr), gdim, gstride, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
    };
    if (make_map(&tm, A, K, lda, gemm::v2::kCK, gemm::v2::kTN, CU_TENSOR_MAP_SWIZZLE_128B)) b.use_tma = 1;
    if (mask && make_map(&tm_e, mask, N, ldm, 128, gemm::v2::kTN, CU_TENSOR_MAP_SWIZZLE_NONE)) b.e_mode = 2;
    else if (residual && make_map(&tm_e, residual, N, ldr, 128, gemm::v2::kTN, CU_TENSOR_MAP_SWIZZLE_NONE)) b.e_mode = 1;
    if (make_map(&tm_c, C, N, ldc, 128, gemm::v2::kTN, CU_TENSOR_MAP_SWIZZLE_NONE)) b.c_tma = 1;
    static bool attr2 = false;
    if (!attr2) {
      cudaFuncSetAttribute(gemm::v2::linear_ts_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           static_cast<int>(gemm::v2::kSmemBytes));
      attr2 = true;
    }
    launch_pdl(gemm::v2::linear_ts_kernel, static_cast<unsigned>(b.n_tiles * cpt2), gemm::v2::kThreads, gemm::v2::kSmemBytes,
               static_cast<cudaStream_t>(stream), b, tm, tm_e, tm_c);
    return launch_status();
  }
  gemm::Args a{};
  a.A = static_cast<const float*>(A); a.lda = lda;
  a.W = static_cast<const float*>(W); a.sw_n = sw_n; a.sw_k = sw_k;
  a.bias = static_cast<const float*>(bias);
  a.residual = static_cast<const float*>(residual); a.ldr = ldr;
  a.mask = static_cast<const float*>(mask); a.ldm = ldm;
  a.C = static_cast<float*>(C); a.ldc = ldc;
  a.M = M; a.N = N; a.K = K; a.relu_in = relu_in & 1;
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
  int rc;
  static EncodeTiledFn encode = tensor_map_encoder();
  static const bool no_tma = [] { const char* e = getenv("VCNF_WGRAD_V1"); return e && atoi(e) != 0; }();
  CUtensorMap tm_g{}, tm_x{};
  auto make_map = [&](CUtensorMap* out, const void* ptr, int64_t cols, int64_t ld) {
    if (!encode || ld % 4 != 0 || reinterpret_cast<uintptr_t>(ptr) % 16 != 0 || B >= (1ll << 31)) return false;
    const cuuint64_t gdim[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(B)};
    const cuuint64_t gstride[1] = {static_cast<cuuint64_t>(ld) * 4};
    const cuuint32_t box[2] = {32, static_cast<cuuint32_t>(gemm::wg2::kR)};
    const cuuint32_t estr[2] = {1, 1};
    return encode(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(ptr), gdim, gstride, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
  };
  if (!no_tma && make_map(&tm_g, gY, N, ldg) && make_map(&tm_x, X, K, ldx)) {
    gemm::wg2::Args w{};
    w.partial = a.partial; w.B = B; w.N = N; w.K = K; w.relu_in = relu_in;
    w.o_tiles = a.o_tiles; w.splits = a.splits; w.chunks = a.chunks;
    w.gx = (K + 31) / 32;
    const size_t smem = 1024 + (static_cast<size_t>(gemm::wg2::kStages) * (4 + w.gx) + static_cast<size_t>(gemm::wg2::kLoBufs) * w.gx) *
                                   gemm::wg2::kGroup;
    static bool attr_set2 = false;
    if (!attr_set2) {
      cudaFuncSetAttribute(gemm::wg2::linear_wgrad_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024);
      attr_set2 = true;
    }
    launch_pdl(gemm::wg2::linear_wgrad_tma_kernel, static_cast<unsigned>(w.o_tiles * w.splits), gemm::kThreads, smem, st, w,
               tm_g, tm_x);
    rc = launch_status();
  } else {
    const size_t smem = 1024 + static_cast<size_t>(gemm::kWStages) * 2 * (128 / 8 + static_cast<size_t>(a.NX) / 8) * gemm::kWSbo;
    static bool attr_set = false;
    if (!attr_set) {
      cudaFuncSetAttribute(gemm::linear_wgrad_tf32x3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
      attr_set = true;
    }
    gemm::linear_wgrad_tf32x3_kernel<<<static_cast<unsigned>(a.o_tiles * a.splits), gemm::kThreads, smem, st>>>(a);
    rc = launch_status();
  }
  if (rc) return rc;
  const int64_t total = static_cast<int64_t>(N) * (K + 1);
  int grid = static_cast<int>((total + 31) / 32);
  if (grid > 1184) grid = 1184;
  launch_pdl(gemm::wgrad_reduce_kernel, static_cast<unsigned>(grid), 256u, 0, st, static_cast<const float*>(a.partial), a.splits,
             a.o_tiles * 128, N, K, static_cast<float*>(gW), static_cast<float*>(gb));
  return launch_status();
}
