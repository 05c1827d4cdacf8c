// Convolutional conditioner of the Glow coupling layers (reference: normflow/nets/cnn.py:7-47 called from
// flows/affine/glow.py:41-56) for sm_100a.
//
//   * im2col / col2im  : patch matrices for the few-channel ends of the net (the 3x3 convolution that reads the
//                        C/2-channel coupling input, and the gradient w.r.t. the C-channel parameter output) and for
//                        the fp32 parity route, where every convolution is  im2col -> vcnf_linear_tf32x3.
//   * conv_gemm_bf16   : implicit-GEMM convolution on tcgen05, bf16 operands, fp32 accumulation in tensor memory.
//                        Activations are NHWC; the 3x3 taps are NINE SHIFTED TMA LOADS of the same tensor (4-D
//                        tensor map, out-of-bounds rows / columns come back as zeros = the padding), no patch matrix.
//                        out[p, n] = epi( sum_t sum_c act[p (+) t, c] * Wp[t][n][c] )
//                        Roles: warp 0 TMA producer, warp 1 MMA issue, warps 2-5 epilogue (bias, LeakyReLU or the
//                        LeakyReLU-adjoint mask, store NHWC bf16 or NCHW fp32); accumulators double-buffered in TMEM.
//   * conv_wgrad_bf16  : G[m, n] = sum_p X[p, m] Y[p, n]  (both operands pixel-major = MN-major for the tensor core,
//                        SWIZZLE_128B tiles exactly as TMA writes them), split over the pixels across CTAs, partials
//                        summed in a fixed order (deterministic).
#include <cstdlib>
#include <cuda.h>
#include <cuda_bf16.h>
#include "common.cuh"
#include "../../include/vcnf_b200.h"

namespace vcnf {
namespace conv {

// ------------------------------------------------------------------------------------------------ im2col / col2im
// col[p, t*C + c] = act(src[b, c, h + s*dy(t), w + s*dx(t)])   (0 outside the image; columns [T*C, Kp) are zero)
// p = (b*H + h)*W + w, t = ky*3 + kx, (dy, dx) = (ky-1, kx-1), s = +1 (flip = 0) or -1 (flip = 1: the patch
// matrix of the transposed convolution).  src has arbitrary element strides (NCHW or NHWC).
template <typename OutT>
__device__ __forceinline__ OutT to_out(float v);
template <> __device__ __forceinline__ float to_out<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 to_out<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

template <typename OutT>
__global__ void __launch_bounds__(256) im2col_kernel(const float* __restrict__ src, int64_t sb, int64_t sc, int64_t sh,
                                                    int64_t sw, OutT* __restrict__ col, int64_t ldcol, int64_t total,
                                                    int C, int H, int W, int T, int Kp, int flip, int act, float slope) {
  for (int64_t e = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; e < total;
       e += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t p = e / Kp;
    const int k = static_cast<int>(e - p * Kp);
    float v = 0.f;
    if (k < T * C) {
      const int t = k / C, c = k - t * C;
      int dy = 0, dx = 0;
      if (T == 9) { dy = t / 3 - 1; dx = t % 3 - 1; }
      if (flip) { dy = -dy; dx = -dx; }
      const int w = static_cast<int>(p % W);
      const int64_t bh = p / W;
      const int h = static_cast<int>(bh % H);
      const int64_t b = bh / H;
      const int hh = h + dy, ww = w + dx;
      if (hh >= 0 && hh < H && ww >= 0 && ww < W) {
        v = __ldg(src + b * sb + c * sc + hh * sh + ww * sw);
        if (act && v < 0.f) v *= slope;
      }
    }
    col[p * ldcol + k] = to_out<OutT>(v);
  }
}

// bf16 patch matrix, one thread per (pixel, 8 adjacent columns): blockIdx.y = column group, consecutive threads =
// consecutive pixels, so for an NCHW source each of the 8 gathers is a coalesced read of the warp and the 8 values leave
// as one 16 B store (the element-per-thread kernel above: 64-bit divisions and 2 B stores, 48 us for 16 MB).
__global__ void __launch_bounds__(256) im2col_bf16_vec_kernel(const float* __restrict__ src, int64_t sb, int64_t sc, int64_t sh,
                                                             int64_t sw, __nv_bfloat16* __restrict__ col, int64_t ldcol,
                                                             int P, int C, int H, int W, int T, int flip, int act,
                                                             float slope) {
  const int k0 = blockIdx.y * 8;
  int t0 = k0 / C, c0 = k0 - t0 * C;
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < P; p += gridDim.x * blockDim.x) {
    const int w = p % W, bh = p / W;
    const int h = bh % H, b = bh / H;
    const float* base = src + static_cast<int64_t>(b) * sb;
    int t = t0, c = c0;
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      v[j] = 0.f;
      if (t < T) {
        int dy = 0, dx = 0;
        if (T == 9) { dy = t / 3 - 1; dx = t - (t / 3) * 3 - 1; }
        if (flip) { dy = -dy; dx = -dx; }
        const int hh = h + dy, ww = w + dx;
        if (hh >= 0 && hh < H && ww >= 0 && ww < W) {
          float x = __ldg(base + c * sc + hh * sh + ww * sw);
          if (act && x < 0.f) x *= slope;
          v[j] = x;
        }
      }
      if (++c == C) { c = 0; ++t; }
    }
    uint4 o;
    __nv_bfloat162* ho = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
    for (int j = 0; j < 4; ++j) ho[j] = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
    *reinterpret_cast<uint4*>(col + static_cast<int64_t>(p) * ldcol + k0) = o;
  }
}

// Adjoint of im2col (fp32): gsrc[b, c, h, w] = act'(src) * sum_t gcol[p - s*off(t), t*C + c]  (gather form: every
// output element is written once, no atomics).
__global__ void __launch_bounds__(256) col2im_kernel(const float* __restrict__ gcol, int64_t ldcol,
                                                    const float* __restrict__ src, float* __restrict__ gsrc, int64_t sb,
                                                    int64_t sc, int64_t sh, int64_t sw, int64_t total, int C, int H, int W,
                                                    int T, int flip, int act, float slope) {
  for (int64_t e = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; e < total;
       e += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    // e enumerates (p, c) with c fastest
    const int64_t p = e / C;
    const int c = static_cast<int>(e - p * C);
    const int w = static_cast<int>(p % W);
    const int64_t bh = p / W;
    const int h = static_cast<int>(bh % H);
    const int64_t b = bh / H;
    float acc = 0.f;
    for (int t = 0; t < T; ++t) {
      int dy = 0, dx = 0;
      if (T == 9) { dy = t / 3 - 1; dx = t % 3 - 1; }
      if (flip) { dy = -dy; dx = -dx; }
      // col[q, t, c] reads src at q + off  =>  src position p receives from q = p - off
      const int hh = h - dy, ww = w - dx;
      if (hh >= 0 && hh < H && ww >= 0 && ww < W)
        acc += __ldg(gcol + ((b * H + hh) * W + ww) * ldcol + t * C + c);
    }
    const int64_t o = b * sb + c * sc + h * sh + w * sw;
    if (act && __ldg(src + o) < 0.f) acc *= slope;
    gsrc[o] = acc;
  }
}

// ------------------------------------------------------------------------------------------------ column sums
// out[c] = sum_p X[p, c]  (bias gradients).  Two stages, fixed order.  X bf16 [P, ld] or fp32 NCHW [B, C, S].
constexpr int kSumChunks = 296;   // 2 x 148
__global__ void __launch_bounds__(256) colsum_bf16_kernel(const __nv_bfloat16* __restrict__ X, int64_t ld, int64_t P, int C,
                                                         int64_t rows_per_chunk, float* __restrict__ partial) {
  const int chunk = blockIdx.x;
  const int64_t p0 = chunk * rows_per_chunk, p1 = p0 + rows_per_chunk < P ? p0 + rows_per_chunk : P;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float acc = 0.f;
    for (int64_t p = p0; p < p1; ++p) acc += __bfloat162float(X[p * ld + c]);
    partial[static_cast<int64_t>(chunk) * C + c] = acc;
  }
}
// Same sums for C % 8 == 0, C <= 256, 16 B aligned rows: a thread owns 8 adjacent channels (one 16 B load per row) and
// every (256 / (C / 8))-th row of the chunk, four rows in flight; the row groups are then added in a fixed order.
__global__ void __launch_bounds__(256) colsum_bf16_vec_kernel(const __nv_bfloat16* __restrict__ X, int64_t ld, int64_t P,
                                                             int C, int64_t rows_per_chunk, float* __restrict__ partial) {
  __shared__ float red[256 * 8];
  const int chunk = blockIdx.x, lanes = C >> 3, groups = 256 / lanes;
  const int cl = threadIdx.x % lanes, g = threadIdx.x / lanes;
  const int64_t p0 = chunk * rows_per_chunk, p1 = p0 + rows_per_chunk < P ? p0 + rows_per_chunk : P;
  float acc[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) acc[k] = 0.f;
  if (g < groups) {
    const __nv_bfloat16* col = X + cl * 8;
    int64_t p = p0 + g;
    for (; p + 3 * groups < p1; p += 4 * groups) {
      uint4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = *reinterpret_cast<const uint4*>(col + (p + static_cast<int64_t>(u) * groups) * ld);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v[u]);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float2 f = __bfloat1622float2(h[k]);
          acc[2 * k] += f.x; acc[2 * k + 1] += f.y;
        }
      }
    }
    for (; p < p1; p += groups) {
      const uint4 v = *reinterpret_cast<const uint4*>(col + p * ld);
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float2 f = __bfloat1622float2(h[k]);
        acc[2 * k] += f.x; acc[2 * k + 1] += f.y;
      }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) red[(g * lanes + cl) * 8 + k] = acc[k];
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float t = 0.f;
    for (int gg = 0; gg < groups; ++gg) t += red[gg * C + c];
    partial[static_cast<int64_t>(chunk) * C + c] = t;
  }
}
__global__ void __launch_bounds__(256) chansum_nchw_kernel(const float* __restrict__ X, int64_t B, int C, int S,
                                                          int64_t b_per_chunk, float* __restrict__ partial) {
  __shared__ float red[256];
  const int chunk = blockIdx.x, c = blockIdx.y;
  const int64_t b0 = chunk * b_per_chunk, b1 = b0 + b_per_chunk < B ? b0 + b_per_chunk : B;
  float acc = 0.f;
  for (int64_t b = b0; b < b1; ++b)
    for (int i = threadIdx.x; i < S; i += blockDim.x) acc += X[(b * C + c) * S + i];
  red[threadIdx.x] = acc;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (static_cast<int>(threadIdx.x) < o) red[threadIdx.x] += red[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) partial[static_cast<int64_t>(chunk) * C + c] = red[0];
}
__global__ void sum_partials_kernel(const float* __restrict__ partial, int n_chunks, int64_t n, float* __restrict__ out) {
  const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= n) return;
  float acc = 0.f;
  for (int k = 0; k < n_chunks; ++k) acc += partial[static_cast<int64_t>(k) * n + i];
  out[i] = acc;
}


// ------------------------------------------------------------------------------------------------ weight packs
// All bf16 operand layouts of one ConvNet2d (3x3 C0 -> Hd, 1x1 Hd -> Hd, 3x3 Hd -> C3; nets/cnn.py:29-47) from its
// three fp32 weight tensors in ONE launch -- as library ops (permute, reshape, pad, flip, cast) they were ~20
// launches per ConvNet2d call and step, 1 000 of the 5 100 launches of a cfg-5 step.  Segments of `out`, in order:
//   w1p [Hd][K1p]      w1p[o][t C0 + c] = w1[o][c][ky][kx]            (forward conv 1 on the patch matrix of x)
//   w2p [Hd][Hd]       w2p[o][i] = w2[o][i]
//   w3p [9][Np3][Hd]   w3p[t][o][i] = w3[o][i][ky][kx]                (implicit-GEMM conv 3)
//   wd3 [Hd][K3p]      wd3[i][t C3 + o] = w3[o][i][ky][kx]            (input gradient of conv 3 on the flipped patches of g)
//   wd2 [Hd][Hd]       wd2[i][o] = w2[o][i]
//   wd1 [9][Np1][Hd]   wd1[t][c][o] = w1[o][c][2 - ky][2 - kx]         (input gradient of conv 1, implicit GEMM)
// t = 3 ky + kx; padding entries are zero.  b3p [Np3] = b3 zero-padded (fp32).
struct PackArgs {
  const float* w1; const float* w2; const float* w3; const float* b3;
  __nv_bfloat16* out; float* b3p;
  int C0, Hd, C3, K1p, Np3, K3p, Np1, with_bwd;
};
__global__ void __launch_bounds__(256) convnet3_pack_kernel(const PackArgs a) {
  const int Hd = a.Hd, C0 = a.C0, C3 = a.C3;
  const int64_t n1 = static_cast<int64_t>(Hd) * a.K1p, n2 = static_cast<int64_t>(Hd) * Hd,
                n3 = static_cast<int64_t>(9) * a.Np3 * Hd, n4 = static_cast<int64_t>(Hd) * a.K3p, n5 = n2,
                n6 = static_cast<int64_t>(9) * a.Np1 * Hd;
  const int64_t total = n1 + n2 + n3 + (a.with_bwd ? n4 + n5 + n6 : 0);
  for (int64_t e = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; e < total;
       e += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    int64_t r = e;
    float v = 0.f;
    if (r < n1) {
      const int o = static_cast<int>(r / a.K1p), k = static_cast<int>(r - static_cast<int64_t>(o) * a.K1p);
      if (k < 9 * C0) { const int t = k / C0, c = k - t * C0; v = a.w1[(static_cast<int64_t>(o) * C0 + c) * 9 + t]; }
    } else if ((r -= n1) < n2) {
      v = a.w2[r];
    } else if ((r -= n2) < n3) {
      const int i = static_cast<int>(r % Hd); const int64_t q = r / Hd;
      const int o = static_cast<int>(q % a.Np3), t = static_cast<int>(q / a.Np3);
      if (o < C3) v = a.w3[(static_cast<int64_t>(o) * Hd + i) * 9 + t];
    } else if ((r -= n3) < n4) {
      const int i = static_cast<int>(r / a.K3p), k = static_cast<int>(r - static_cast<int64_t>(i) * a.K3p);
      if (k < 9 * C3) { const int t = k / C3, o = k - t * C3; v = a.w3[(static_cast<int64_t>(o) * Hd + i) * 9 + t]; }
    } else if ((r -= n4) < n5) {
      const int i = static_cast<int>(r / Hd), o = static_cast<int>(r - static_cast<int64_t>(i) * Hd);
      v = a.w2[static_cast<int64_t>(o) * Hd + i];
    } else {
      r -= n5;
      const int o = static_cast<int>(r % Hd); const int64_t q = r / Hd;
      const int c = static_cast<int>(q % a.Np1), t = static_cast<int>(q / a.Np1);
      if (c < C0) v = a.w1[(static_cast<int64_t>(o) * C0 + c) * 9 + (8 - t)];
    }
    a.out[e] = __float2bfloat16_rn(v);
  }
  if (blockIdx.x == 0)
    for (int o = threadIdx.x; o < a.Np3; o += blockDim.x) a.b3p[o] = (o < C3 && a.b3) ? a.b3[o] : 0.f;
}

// ------------------------------------------------------------------------------------------------ tcgen05 helpers
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait_sleep(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) __nanosleep(20);
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
__device__ __forceinline__ bool elect_one() {
  uint32_t leader;
  asm volatile("{.reg .b32 rx; .reg .pred px; elect.sync rx|px, 0xffffffff; selp.b32 %0, 1, 0, px;}" : "=r"(leader));
  return leader != 0;
}
// Shared-memory matrix descriptor (cute/arch/mma_sm100_desc.hpp SmemDescriptor): start >> 4 at [0,14), LBO >> 4 at
// [16,30), SBO >> 4 at [32,46), version 1 at [46,48), layout type at [61,64) (2 = SWIZZLE_128B).
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((saddr >> 4) & 0x3fffu);
  d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3fffu) << 16;
  d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3fffu) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(2) << 61;
  return d;
}
// Instruction descriptor, kind::f16: D = F32 (1 at [4,6)), A = B = BF16 (1 at [7,10) and [10,13)), majorness of A / B
// at bits 15 / 16 (0 = K-major, 1 = MN-major), N >> 3 at [17,23), M >> 4 at [24,29).
__host__ __device__ constexpr uint32_t instr_desc_bf16(int M, int N, int a_mn, int b_mn) {
  return (1u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(a_mn) << 15) | (static_cast<uint32_t>(b_mn) << 16) |
         (static_cast<uint32_t>(N >> 3) << 17) | (static_cast<uint32_t>(M >> 4) << 24);
}
// Four K = 16 steps of one 64-wide k-chunk, issued back to back by one lane.  `a_step` / `b_step` = descriptor
// start-address advance per step in 16 B units (K-major SW128: 32 B inside the swizzle row -> 2; MN-major: 16 pixel
// rows of 128 B -> 128).
__device__ __forceinline__ void mma_bf16_x4(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                            uint32_t acc, uint64_t a_step, uint64_t b_step) {
  asm volatile(
      "{\n\t"
      ".reg .pred pa, pt;\n\t"
      ".reg .b64 a1, a2, a3, b1, b2, b3;\n\t"
      "setp.ne.b32 pa, %4, 0;\n\t"
      "setp.eq.b32 pt, %3, %3;\n\t"
      "add.u64 a1, %1, %5;\n\t add.u64 a2, a1, %5;\n\t add.u64 a3, a2, %5;\n\t"
      "add.u64 b1, %2, %6;\n\t add.u64 b2, b1, %6;\n\t add.u64 b3, b2, %6;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, pa;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], a1, b1, %3, pt;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], a2, b2, %3, pt;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], a3, b3, %3, pt;\n\t"
      "}\n" ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(acc), "l"(a_step), "l"(b_step)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* map, int c0, int c1, int c2, int c3,
                                            uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(smem_u32(bar))
      : "memory");
}

// ------------------------------------------------------------------------------------------------ implicit GEMM
constexpr int kStages = 4;
constexpr int kEpiWarps = 8;                  // two warps per TMEM lane quarter, alternating 32-column chunks
constexpr int kThreads = (2 + kEpiWarps) * 32;   // warp 0 producer, warp 1 MMA, warps 2-9 epilogue
constexpr int kWgThreads = 192;               // weight-gradient kernel: warp 0 producer, warp 1 MMA, warps 2-5 epilogue
constexpr uint32_t kATile = 128 * 128;        // 128 pixels x 64 bf16 channels
constexpr uint32_t kWTileMax = 256 * 128;     // <= 256 output channels x 64 bf16 channels
constexpr uint32_t kStageBytes = kATile + kWTileMax;
constexpr size_t kSmemGemm = 1024 + static_cast<size_t>(kStages) * kStageBytes + 256;
constexpr uint32_t kTmemCols = 512;

struct GemmArgs {
  int64_t B;
  int H, W, Cp, taps;
  int Np, NT, n_tiles;
  int64_t m_tiles;
  int BW, BH, BB, hb;            // tile = BB images x BH rows x BW (= W) columns = 128 pixels; hb = ceil(H / BH)
  const float* bias;             // [>= n_tiles*NT] or NULL
  const __nv_bfloat16* mask;     // saved post-activation output [P, ldm] (LeakyReLU adjoint) or NULL
  int64_t ldm;
  float slope;
  int epi;                       // 0 none, 1 LeakyReLU(slope) after the bias, 2 multiply by (mask > 0 ? 1 : slope)
  void* out;
  int out_mode;                  // 0: NHWC bf16 [P, ldo];  1: NCHW fp32 [B, Cout, H, W]
  int Cout;
  int64_t ldo;
};

__global__ void __launch_bounds__(kThreads, 1) conv_gemm_bf16_kernel(const GemmArgs a,
                                                                      const __grid_constant__ CUtensorMap tm_a,
                                                                      const __grid_constant__ CUtensorMap tm_w) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + static_cast<size_t>(kStages) * kStageBytes);
  uint64_t* full = bars;                  // [kStages]
  uint64_t* empty = bars + kStages;       // [kStages]
  uint64_t* tfull = bars + 2 * kStages;   // [2]
  uint64_t* tempty = tfull + 2;           // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(tfull + s, 1); mbar_init(tempty + s, kEpiWarps); }
    mbar_init_fence();
  }
  if (warp == 1) tmem_alloc(tmem_slot, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int64_t total_tiles = a.m_tiles * a.n_tiles;
  const int kchunks = a.Cp / 64;
  const int ksteps = a.taps * kchunks;
  const uint32_t stage_tx = kATile + static_cast<uint32_t>(a.NT) * 128u;

  if (warp == 0) {
    // ---------------- TMA producer
    if (elect_one()) {
      uint32_t it = 0;
      for (int64_t tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int64_t m = tile / a.n_tiles;
        const int n = static_cast<int>(tile - m * a.n_tiles);
        const int b0 = static_cast<int>((m / a.hb) * a.BB);
        const int h0 = static_cast<int>(m % a.hb) * a.BH;
        for (int t = 0; t < a.taps; ++t) {
          int dy = 0, dx = 0;
          if (a.taps == 9) { dy = t / 3 - 1; dx = t % 3 - 1; }
          for (int kc = 0; kc < kchunks; ++kc, ++it) {
            const int s = it % kStages;
            const uint32_t ph = (it / kStages) & 1;
            mbar_wait(empty + s, ph ^ 1);
            unsigned char* st = smem + static_cast<size_t>(s) * kStageBytes;
            mbar_expect_tx(full + s, stage_tx);
            tma_load_4d(smem_u32(st), &tm_a, kc * 64, dx, h0 + dy, b0, full + s);
            tma_load_2d(smem_u32(st + kATile), &tm_w, kc * 64, t * a.Np + n * a.NT, full + s);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ---------------- MMA issue (whole warp convergent, one elected lane issues)
    const uint32_t idesc = instr_desc_bf16(128, a.NT, 0, 0);
    uint32_t it = 0, tcount = 0;
    for (int64_t tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++tcount) {
      const uint32_t buf = tcount & 1;
      mbar_wait(tempty + buf, ((tcount >> 1) & 1) ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + buf * 256;
      for (int ks = 0; ks < ksteps; ++ks, ++it) {
        const int s = it % kStages;
        const uint32_t ph = (it / kStages) & 1;
        mbar_wait(full + s, ph);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + static_cast<size_t>(s) * kStageBytes);
        const uint64_t a_desc = smem_desc_sw128(sa, 16, 1024);
        const uint64_t b_desc = smem_desc_sw128(sa + kATile, 16, 1024);
        if (elect_one()) {
          mma_bf16_x4(d_tmem, a_desc, b_desc, idesc, ks > 0 ? 1u : 0u, 2, 2);
          tc_commit(empty + s);
          if (ks == ksteps - 1) tc_commit(tfull + buf);
        }
        __syncwarp();
      }
    }
  } else {
    // ---------------- epilogue: warp w may read TMEM lanes [32 (w % 4), 32 (w % 4) + 32); the two warps of a lane
    // quarter take alternate 32-column chunks.  (With one warp per scheduler the epilogue -- ~300 dependent
    // instructions per chunk -- was the whole kernel: 15 k cycles per 128 x 256 tile against ~2 k for the MMAs; ncu:
    // tensor pipe 9.5 % active, issue 11.6 %.  Vector bias / mask loads, fmax for slope 0 and twice the warps.)
    const int q = warp & 3;
    const int half = (warp - 2) >> 2;          // 0 or 1
    const int r = q * 32 + lane;               // pixel row of the tile
    const int w_ = r % a.BW;
    const int h_ = (r / a.BW) % a.BH;
    const int bb = r / (a.BW * a.BH);
    const bool relu = a.slope == 0.f;
    uint32_t tcount = 0;
    for (int64_t tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++tcount) {
      const uint32_t buf = tcount & 1;
      const int64_t m = tile / a.n_tiles;
      const int n = static_cast<int>(tile - m * a.n_tiles);
      const int64_t b = (m / a.hb) * a.BB + bb;
      const int h = static_cast<int>(m % a.hb) * a.BH + h_;
      const bool live = b < a.B && h < a.H;
      const int64_t p = (b * a.H + h) * a.W + w_;
      mbar_wait_sleep(tfull + buf, (tcount >> 1) & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + buf * 256 + (static_cast<uint32_t>(q * 32) << 16);
      for (int c0 = half * 32; c0 < a.NT; c0 += 64) {
        uint32_t v[32];
        tmem_ld32(taddr + c0, v);
        const int nbase = n * a.NT + c0;
        const int ncol = a.NT - c0 < 32 ? a.NT - c0 : 32;      // 16 or 32 (NT is a multiple of 16)
        float4 bv[8];
        if (a.bias) {
          const float4* bp = reinterpret_cast<const float4*>(a.bias + nbase);
#pragma unroll
          for (int g = 0; g < 8; ++g) bv[g] = (g * 4 < ncol) ? __ldg(bp + g) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        uint4 mv[4];
        if (a.epi == 2 && live) {
          const uint4* mp = reinterpret_cast<const uint4*>(a.mask + p * a.ldm + nbase);
#pragma unroll
          for (int g = 0; g < 4; ++g) mv[g] = (g * 8 < ncol) ? __ldg(mp + g) : make_uint4(0u, 0u, 0u, 0u);
        }
        tmem_ld_wait();
        float f[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
        if (a.bias) {
#pragma unroll
          for (int g = 0; g < 8; ++g) {
            f[g * 4 + 0] += bv[g].x; f[g * 4 + 1] += bv[g].y; f[g * 4 + 2] += bv[g].z; f[g * 4 + 3] += bv[g].w;
          }
        }
        if (a.epi == 1) {
          if (relu) {
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.f);
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = f[j] > 0.f ? f[j] : f[j] * a.slope;
          }
        } else if (a.epi == 2 && live) {
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const uint32_t mw[4] = {mv[g].x, mv[g].y, mv[g].z, mv[g].w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              // bf16 sign / zero test on the raw halves: value > 0  <=>  sign clear and magnitude non-zero
              const uint32_t lo = mw[j] & 0xffffu, hi = mw[j] >> 16;
              if (!(lo != 0u && lo < 0x8000u)) f[g * 8 + 2 * j] *= a.slope;
              if (!(hi != 0u && hi < 0x8000u)) f[g * 8 + 2 * j + 1] *= a.slope;
            }
          }
        }
        if (live) {
          if (a.out_mode == 0) {
            __nv_bfloat16* op = reinterpret_cast<__nv_bfloat16*>(a.out) + p * a.ldo + nbase;
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              if (g * 8 < ncol) {
                uint4 o;
                __nv_bfloat162 t0 = __floats2bfloat162_rn(f[g * 8 + 0], f[g * 8 + 1]);
                __nv_bfloat162 t1 = __floats2bfloat162_rn(f[g * 8 + 2], f[g * 8 + 3]);
                __nv_bfloat162 t2 = __floats2bfloat162_rn(f[g * 8 + 4], f[g * 8 + 5]);
                __nv_bfloat162 t3 = __floats2bfloat162_rn(f[g * 8 + 6], f[g * 8 + 7]);
                o.x = *reinterpret_cast<uint32_t*>(&t0); o.y = *reinterpret_cast<uint32_t*>(&t1);
                o.z = *reinterpret_cast<uint32_t*>(&t2); o.w = *reinterpret_cast<uint32_t*>(&t3);
                *reinterpret_cast<uint4*>(op + g * 8) = o;
              }
            }
          } else {
            float* op = reinterpret_cast<float*>(a.out);
            const int64_t hw = static_cast<int64_t>(a.H) * a.W;
            const int64_t base = b * a.Cout * hw + static_cast<int64_t>(h) * a.W + w_;
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (j < ncol && nbase + j < a.Cout) op[base + (nbase + j) * hw] = f[j];
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty + buf);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, kTmemCols);
}

// ------------------------------------------------------------------------------------------------ weight gradient
constexpr int kWgPix = 64;                                 // pixels per stage
constexpr uint32_t kWgGroup = kWgPix * 128;                // one 64-channel group of a stage (8 KB)
constexpr uint32_t kWgStage = 6 * kWgGroup;                // X: 2 groups (128 channels), Y: <= 4 groups (256 channels)
constexpr size_t kSmemWg = 1024 + static_cast<size_t>(kStages) * kWgStage + 256;

struct WgArgs {
  int64_t P;
  int Mx, Ny, NT;                // NT = min(Ny, 256), multiple of 64
  int m_tiles, n_tiles, splits;
  int64_t pix_per_split;         // multiple of kWgPix
  float* partial;                // [splits][m_tiles*128][Ny]
};

__global__ void __launch_bounds__(kWgThreads, 1) conv_wgrad_bf16_kernel(const WgArgs a,
                                                                       const __grid_constant__ CUtensorMap tm_x,
                                                                       const __grid_constant__ CUtensorMap tm_y) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + static_cast<size_t>(kStages) * kWgStage);
  uint64_t* full = bars;
  uint64_t* empty = bars + kStages;
  uint64_t* tfull = bars + 2 * kStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tfull + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, 1); }
    mbar_init(tfull, 1);
    mbar_init_fence();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int tile = blockIdx.x / a.splits, sp = blockIdx.x % a.splits;
  const int mt = tile / a.n_tiles, nt = tile % a.n_tiles;
  const int64_t p0 = sp * a.pix_per_split;
  int64_t p1 = p0 + a.pix_per_split;
  if (p1 > a.P) p1 = a.P;
  const int nsteps = p1 > p0 ? static_cast<int>((p1 - p0 + kWgPix - 1) / kWgPix) : 0;
  const int ygroups = a.NT / 64;
  const uint32_t stage_tx = (2 + ygroups) * kWgGroup;

  if (warp == 0) {
    if (elect_one()) {
      for (int it = 0; it < nsteps; ++it) {
        const int s = it % kStages;
        const uint32_t ph = (it / kStages) & 1;
        mbar_wait(empty + s, ph ^ 1);
        unsigned char* st = smem + static_cast<size_t>(s) * kWgStage;
        const int pix = static_cast<int>(p0 + static_cast<int64_t>(it) * kWgPix);
        mbar_expect_tx(full + s, stage_tx);
        for (int g = 0; g < 2; ++g) tma_load_2d(smem_u32(st + g * kWgGroup), &tm_x, mt * 128 + g * 64, pix, full + s);
        for (int g = 0; g < ygroups; ++g)
          tma_load_2d(smem_u32(st + (2 + g) * kWgGroup), &tm_y, nt * a.NT + g * 64, pix, full + s);
      }
    }
  } else if (warp == 1) {
    const uint32_t idesc = instr_desc_bf16(128, a.NT, 1, 1);
    for (int it = 0; it < nsteps; ++it) {
      const int s = it % kStages;
      const uint32_t ph = (it / kStages) & 1;
      mbar_wait(full + s, ph);
      tc_fence_after();
      const uint32_t sa = smem_u32(smem + static_cast<size_t>(s) * kWgStage);
      // MN-major SWIZZLE_128B: LBO = next 64-channel group, SBO = next 8 pixel rows
      const uint64_t a_desc = smem_desc_sw128(sa, kWgGroup, 1024);
      const uint64_t b_desc = smem_desc_sw128(sa + 2 * kWgGroup, kWgGroup, 1024);
      if (elect_one()) {
        mma_bf16_x4(tmem_base, a_desc, b_desc, idesc, it > 0 ? 1u : 0u, 128, 128);
        tc_commit(empty + s);
        if (it == nsteps - 1) tc_commit(tfull);
      }
      __syncwarp();
    }
  } else {
    const int q = warp & 3;
    const int m = mt * 128 + q * 32 + lane;
    float* dst = a.partial + (static_cast<int64_t>(sp) * a.m_tiles * 128 + m) * a.Ny + nt * a.NT;
    if (nsteps > 0) {
      mbar_wait_sleep(tfull, 0);
      tc_fence_after();
    }
    const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    for (int c0 = 0; c0 < a.NT; c0 += 32) {
      uint32_t v[32];
      if (nsteps > 0) {
        tmem_ld32(taddr + c0, v);
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = 0u;
      }
#pragma unroll
      for (int g = 0; g < 8; ++g)
        *reinterpret_cast<uint4*>(dst + c0 + g * 4) = make_uint4(v[g * 4], v[g * 4 + 1], v[g * 4 + 2], v[g * 4 + 3]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 256);
}

// G[m, n] = sum over splits, fixed order; rows m >= Mx are padding and dropped.
__global__ void __launch_bounds__(256) wgrad_sum_kernel(const float* __restrict__ partial, int splits, int64_t rows_pad,
                                                       int Mx, int Ny, float* __restrict__ G) {
  const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= static_cast<int64_t>(Mx) * Ny) return;
  float acc = 0.f;
  for (int s = 0; s < splits; ++s) acc += partial[static_cast<int64_t>(s) * rows_pad * Ny + i];
  G[i] = acc;
}

}  // namespace conv
}  // namespace vcnf

using namespace vcnf;

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn conv_map_encoder() {
  static EncodeTiledFn cached = [] {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      return static_cast<EncodeTiledFn>(nullptr);
    return reinterpret_cast<EncodeTiledFn>(fn);
  }();
  return cached;
}
static bool bf16_map_2d(CUtensorMap* out, const void* ptr, int64_t cols, int64_t rows, int64_t ld, uint32_t box_c,
                        uint32_t box_r) {
  EncodeTiledFn enc = conv_map_encoder();
  if (!enc || reinterpret_cast<uintptr_t>(ptr) % 16 != 0 || ld % 8 != 0) return false;
  const cuuint64_t gdim[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
  const cuuint64_t gstride[1] = {static_cast<cuuint64_t>(ld) * 2};
  const cuuint32_t box[2] = {box_c, box_r};
  const cuuint32_t estr[2] = {1, 1};
  return enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), gdim, gstride, box, estr,
             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

extern "C" {

int vcnf_im2col(const void* src, int64_t sb, int64_t sc, int64_t sh, int64_t sw, void* col, int64_t ldcol, int64_t B,
                int C, int H, int W, int ksize, int Kp, int flip, int act, double slope, int out_bf16, void* stream) {
  if (B == 0) return 0;
  const int T = ksize * ksize;
  if (!src || !col || B < 0 || C <= 0 || H <= 0 || W <= 0 || (ksize != 1 && ksize != 3) || Kp < T * C || ldcol < Kp)
    return VCNF_E_BADARG;
  const int64_t total = B * H * W * Kp;
  int64_t blocks = (total + 255) / 256;
  const int64_t cap = static_cast<int64_t>(sm_count() > 0 ? sm_count() : VCNF_SM_COUNT_DEFAULT) * 16;
  if (blocks > cap) blocks = cap;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (out_bf16 && Kp % 8 == 0 && ldcol % 8 == 0 && reinterpret_cast<uintptr_t>(col) % 16 == 0 &&
      B * H * W < (1ll << 31)) {
    const int P = static_cast<int>(B * H * W);
    int bx = (P + 255) / 256;
    if (bx > 2048) bx = 2048;
    conv::im2col_bf16_vec_kernel<<<dim3(static_cast<unsigned>(bx), static_cast<unsigned>(Kp / 8)), 256, 0, st>>>(
        static_cast<const float*>(src), sb, sc, sh, sw, static_cast<__nv_bfloat16*>(col), ldcol, P, C, H, W, T, flip, act,
        static_cast<float>(slope));
  } else if (out_bf16)
    conv::im2col_kernel<__nv_bfloat16><<<static_cast<unsigned>(blocks), 256, 0, st>>>(
        static_cast<const float*>(src), sb, sc, sh, sw, static_cast<__nv_bfloat16*>(col), ldcol, total, C, H, W, T, Kp,
        flip, act, static_cast<float>(slope));
  else
    conv::im2col_kernel<float><<<static_cast<unsigned>(blocks), 256, 0, st>>>(
        static_cast<const float*>(src), sb, sc, sh, sw, static_cast<float*>(col), ldcol, total, C, H, W, T, Kp, flip, act,
        static_cast<float>(slope));
  return launch_status();
}

int vcnf_col2im(const void* gcol, int64_t ldcol, const void* src, void* gsrc, int64_t sb, int64_t sc, int64_t sh,
                int64_t sw, int64_t B, int C, int H, int W, int ksize, int flip, int act, double slope, void* stream) {
  if (B == 0) return 0;
  const int T = ksize * ksize;
  if (!gcol || !gsrc || (act && !src) || B < 0 || C <= 0 || H <= 0 || W <= 0 || (ksize != 1 && ksize != 3) ||
      ldcol < T * C)
    return VCNF_E_BADARG;
  const int64_t total = B * H * W * C;
  int64_t blocks = (total + 255) / 256;
  const int64_t cap = static_cast<int64_t>(sm_count() > 0 ? sm_count() : VCNF_SM_COUNT_DEFAULT) * 16;
  if (blocks > cap) blocks = cap;
  conv::col2im_kernel<<<static_cast<unsigned>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const float*>(gcol), ldcol, static_cast<const float*>(src), static_cast<float*>(gsrc), sb, sc, sh, sw,
      total, C, H, W, T, flip, act, static_cast<float>(slope));
  return launch_status();
}

int64_t vcnf_convnet3_pack_elems(int C0, int Hd, int C3, int with_bwd) {
  if (C0 <= 0 || Hd <= 0 || C3 <= 0) return 0;
  const int64_t K1p = (9 * C0 + 63) / 64 * 64, Np3 = (C3 + 15) / 16 * 16, K3p = (9 * C3 + 63) / 64 * 64,
                Np1 = (C0 + 15) / 16 * 16;
  int64_t n = static_cast<int64_t>(Hd) * K1p + static_cast<int64_t>(Hd) * Hd + 9 * Np3 * Hd;
  if (with_bwd) n += static_cast<int64_t>(Hd) * K3p + static_cast<int64_t>(Hd) * Hd + 9 * Np1 * Hd;
  return n;
}

int vcnf_convnet3_pack(const void* w1, const void* w2, const void* w3, const void* b3, int C0, int Hd, int C3,
                       int with_bwd, void* out_bf16, void* b3p, void* stream) {
  if (!w1 || !w2 || !w3 || !out_bf16 || !b3p || C0 <= 0 || Hd <= 0 || C3 <= 0) return VCNF_E_BADARG;
  conv::PackArgs a{};
  a.w1 = static_cast<const float*>(w1); a.w2 = static_cast<const float*>(w2); a.w3 = static_cast<const float*>(w3);
  a.b3 = static_cast<const float*>(b3);
  a.out = static_cast<__nv_bfloat16*>(out_bf16); a.b3p = static_cast<float*>(b3p);
  a.C0 = C0; a.Hd = Hd; a.C3 = C3;
  a.K1p = (9 * C0 + 63) / 64 * 64; a.Np3 = (C3 + 15) / 16 * 16; a.K3p = (9 * C3 + 63) / 64 * 64; a.Np1 = (C0 + 15) / 16 * 16;
  a.with_bwd = with_bwd ? 1 : 0;
  const int64_t total = vcnf_convnet3_pack_elems(C0, Hd, C3, with_bwd);
  int64_t blocks = (total + 255) / 256;
  if (blocks > 1184) blocks = 1184;
  conv::convnet3_pack_kernel<<<static_cast<unsigned>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(a);
  return launch_status();
}

int64_t vcnf_colsum_workspace(int C) { return static_cast<int64_t>(conv::kSumChunks) * C * 4; }

int vcnf_colsum(const void* X, int layout, int64_t rows, int C, int64_t ld_or_S, void* out, void* workspace,
                void* stream) {
  if (!out || !workspace || C <= 0 || rows < 0 || (rows > 0 && !X) || layout < 0 || layout > 1) return VCNF_E_BADARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (rows == 0) {
    cudaMemsetAsync(out, 0, static_cast<size_t>(C) * 4, st);
    return launch_status();
  }
  int64_t chunks = rows < conv::kSumChunks ? rows : conv::kSumChunks;
  const int64_t rpc = (rows + chunks - 1) / chunks;
  chunks = (rows + rpc - 1) / rpc;
  if (layout == 0) {
    if (ld_or_S < C) return VCNF_E_BADARG;
    if (C % 8 == 0 && C <= 256 && ld_or_S % 8 == 0 && reinterpret_cast<uintptr_t>(X) % 16 == 0)
      conv::colsum_bf16_vec_kernel<<<static_cast<unsigned>(chunks), 256, 0, st>>>(
          static_cast<const __nv_bfloat16*>(X), ld_or_S, rows, C, rpc, static_cast<float*>(workspace));
    else
      conv::colsum_bf16_kernel<<<static_cast<unsigned>(chunks), 256, 0, st>>>(static_cast<const __nv_bfloat16*>(X), ld_or_S,
                                                                              rows, C, rpc, static_cast<float*>(workspace));
  } else {
    if (ld_or_S <= 0 || C > 65535) return VCNF_E_BADARG;
    conv::chansum_nchw_kernel<<<dim3(static_cast<unsigned>(chunks), C), 256, 0, st>>>(
        static_cast<const float*>(X), rows, C, static_cast<int>(ld_or_S), rpc, static_cast<float*>(workspace));
  }
  int rc = launch_status();
  if (rc) return rc;
  launch_warp_reduce<float>(static_cast<const float*>(workspace), static_cast<int>(chunks), C, C,
                            static_cast<float*>(out), st);
  return launch_status();
}

int vcnf_conv_gemm_bf16(const void* act, int64_t B, int H, int W, int Cp, int taps, const void* Wp, int Np,
                        const void* bias, const void* mask, int64_t ldm, double slope, int epi, void* out, int out_mode,
                        int Cout, int64_t ldo, void* stream) {
  if (B == 0) return 0;
  if (!act || !Wp || !out || B < 0 || H <= 0 || W <= 0 || Cp <= 0 || Cp % 64 != 0 || (taps != 1 && taps != 9) ||
      Np <= 0 || Np % 16 != 0 || epi < 0 || epi > 2 || (epi == 2 && !mask) || out_mode < 0 || out_mode > 1 ||
      Cout <= 0 || Cout > Np)
    return VCNF_E_BADARG;
  // tile = 128 pixels = BB images x BH rows x W columns
  if (W > 128 || (128 % W) != 0) return VCNF_E_UNSUPPORTED;
  conv::GemmArgs a{};
  a.BW = W;
  a.BH = 128 / W < H ? 128 / W : H;
  if (128 % (a.BW * a.BH) != 0) return VCNF_E_UNSUPPORTED;
  a.BB = 128 / (a.BW * a.BH);
  a.hb = (H + a.BH - 1) / a.BH;
  if (a.BB > 1 && a.hb != 1) return VCNF_E_UNSUPPORTED;
  if (B >= (1ll << 31)) return VCNF_E_UNSUPPORTED;
  a.B = B; a.H = H; a.W = W; a.Cp = Cp; a.taps = taps; a.Np = Np;
  a.m_tiles = ((B + a.BB - 1) / a.BB) * a.hb;
  const int sms = sm_count();
  if (sms <= 0) return static_cast<int>(cudaErrorNoDevice);
  // column tile: the widest of {256, 128, 64} (or Np itself below 64) that still gives every SM a tile
  int NT = Np;
  if (Np > 64) {
    NT = Np % 256 == 0 ? 256 : (Np % 128 == 0 ? 128 : (Np % 64 == 0 ? 64 : 0));
    if (NT == 0) {
      if (Np <= 256) NT = Np; else return VCNF_E_UNSUPPORTED;
    }
    while (NT > 64 && NT % 2 == 0 && (NT / 2) % 32 == 0 && a.m_tiles * (Np / NT) < sms) NT /= 2;
  }
  if (out_mode == 0 && (NT % 32 != 0 || ldo % 8 != 0)) return VCNF_E_UNSUPPORTED;
  if (epi == 2 && (NT % 32 != 0 || ldm % 8 != 0)) return VCNF_E_UNSUPPORTED;
  a.NT = NT;
  a.n_tiles = Np / NT;
  if (bias && reinterpret_cast<uintptr_t>(bias) % 16 != 0) return VCNF_E_UNSUPPORTED;   // read as float4
  if (mask && reinterpret_cast<uintptr_t>(mask) % 16 != 0) return VCNF_E_UNSUPPORTED;
  a.bias = static_cast<const float*>(bias);
  a.mask = static_cast<const __nv_bfloat16*>(mask);
  a.ldm = ldm;
  a.slope = static_cast<float>(slope);
  a.epi = epi;
  a.out = out; a.out_mode = out_mode; a.Cout = Cout; a.ldo = ldo;

  EncodeTiledFn enc = conv_map_encoder();
  if (!enc || reinterpret_cast<uintptr_t>(act) % 16 != 0 || reinterpret_cast<uintptr_t>(Wp) % 16 != 0)
    return VCNF_E_UNSUPPORTED;
  CUtensorMap tm_a{}, tm_w{};
  {
    const cuuint64_t gdim[4] = {static_cast<cuuint64_t>(Cp), static_cast<cuuint64_t>(W), static_cast<cuuint64_t>(H),
                                static_cast<cuuint64_t>(B)};
    const cuuint64_t gstride[3] = {static_cast<cuuint64_t>(Cp) * 2, static_cast<cuuint64_t>(W) * Cp * 2,
                                   static_cast<cuuint64_t>(H) * W * Cp * 2};
    const cuuint32_t box[4] = {64, static_cast<cuuint32_t>(a.BW), static_cast<cuuint32_t>(a.BH),
                               static_cast<cuuint32_t>(a.BB)};
    const cuuint32_t estr[4] = {1, 1, 1, 1};
    if (enc(&tm_a, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(act), gdim, gstride, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return VCNF_E_UNSUPPORTED;
  }
  if (!bf16_map_2d(&tm_w, Wp, Cp, static_cast<int64_t>(taps) * Np, Cp, 64, static_cast<uint32_t>(NT)))
    return VCNF_E_UNSUPPORTED;
  static bool attr = false;
  if (!attr) {
    cudaFuncSetAttribute(conv::conv_gemm_bf16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         static_cast<int>(conv::kSmemGemm));
    attr = true;
  }
  const int64_t tiles = a.m_tiles * a.n_tiles;
  const unsigned grid = static_cast<unsigned>(tiles < sms ? tiles : sms);
  conv::conv_gemm_bf16_kernel<<<grid, conv::kThreads, conv::kSmemGemm, static_cast<cudaStream_t>(stream)>>>(a, tm_a, tm_w);
  return launch_status();
}

static void wgrad_plan(int64_t P, int Mx, int Ny, conv::WgArgs* a) {
  a->P = P; a->Mx = Mx; a->Ny = Ny;
  a->NT = Ny < 256 ? Ny : 256;
  a->m_tiles = (Mx + 127) / 128;
  a->n_tiles = (Ny + a->NT - 1) / a->NT;
  const int sms = sm_count() > 0 ? sm_count() : VCNF_SM_COUNT_DEFAULT;
  int splits = sms / (a->m_tiles * a->n_tiles);
  if (splits < 1) splits = 1;
  const int64_t steps = (P + conv::kWgPix - 1) / conv::kWgPix;
  if (splits > steps) splits = static_cast<int>(steps > 0 ? steps : 1);
  const int64_t steps_per = (steps + splits - 1) / splits;
  a->pix_per_split = steps_per * conv::kWgPix;
  a->splits = static_cast<int>((steps + steps_per - 1) / (steps_per > 0 ? steps_per : 1));
  if (a->splits < 1) a->splits = 1;
}

int64_t vcnf_conv_wgrad_workspace(int64_t P, int Mx, int Ny) {
  if (Mx <= 0 || Ny <= 0 || P < 0) return 0;
  conv::WgArgs a{};
  wgrad_plan(P, Mx, Ny, &a);
  return static_cast<int64_t>(a.splits) * a.m_tiles * 128 * Ny * 4;
}

int vcnf_conv_wgrad_bf16(const void* X, int64_t ldx, int Mx, const void* Y, int64_t ldy, int Ny, int64_t P, void* G,
                         void* workspace, void* stream) {
  if (!G || !workspace || Mx <= 0 || Ny <= 0 || P < 0 || Mx % 64 != 0 || Ny % 64 != 0 || (Ny > 256 && Ny % 256 != 0) ||
      (P > 0 && (!X || !Y)) || ldx < Mx || ldy < Ny)
    return VCNF_E_BADARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (P == 0) {
    cudaMemsetAsync(G, 0, static_cast<size_t>(Mx) * Ny * 4, st);
    return launch_status();
  }
  if (P >= (1ll << 31)) return VCNF_E_UNSUPPORTED;
  conv::WgArgs a{};
  wgrad_plan(P, Mx, Ny, &a);
  a.partial = static_cast<float*>(workspace);
  CUtensorMap tm_x{}, tm_y{};
  if (!bf16_map_2d(&tm_x, X, Mx, P, ldx, 64, conv::kWgPix) || !bf16_map_2d(&tm_y, Y, Ny, P, ldy, 64, conv::kWgPix))
    return VCNF_E_UNSUPPORTED;
  static bool attr = false;
  if (!attr) {
    cudaFuncSetAttribute(conv::conv_wgrad_bf16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         static_cast<int>(conv::kSmemWg));
    attr = true;
  }
  const unsigned grid = static_cast<unsigned>(a.m_tiles * a.n_tiles * a.splits);
  conv::conv_wgrad_bf16_kernel<<<grid, conv::kWgThreads, conv::kSmemWg, st>>>(a, tm_x, tm_y);
  int rc = launch_status();
  if (rc) return rc;
  const int64_t n = static_cast<int64_t>(Mx) * Ny;
  conv::wgrad_sum_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, st>>>(
      a.partial, a.splits, static_cast<int64_t>(a.m_tiles) * 128, Mx, Ny, static_cast<float*>(G));
  return launch_status();
}

}  // extern "C"
