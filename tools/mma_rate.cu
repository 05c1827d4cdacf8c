// tcgen05.mma issue/throughput probe: one CTA, one thread issues R back-to-back MMAs, clock64 around
// issue..completion.  Variants: kind (tf32 / bf16), N, A from smem or TMEM, operand swizzle, number of
// accumulators rotated.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/build/mma_rate tools/mma_rate.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t desc(uint32_t addr, uint32_t lbo, uint32_t sbo, uint32_t layout) {
  return (uint64_t)((addr >> 4) & 0x3fff) | ((uint64_t)((lbo >> 4) & 0x3fff) << 16) | ((uint64_t)((sbo >> 4) & 0x3fff) << 32) |
         (1ull << 46) | ((uint64_t)layout << 61);
}
struct P { int kind, N, a_tmem, swz, nacc, R, ctas; long long* out; };
__global__ void __launch_bounds__(128, 1) probe(P p) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int tid = threadIdx.x;
  for (int i = tid; i < 96 * 1024 / 4; i += 128) reinterpret_cast<float*>(smem)[i] = 0.f;
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  if (tid < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(s32(&slot)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("fence.proxy.async.shared::cta;");
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t tm = slot;
  if (tid < 32) {
    const uint32_t fmt = p.kind != 1 ? 2u : 1u;     // tf32 : bf16
    const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(p.N >> 3) << 17) | ((128u >> 4) << 24);
    const uint32_t a_addr = s32(smem), b_addr = s32(smem) + 32768;
    const uint64_t ad = p.swz ? desc(a_addr, 16, 1024, 2) : desc(a_addr, 128, 256, 0);
    const uint64_t bd = p.swz ? desc(b_addr, 16, 1024, 2) : desc(b_addr, 128, 256, 0);
    const long long t0 = clock64();
    for (int r = 0; r < p.R; ++r) {
      const uint32_t d = tm + 256 + (p.nacc > 1 ? (r & 1) * 128 : 0);
      uint32_t leader;
      asm volatile("{.reg .b32 rx; .reg .pred px; elect.sync rx|px, 0xffffffff; selp.b32 %0, 1, 0, px;}" : "=r"(leader));
      if (!leader) continue;
      if (p.kind == 3 || p.kind == 4) {   // B (kind 3) or both operands (kind 4) MN-major, SWIZZLE_128B_BASE32B
        const uint32_t idm = idesc | (1u << 16) | (p.kind == 4 ? (1u << 15) : 0u);
        const uint64_t bm = desc(b_addr, 4096, 512, 1), am = desc(a_addr, 4096, 512, 1);
        if (p.kind == 3)
          asm volatile("{.reg .pred q; setp.ne.b32 q, %4, 0;\n"
                       "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, q;\n"
                       "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, q;\n"
                       "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, q;\n"
                       "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, q;\n"
                       "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, q;\n"
                       "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, q;\n"
                       "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, q;\n"
                       "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, q;}" ::"r"(d), "r"(tm), "l"(bm), "r"(idm), "r"(1));
        else
          asm volatile("{.reg .pred q; setp.ne.b32 q, %4, 0;\n"
                       "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, q;\n"
                       "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, q;\n"
                       "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, q;\n"
                       "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, q;\n"
                       "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, q;\n"
                       "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, q;\n"
                       "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, q;\n"
                       "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, q;}" ::"r"(d), "l"(am), "l"(bm), "r"(idm), "r"(1));
        continue;
      }
      if (p.kind == 2) {
        asm volatile("{.reg .pred q; setp.ne.b32 q, %4, 0;\n"
                     "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, q;\n"
                     "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, q;\n"
                     "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, q;\n"
                     "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, q;\n"
                     "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, q;\n"
                     "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, q;\n"
                     "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, q;\n"
                     "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, q;}" ::"r"(d), "r"(tm), "l"(bd), "r"(idesc), "r"(1));
        continue;
      }
      if (p.a_tmem) {
        if (p.kind == 0)
          asm volatile("{.reg .pred q; setp.ne.b32 q, %4, 0; tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, q;}" ::"r"(d), "r"(tm), "l"(bd), "r"(idesc), "r"(1));
        else
          asm volatile("{.reg .pred q; setp.ne.b32 q, %4, 0; tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, q;}" ::"r"(d), "r"(tm), "l"(bd), "r"(idesc), "r"(1));
      } else {
        if (p.kind == 0)
          asm volatile("{.reg .pred q; setp.ne.b32 q, %4, 0; tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, q;}" ::"r"(d), "l"(ad), "l"(bd), "r"(idesc), "r"(1));
        else
          asm volatile("{.reg .pred q; setp.ne.b32 q, %4, 0; tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, q;}" ::"r"(d), "l"(ad), "l"(bd), "r"(idesc), "r"(1));
      }
    }
    __syncwarp();
    const long long t1 = clock64();
    if (tid == 0)
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(s32(&bar)));
    uint32_t done = 0;
    while (!done)
      asm volatile("{.reg .pred q; mbarrier.try_wait.parity.shared::cta.b64 q, [%1], 0; selp.b32 %0, 1, 0, q;}" : "=r"(done) : "r"(s32(&bar)));
    const long long t2 = clock64();
    if (tid == 0) {
    p.out[blockIdx.x * 2] = t1 - t0;
    p.out[blockIdx.x * 2 + 1] = t2 - t0;
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (tid < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tm));
}
int main() {
  long long* out; cudaMalloc(&out, 148 * 16);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
  const int R = 512;
  printf("kind N a_tmem swz nacc ctas : issue clk/MMA, total clk/MMA (floor = N/2 clk at M=128)\n");
  for (int N : {16, 32, 64, 128, 256}) {
    P p{2, N, 1, 1, 1, R / 8, 148, out};
    probe<<<148, 128, 96 * 1024>>>(p);
    cudaDeviceSynchronize();
    long long h[2]; cudaMemcpy(h, out, 16, cudaMemcpyDeviceToHost);
    printf("tf32 x8-unrolled N=%3d a_tmem=1: %6.1f  %6.1f clk/MMA\n", N, (double)h[0] / R, (double)h[1] / R);
  }
  for (int kind : {3, 4}) for (int N : {32, 64, 128, 160}) {
    P p{kind, N, 1, 1, 1, R / 8, 148, out};
    probe<<<148, 128, 96 * 1024>>>(p);
    cudaDeviceSynchronize();
    long long h[2]; cudaMemcpy(h, out, 16, cudaMemcpyDeviceToHost);
    printf("tf32 x8-unrolled N=%3d %s MN-major (128B_BASE32B): %6.1f  %6.1f clk/MMA\n", N, kind == 3 ? "A tmem, B" : "A and B", (double)h[0] / R, (double)h[1] / R);
  }
  return 0;
  for (int ctas : {148})
    for (int kind : {0})
      for (int N : {64, 128, 256})
        for (int a_tmem : {0, 1})
          for (int swz : {0, 1})
            for (int nacc : {1, 2}) {
              if (N == 256 && nacc == 2) continue;
              P p{kind, N, a_tmem, swz, nacc, R, ctas, out};
              probe<<<ctas, 128, 96 * 1024>>>(p);
              cudaError_t e = cudaDeviceSynchronize();
              if (e) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
              long long h[2]; cudaMemcpy(h, out, 16, cudaMemcpyDeviceToHost);
              printf("%s N=%3d a_tmem=%d swz=%d nacc=%d ctas=%3d : %6.1f  %6.1f\n", kind ? "bf16" : "tf32", N, a_tmem, swz, nacc, ctas,
                     (double)h[0] / R, (double)h[1] / R);
            }
  return 0;
}
