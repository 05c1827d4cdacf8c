// Probe: tcgen05.mma kind::tf32 with MN-major (feature-contiguous) SWIZZLE_128B operands in shared memory,
// i.e. exactly what a TMA box {32 features, 8 samples} of a row-major [samples, features] matrix produces.
//   D[m, n] = sum_k A[m, k] * B[n, k],  M = 128, N = 32, K = 8 (one MMA) and K = 32 (4 k-steps)
// Layout (bytes): element (f, k) of an operand with F features: group g = f / 32 at g * LBO, sample row k at
// k * 128 (8-row atoms: SBO = 1024), 16 B chunk c = (f % 32) / 4 stored at chunk (c ^ (k & 7)).
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cmath>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t desc(uint32_t addr, uint32_t lbo, uint32_t sbo, uint32_t layout) {
  return (uint64_t)((addr >> 4) & 0x3fff) | ((uint64_t)((lbo >> 4) & 0x3fff) << 16) | ((uint64_t)((sbo >> 4) & 0x3fff) << 32) |
         (1ull << 46) | ((uint64_t)layout << 61);
}
constexpr int KS = 32;   // samples staged
__global__ void __launch_bounds__(128, 1) probe(const float* A, const float* B, float* D, int N, int ksteps, int mode, int sbo, int kadv) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int tid = threadIdx.x;
  unsigned char* sa = smem;                       // 4 groups x KS rows x 128 B
  unsigned char* sb = smem + 4 * KS * 128;        // N/32 groups
  const uint32_t lbo = KS * 128;
  for (int e = tid; e < 128 * KS; e += 128) {     // A[m][k] row-major [128][KS] in global
    const int m = e / KS, k = e % KS;
    const uint32_t off = mode == 0 ? (m / 32) * lbo + k * 128 + ((((m % 32) / 4) ^ (k & 7)) << 4) + (m % 4) * 4
                                   : (m / 32) * lbo + k * 128 + ((((m % 32) / 8) ^ (k & 3)) << 5) + (m % 8) * 4;
    *reinterpret_cast<float*>(sa + off) = A[e];
  }
  for (int e = tid; e < N * KS; e += 128) {
    const int n = e / KS, k = e % KS;
    const uint32_t off = mode == 0 ? (n / 32) * lbo + k * 128 + ((((n % 32) / 4) ^ (k & 7)) << 4) + (n % 4) * 4
                                   : (n / 32) * lbo + k * 128 + ((((n % 32) / 8) ^ (k & 3)) << 5) + (n % 8) * 4;
    *reinterpret_cast<float*>(sb + off) = B[e];
  }
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  if (tid < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" ::"r"(s32(&slot)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("fence.proxy.async.shared::cta;");
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t tm = slot;
  if (tid == 0) {
    // idesc: D f32, A/B tf32, a_major = b_major = MN (bits 15, 16), N >> 3 at 17, M >> 4 at 24
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(N >> 3) << 17) | ((128u >> 4) << 24);
    for (int kk = 0; kk < ksteps; ++kk) {
      const uint64_t ad = desc(s32(sa) + kk * kadv, lbo, sbo, mode == 0 ? 2 : 1);
      const uint64_t bd = desc(s32(sb) + kk * kadv, lbo, sbo, mode == 0 ? 2 : 1);
      asm volatile("{.reg .pred q; setp.ne.b32 q, %4, 0; tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, q;}" ::"r"(tm), "l"(ad), "l"(bd), "r"(idesc), "r"(kk > 0 ? 1 : 0));
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(s32(&bar)));
  }
  uint32_t done = 0;
  while (!done)
    asm volatile("{.reg .pred q; mbarrier.try_wait.parity.shared::cta.b64 q, [%1], 0; selp.b32 %0, 1, 0, q;}" : "=r"(done) : "r"(s32(&bar)));
  asm volatile("tcgen05.fence::after_thread_sync;");
  const int warp = tid >> 5, lane = tid & 31;
  for (int cb = 0; cb < N; cb += 32) {
    uint32_t r[32];
    const uint32_t taddr = tm + ((uint32_t)(warp * 32) << 16) + cb;
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;");
    for (int j = 0; j < 32; ++j) D[(warp * 32 + lane) * N + cb + j] = __uint_as_float(r[j]);
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (tid < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" ::"r"(tm));
}
int main() {
  for (int mode : {0, 1}) for (int sbo : {512, 1024}) for (int N : {32, 144}) for (int ksteps : {1, 4}) {
    const int K = ksteps * 8;
    const int NP = (N + 31) / 32 * 32;
    float *hA = new float[128 * KS](), *hB = new float[NP * KS](), *hD = new float[128 * N];
    srand(1);
    for (int m = 0; m < 128; ++m) for (int k = 0; k < K; ++k) hA[m * KS + k] = (float)(rand() % 17 - 8);
    for (int n = 0; n < N; ++n) for (int k = 0; k < K; ++k) hB[n * KS + k] = (float)(rand() % 13 - 6);
    float *A, *B, *D;
    cudaMalloc(&A, 128 * KS * 4); cudaMalloc(&B, NP * KS * 4); cudaMalloc(&D, 128 * N * 4);
    cudaMemcpy(A, hA, 128 * KS * 4, cudaMemcpyHostToDevice); cudaMemcpy(B, hB, NP * KS * 4, cudaMemcpyHostToDevice);
    const size_t smem = 4 * KS * 128 + (NP / 32) * KS * 128;
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    probe<<<1, 128, smem>>>(A, B, D, N, ksteps, mode, sbo, 1024);
    cudaError_t e = cudaDeviceSynchronize();
    cudaMemcpy(hD, D, 128 * N * 4, cudaMemcpyDeviceToHost);
    int bad = 0; double worst = 0;
    for (int m = 0; m < 128; ++m) for (int n = 0; n < N; ++n) {
      double ref = 0; for (int k = 0; k < K; ++k) ref += (double)hA[m * KS + k] * hB[n * KS + k];
      const double err = fabs(ref - hD[m * N + n]); if (err > 1e-3) ++bad; if (err > worst) worst = err;
    }
    printf("MN-major mode=%d sbo=%d tf32: N=%d K=%d: %s  bad=%d worst=%g  (D[0][0..3] = %g %g %g %g)\n", mode, sbo, N, K, cudaGetErrorString(e), bad, worst, hD[0], hD[1], hD[2], hD[3]);
  }
  return 0;
}
