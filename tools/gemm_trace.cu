// Phase timeline of linear_ts_kernel (clock64 stamps per CTA): compiles the kernel source directly with
// -DVCNF_GEMM_TRACE.  Usage: gemm_trace [M] [K]
#define VCNF_GEMM_TRACE 1
#include "../vcnf_b200/csrc/gemm_tf32x3.cu"
#include <cstdio>
#include <vector>
int main(int argc, char** argv) {
  const int64_t M = argc > 1 ? atoll(argv[1]) : 65536;
  const int K = argc > 2 ? atoi(argv[2]) : 128, N = 128;
  const int do_flush = argc > 3 ? atoi(argv[3]) : 1;
  float *A, *W, *C, *bias;
  cudaMalloc(&A, M * K * 4); cudaMalloc(&W, N * K * 4); cudaMalloc(&C, M * N * 4); cudaMalloc(&bias, N * 4);
  cudaMemset(A, 0, M * K * 4); cudaMemset(W, 0, N * K * 4); cudaMemset(bias, 0, N * 4);
  char* flush; cudaMalloc(&flush, 256 << 20);
  for (int r = 0; r < 3; ++r) {
    if (do_flush) cudaMemset(flush, r, 256 << 20);
    int rc = vcnf_linear_tf32x3(A, K, W, K, 1, bias, A, K, nullptr, 0, C, N, M, N, K, 1, nullptr);
    cudaError_t e = cudaDeviceSynchronize();
    if (rc || e) { printf("rc=%d err=%s\n", rc, cudaGetErrorString(e)); return 1; }
  }
  static long long h[160][16];
  cudaMemcpyFromSymbol(h, vcnf::gemm::v2::g_trace, sizeof(h));
  const char* names[12] = {"entry", "alloc+sync", "W in TMEM", "stage0 converted", "MMA sees stage0", "acc0 full",
                           "tile0 stored", "producer done", "MMA issued all", "epilogue done", "exit", "prologue copies issued"};
  for (int cta : {0, 73, 147}) {
    printf("CTA %d (us from entry @1.965 GHz):", cta);
    for (int i = 0; i < 12; ++i) printf("  %s=%.2f", names[i], (h[cta][i] - h[cta][0]) / 1965.0);
    printf("\n");
  }
  double avg[12] = {0};
  for (int c = 0; c < 148; ++c) for (int i = 0; i < 12; ++i) avg[i] += (h[c][i] - h[c][0]) / 1965.0 / 148;
  printf("mean over CTAs:");
  for (int i = 0; i < 12; ++i) printf("  %s=%.2f", names[i], avg[i]);
  printf("\n");
  return 0;
}
