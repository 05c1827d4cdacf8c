// Phase timeline of the TMA-fed weight-gradient kernel (clock64 stamps), built with -DVCNF_GEMM_TRACE.
#define VCNF_GEMM_TRACE 1
#include "../vcnf_b200/csrc/gemm_tf32x3.cu"
#include <cstdio>
int main(int argc, char** argv) {
  const int64_t B = argc > 1 ? atoll(argv[1]) : 65536;
  const int K = 128, N = 128;
  float *G, *X, *gW, *gb; void* ws;
  cudaMalloc(&G, B * N * 4); cudaMalloc(&X, B * K * 4); cudaMalloc(&gW, N * K * 4); cudaMalloc(&gb, N * 4);
  cudaMemset(G, 0, B * N * 4); cudaMemset(X, 0, B * K * 4);
  cudaMalloc(&ws, vcnf_linear_wgrad_workspace(B, N, K));
  char* flush; cudaMalloc(&flush, 256 << 20);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float ms = 0;
  for (int r = 0; r < 3; ++r) {
    cudaMemset(flush, r, 256 << 20);
    cudaEventRecord(e0);
    int rc = vcnf_linear_wgrad_tf32x3(G, N, X, K, gW, gb, ws, B, N, K, 1, nullptr);
    cudaEventRecord(e1);
    cudaError_t e = cudaDeviceSynchronize();
    if (rc || e) { printf("rc=%d err=%s\n", rc, cudaGetErrorString(e)); return 1; }
    cudaEventElapsedTime(&ms, e0, e1);
  }
  static long long h[160][16];
  cudaMemcpyFromSymbol(h, vcnf::gemm::wg2::g_wtrace, sizeof(h));
  const char* names[9] = {"entry", "alloc+sync", "stage0 landed", "stage0 converted", "converter done", "MMA issued all",
                          "acc full (epilogue)", "partial stored", "exit"};
  double avg[9] = {0};
  for (int c = 0; c < 148; ++c) for (int i = 0; i < 9; ++i) avg[i] += (h[c][i] - h[c][0]) / 1965.0 / 148;
  printf("B=%lld event time (kernel + reduce) %.1f us; mean over CTAs (us):", (long long)B, ms * 1e3);
  for (int i = 0; i < 9; ++i) printf("  %s=%.2f", names[i], avg[i]);
  printf("\n");
  const char* n2[7] = {"X conv: wait full begin", "full seen", "converted+arrived", "next TMA issued", "MMA: ready seen", "MMA issued", "gY warp: full seen"};
  printf("stage 8 of CTA 0/73 (us from entry):");
  for (int c : {0, 73}) { for (int i = 0; i < 7; ++i) printf("  %s=%.2f", n2[i], (h[c][9 + i] - h[c][0]) / 1965.0); printf(" |"); }
  printf("\n");
  return 0;
}
