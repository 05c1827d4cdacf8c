// Timing experiments on linear_ts_kernel: compile with -DVCNF_EXP_SKIP_CONVERT / _MMA / _EPI to drop one pipeline
// role's work (results are then wrong -- timing only) and see which role bounds the per-tile rate.
#include "../vcnf_b200/csrc/gemm_tf32x3.cu"
#include <cstdio>
#include <algorithm>
#include <vector>
static float time_one(int flush_on, char* flush, const float* A, int64_t lda, const float* W, const float* bias, const float* res,
                      int64_t ldr, float* C, int64_t ldc, int64_t M, int N, int K) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  std::vector<float> ts;
  for (int r = 0; r < 12; ++r) {
    if (flush_on) cudaMemsetAsync(flush, r, 256 << 20);
    cudaEventRecord(e0);
    int rc = vcnf_linear_tf32x3(A, lda, W, K, 1, bias, res, ldr, nullptr, 0, C, ldc, M, N, K, 1, nullptr);
    cudaEventRecord(e1);
    cudaError_t e = cudaDeviceSynchronize();
    if (rc || e) { printf("rc=%d err=%s\n", rc, cudaGetErrorString(e)); exit(1); }
    float ms; cudaEventElapsedTime(&ms, e0, e1); ts.push_back(ms * 1e3f);
  }
  std::sort(ts.begin(), ts.end());
  return ts[ts.size() / 2];
}
int main() {
  const int64_t M = 65536; const int K = 128;
  float *A, *W, *C, *bias, *R;
  cudaMalloc(&A, M * K * 4); cudaMalloc(&R, M * K * 4); cudaMalloc(&W, 384 * K * 4); cudaMalloc(&C, M * 324 * 4); cudaMalloc(&bias, 384 * 4);
  cudaMemset(A, 0, M * K * 4); cudaMemset(R, 0, M * K * 4); cudaMemset(W, 0, 384 * K * 4); cudaMemset(bias, 0, 384 * 4);
  char* flush; cudaMalloc(&flush, 256 << 20);
  printf("hidden 128x128 +residual: flushed %.1f us, warm L2 %.1f us\n",
         time_one(1, flush, A, K, W, bias, R, K, C, 128, M, 128, K), time_one(0, flush, A, K, W, bias, R, K, C, 128, M, 128, K));
  printf("hidden 128x128 no residual: flushed %.1f us, warm L2 %.1f us\n",
         time_one(1, flush, A, K, W, bias, nullptr, 0, C, 128, M, 128, K), time_one(0, flush, A, K, W, bias, nullptr, 0, C, 128, M, 128, K));
  printf("head 128->322 (ldc 324): flushed %.1f us, warm L2 %.1f us\n",
         time_one(1, flush, A, K, W, bias, nullptr, 0, C, 324, M, 322, K), time_one(0, flush, A, K, W, bias, nullptr, 0, C, 324, M, 322, K));
  return 0;
}
