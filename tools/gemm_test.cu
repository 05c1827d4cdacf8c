// Standalone check of vcnf_linear_tf32x3 against a CPU double reference (GPU box; no torch).
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <cuda_runtime.h>
#include "../include/vcnf_b200.h"

static float frand() { return (float)rand() / RAND_MAX * 2.f - 1.f; }

static int run(int64_t M, int N, int K, int64_t lda, bool transposed_w, bool relu_in, bool use_bias, bool use_res,
               bool use_mask, int reps) {
  std::vector<float> A(M * lda), W((size_t)N * K), bias(N), R(M * N), Mk(M * N), C(M * N, -7.f);
  for (auto& v : A) v = frand();
  for (auto& v : W) v = frand() * 0.3f;
  for (auto& v : bias) v = frand();
  for (auto& v : R) v = frand();
  for (auto& v : Mk) v = frand();
  // logical W[n][k]; stored [N,K] (sw_n=K, sw_k=1) or transposed [K,N] (sw_n=1, sw_k=N)
  std::vector<float> Ws((size_t)N * K);
  for (int n = 0; n < N; ++n)
    for (int k = 0; k < K; ++k) Ws[transposed_w ? (size_t)k * N + n : (size_t)n * K + k] = W[(size_t)n * K + k];
  float *dA, *dW, *db, *dR, *dM, *dC;
  cudaMalloc(&dA, A.size() * 4); cudaMalloc(&dW, Ws.size() * 4); cudaMalloc(&db, N * 4);
  cudaMalloc(&dR, R.size() * 4); cudaMalloc(&dM, Mk.size() * 4); cudaMalloc(&dC, C.size() * 4);
  cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(dW, Ws.data(), Ws.size() * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(db, bias.data(), N * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(dR, R.data(), R.size() * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(dM, Mk.data(), Mk.size() * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(dC, C.data(), C.size() * 4, cudaMemcpyHostToDevice);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  int rc = 0;
  float ms = 0;
  for (int r = 0; r < reps; ++r) {
    if (r == reps - 1) cudaEventRecord(e0);
    rc = vcnf_linear_tf32x3(dA, lda, dW, transposed_w ? 1 : K, transposed_w ? N : 1, use_bias ? db : nullptr,
                            use_res ? dR : nullptr, N, use_mask ? dM : nullptr, N, dC, N, M, N, K, relu_in, nullptr);
    if (r == reps - 1) cudaEventRecord(e1);
  }
  cudaError_t err = cudaDeviceSynchronize();
  cudaEventElapsedTime(&ms, e0, e1);
  printf("M=%lld N=%d K=%d lda=%lld wT=%d relu=%d bias=%d res=%d mask=%d : rc=%d sync=%s  %.3f ms", (long long)M, N, K,
         (long long)lda, transposed_w, relu_in, use_bias, use_res, use_mask, rc, cudaGetErrorString(err), ms);
  if (rc || err != cudaSuccess) { printf("\n"); return 1; }
  cudaMemcpy(C.data(), dC, C.size() * 4, cudaMemcpyDeviceToHost);
  double worst = 0, worst_rel = 0; int64_t wi = -1; int bad = 0;
  const int64_t step = M > 2048 ? M / 1024 : 1;
  for (int64_t m = 0; m < M; m += step) {
    for (int n = 0; n < N; ++n) {
      double acc = 0, mag = 0;
      for (int k = 0; k < K; ++k) {
        double av = A[m * lda + k];
        if (relu_in && av < 0) av = 0;
        acc += av * W[(size_t)n * K + k];
        mag += fabs(av * W[(size_t)n * K + k]);
      }
      if (use_bias) acc += bias[n];
      if (use_mask && !(Mk[m * N + n] > 0)) acc = 0;
      if (use_res) acc += R[m * N + n];
      double e = fabs(acc - C[m * N + n]);
      if (e > worst) { worst = e; wi = m * N + n; }
      double rel = e / (mag + 1e-6);
      if (rel > worst_rel) worst_rel = rel;
      if (rel > 2e-6) ++bad;
    }
  }
  printf("  max|err|=%.3g (at m=%lld n=%lld got %.6f) max err/sum|terms|=%.3g bad=%d %s\n", worst, (long long)(wi / N),
         (long long)(wi % N), wi >= 0 ? C[wi] : 0.f, worst_rel, bad, bad ? "FAIL" : "ok");
  cudaFree(dA); cudaFree(dW); cudaFree(db); cudaFree(dR); cudaFree(dM); cudaFree(dC);
  return bad != 0;
}

int main(int argc, char** argv) {
  int fails = 0;
  fails += run(128, 128, 32, 32, false, false, false, false, false, 1);      // one tile, one k-chunk
  if (argc > 1 && atoi(argv[1]) == 1) return fails;
  if (argc > 1 && atoi(argv[1]) == 2) {   // timing sweep
    run(18944, 128, 128, 128, false, false, true, false, false, 4);   // 148 tiles: one per CTA
    run(65536, 128, 32, 32, false, false, true, false, false, 4);
    run(65536, 128, 64, 64, false, false, true, false, false, 4);
    run(65536, 128, 128, 128, false, false, true, false, false, 4);
    run(262144, 128, 128, 128, false, false, true, false, false, 4);
    return 0;
  }
  if (argc > 1 && atoi(argv[1]) == 3) { run(65536, 128, 128, 128, false, true, true, true, false, 3); return 0; }
  fails += run(128, 128, 128, 128, false, false, true, false, false, 1);
  fails += run(300, 128, 128, 128, false, true, true, true, false, 1);       // ragged M, relu, residual
  fails += run(1000, 128, 15, 15, false, false, true, false, false, 1);      // K=15, unaligned lda
  fails += run(4096, 322, 128, 128, false, false, true, false, false, 1);    // 3 column slices (66 valid in the last)
  fails += run(777, 128, 128, 128, true, false, false, true, true, 1);       // input-gradient form
  fails += run(777, 15, 128, 128, true, false, false, false, false, 1);      // dgrad of the first layer (N=15)
  fails += run(65536, 128, 128, 128, false, true, true, true, false, 4);     // bench shape, timed
  fails += run(65536, 322, 128, 128, false, false, true, false, false, 4);
  printf("%s (%d failing cases)\n", fails ? "GEMM TEST FAILED" : "GEMM TEST PASSED", fails);
  return fails;
}
