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

static int run_wgrad(int64_t B, int N, int K, bool relu, int reps) {
  std::vector<float> G(B * N), X(B * K), gW((size_t)N * K, -3.f), gb(N, -3.f);
  for (auto& v : G) v = frand();
  for (auto& v : X) v = frand();
  float *dG, *dX, *dW, *db, *ws;
  int64_t wsb = vcnf_linear_wgrad_workspace(B, N, K);
  cudaMalloc(&dG, G.size() * 4); cudaMalloc(&dX, X.size() * 4); cudaMalloc(&dW, gW.size() * 4);
  cudaMalloc(&db, N * 4); cudaMalloc(&ws, wsb);
  cudaMemcpy(dG, G.data(), G.size() * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(dX, X.data(), X.size() * 4, cudaMemcpyHostToDevice);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  int rc = 0; float ms = 0;
  for (int r = 0; r < reps; ++r) {
    if (r == reps - 1) cudaEventRecord(e0);
    rc = vcnf_linear_wgrad_tf32x3(dG, N, dX, K, dW, db, ws, B, N, K, relu, nullptr);
    if (r == reps - 1) cudaEventRecord(e1);
  }
  cudaError_t err = cudaDeviceSynchronize();
  cudaEventElapsedTime(&ms, e0, e1);
  printf("WGRAD B=%lld N=%d K=%d relu=%d ws=%lld KB: rc=%d sync=%s %.3f ms", (long long)B, N, K, relu, (long long)wsb / 1024, rc,
         cudaGetErrorString(err), ms);
  if (rc || err != cudaSuccess) { printf("\n"); return 1; }
  cudaMemcpy(gW.data(), dW, gW.size() * 4, cudaMemcpyDeviceToHost);
  cudaMemcpy(gb.data(), db, N * 4, cudaMemcpyDeviceToHost);
  double worst = 0; int bad = 0;
  const int ostep = N > 64 ? 7 : 1, istep = K > 64 ? 5 : 1;
  for (int o = 0; o < N; o += ostep) {
    double sb = 0, sbm = 0;
    for (int64_t b = 0; b < B; ++b) { sb += G[b * N + o]; sbm += fabs(G[b * N + o]); }
    double eb = fabs(sb - gb[o]) / (sbm + 1e-6);
    if (eb > worst) worst = eb;
    if (eb > 2e-6) ++bad;
    for (int i = 0; i < K; i += istep) {
      double acc = 0, mag = 0;
      for (int64_t b = 0; b < B; ++b) {
        double xv = X[b * K + i]; if (relu && xv < 0) xv = 0;
        acc += (double)G[b * N + o] * xv; mag += fabs((double)G[b * N + o] * xv);
      }
      double e = fabs(acc - gW[(size_t)o * K + i]) / (mag + 1e-6);
      if (e > worst) worst = e;
      if (e > 2e-6) ++bad;
    }
  }
  printf("  max err/sum|terms|=%.3g bad=%d %s\n", worst, bad, bad ? "FAIL" : "ok");
  cudaFree(dG); cudaFree(dX); cudaFree(dW); cudaFree(db); cudaFree(ws);
  return bad != 0;
}

static void wgrad_probe(int mode) {
  const int64_t B = 32; const int N = 128, K = 16;
  std::vector<float> G(B * N, 0.f), X(B * K, 0.f), gW((size_t)N * K, -3.f), gb(N, -3.f);
  for (int64_t b = 0; b < 8; ++b) {
    for (int o = 0; o < N; ++o) G[b * N + o] = mode == 0 ? (float)((o + 1) * (1 << b)) : 1.f;
    for (int i = 0; i < K; ++i) X[b * K + i] = mode == 0 ? 1.f : (float)((i + 1) * (1 << b));
  }
  float *dG, *dX, *dW, *db, *ws;
  cudaMalloc(&dG, G.size() * 4); cudaMalloc(&dX, X.size() * 4); cudaMalloc(&dW, gW.size() * 4);
  cudaMalloc(&db, N * 4); cudaMalloc(&ws, vcnf_linear_wgrad_workspace(B, N, K));
  cudaMemcpy(dG, G.data(), G.size() * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(dX, X.data(), X.size() * 4, cudaMemcpyHostToDevice);
  vcnf_linear_wgrad_tf32x3(dG, N, dX, K, dW, db, ws, B, N, K, 0, nullptr);
  cudaDeviceSynchronize();
  cudaMemcpy(gW.data(), dW, gW.size() * 4, cudaMemcpyDeviceToHost);
  cudaMemcpy(gb.data(), db, N * 4, cudaMemcpyDeviceToHost);
  printf("PROBE mode %d: expect gW[o,i] = %s, gb[o] = %s\n", mode, mode == 0 ? "255*(o+1)" : "255*(i+1)",
         mode == 0 ? "255*(o+1)" : "8");
  for (int o = 0; o < 20; ++o) {
    printf("  o=%3d gb=%g :", o, gb[o]);
    for (int i = 0; i < K; ++i) printf(" %g", gW[o * K + i]);
    printf("\n");
  }
}

int main(int argc, char** argv) {
  if (argc > 1 && atoi(argv[1]) == 5) { wgrad_probe(0); wgrad_probe(1); return 0; }
  if (argc > 1 && atoi(argv[1]) == 4) {
    int f = 0;
    f += run_wgrad(32, 128, 128, false, 1);
    f += run_wgrad(1000, 128, 128, true, 1);
    f += run_wgrad(777, 128, 15, false, 1);
    f += run_wgrad(5000, 322, 128, false, 1);
    f += run_wgrad(65536, 128, 128, true, 3);
    f += run_wgrad(65536, 322, 128, false, 3);
    f += run_wgrad(65536, 128, 15, false, 3);
    printf("%s (%d failing)\n", f ? "WGRAD TEST FAILED" : "WGRAD TEST PASSED", f);
    return f;
  }
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
  if (argc > 1 && atoi(argv[1]) == 6) { run(262144, 128, 128, 128, false, false, true, false, false, 2); return 0; }
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
