// Issue-rate probe for the spline kernels' instruction mix on sm_100a: scalar FFMA vs packed FFMA2 /
// FADD2 / FMUL2, MUFU.EX2, and FFMA mixed with ALU-pipe work.  Prints warp-instructions per cycle per
// SM sub-partition (one CTA of `nw` warps per SM, clock64 around an unrolled loop).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipe_rate tools/pipe_rate.cu && ./pipe_rate
#include <cstdio>
#include <cuda_runtime.h>

#define ITER 256
template <int MODE>
__global__ void probe(float* out, long long* cyc, float a, float b) {
  float2 r[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) r[i] = make_float2(threadIdx.x * 0.001f + i, i * 0.5f);
  const float2 A = make_float2(a, a * 0.5f), Bv = make_float2(b, b * 2.f);
  int q[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) q[i] = threadIdx.x + i;
  __syncthreads();
  long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < ITER; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (MODE == 0) { r[i].x = fmaf(r[i].x, A.x, Bv.x); r[i].y = fmaf(r[i].y, A.y, Bv.y); }      // 2 FFMA
      if (MODE == 1) { r[i] = __ffma2_rn(r[i], A, Bv); }                                          // 1 FFMA2
      if (MODE == 2) { r[i].x = exp2f(r[i].x); asm volatile("" ::: "memory"); }                   // MUFU.EX2 (+ range code)
      if (MODE == 3) { r[i].x = fmaf(r[i].x, A.x, Bv.x); q[i] = (q[i] ^ it) + 3; }                // FFMA + LOP3/IADD
      if (MODE == 4) { r[i] = __ffma2_rn(r[i], A, Bv); q[i] = (q[i] ^ it) + 3; }                  // FFMA2 + ALU
      if (MODE == 5) { r[i] = __fadd2_rn(r[i], A); }
      if (MODE == 6) { r[i].x = r[i].x + A.x; r[i].y = r[i].y + A.y; }
      if (MODE == 7) { asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(r[i].x)); }
      if (MODE == 8) { r[i].x = fmaxf(r[i].x, A.x * it); }
    }
  }
  long long t1 = clock64();
  float s = 0; int qs = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) { s += r[i].x + r[i].y; qs += q[i]; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s + qs;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int MODE>
void run(const char* name, int per_iter_instr, int nw) {
  float* out; long long* cyc;
  cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 8);
  probe<MODE><<<148, nw * 32>>>(out, cyc, 1.0001f, 0.5f);
  probe<MODE><<<148, nw * 32>>>(out, cyc, 1.0001f, 0.5f);
  long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  double warp_instr_per_smsp = double(ITER) * 8 * per_iter_instr * (nw / 4.0);
  printf("%-28s nw=%2d cycles=%8lld  -> %.3f listed warp-instr / cycle / SMSP\n", name, nw, h, warp_instr_per_smsp / h);
  cudaFree(out); cudaFree(cyc);
}

int main() {
  for (int nw : {4, 8, 16}) {
    run<0>("2x FFMA (3-reg)", 2, nw);
    run<1>("FFMA2", 1, nw);
    run<6>("2x FADD", 2, nw);
    run<5>("FADD2", 1, nw);
    run<2>("exp2f", 1, nw);
    run<7>("ex2.approx.ftz", 1, nw);
    run<3>("FFMA + 2 ALU", 3, nw);
    run<4>("FFMA2 + 2 ALU", 3, nw);
    run<8>("FMNMX+FMUL", 2, nw);
  }
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
