// Shared device helpers: mbarrier + 1-D bulk async copies (TMA engine, no tensor map needed for
// contiguous tiles), warp reductions, launch bookkeeping.  sm_100a only.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define VCNF_SM_COUNT_DEFAULT 148

namespace vcnf {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
// Make barrier initialisation visible to the async proxy (TMA engine).
__device__ __forceinline__ void mbar_init_fence() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t"
      "}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}
// global -> shared bulk copy, completion counted in bytes on `bar`.  dst/src 16 B aligned,
// bytes a multiple of 16.
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
// shared -> global bulk copy (bulk-group completion).
__device__ __forceinline__ void bulk_s2g(void* dst, const void* src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// Wait until the smem source of all committed bulk stores has been read.
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
// Order generic-proxy smem writes before async-proxy reads of the same smem.
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// out[i] = sum over chunks c of partial[c * stride + i], one WARP per output: lane l adds chunks l, l + 32, ... in
// order, then a fixed xor tree -- the same summation order on every run (bit-reproducible), and 32 independent
// loads in flight per output instead of one thread walking all chunks serially.
template <typename T>
__global__ void __launch_bounds__(256) warp_reduce_chunks_kernel(const T* __restrict__ partial, int chunks, int64_t stride,
                                                                 int64_t total, T* __restrict__ out) {
  const int64_t i = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (i >= total) return;
  T acc = T(0);
  for (int c = lane; c < chunks; c += 32) acc += partial[static_cast<int64_t>(c) * stride + i];
  acc = warp_sum(acc);
  if (lane == 0) out[i] = acc;
}
template <typename T>
inline void launch_warp_reduce(const T* partial, int chunks, int64_t stride, int64_t total, T* out, cudaStream_t st) {
  if (total <= 0) return;
  const int64_t blocks = (total * 32 + 255) / 256;
  warp_reduce_chunks_kernel<T><<<static_cast<unsigned>(blocks), 256, 0, st>>>(partial, chunks, stride, total, out);
}

inline int sm_count() {
  static int n = -1;
  if (n < 0) {
    int dev = 0, v = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) {
      cudaGetLastError();
      return 0;
    }
    n = v;
  }
  return n;
}

inline int launch_status() {
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : static_cast<int>(e);
}

}  // namespace vcnf
