// Microbenchmark: cost of back-to-back tcgen05.mma (M=128, K=16, bf16) as a function of N and of the number of
// independent TMEM accumulators they rotate over.  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_issue mma_issue.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "../../opticalflow_b200/csrc/ptx_sm100.cuh"
using namespace pwc;

__global__ void __launch_bounds__(128, 1) k(int N, int nacc, int iters, unsigned long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t base = (ptx::smem_addr(smem) + 1023u) & ~1023u;
  __shared__ uint64_t bar; __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 48 * 1024 / 4; i += 128) reinterpret_cast<uint32_t*>(smem + (base - ptx::smem_addr(smem)))[i] = 0;
  if (threadIdx.x == 0) { ptx::mbar_init(ptx::smem_addr(&bar), 1); ptx::fence_barrier_init(); ptx::fence_proxy_async(); }
  if (warp == 0) ptx::tmem_alloc(ptx::smem_addr(&slot), 512);
  ptx::tc_fence_before(); __syncthreads(); ptx::tc_fence_after();
  const uint32_t tm = slot;
  if (warp == 0) {
    const uint32_t idesc = ptx::make_idesc_bf16(128, N);
    const uint64_t da = ptx::make_kmajor_desc(base, 1024, 2u), db = ptx::make_kmajor_desc(base + 16384, 1024, 2u);
    unsigned long long t0 = 0, t1 = 0, t2 = 0;
    if (ptx::elect_one()) {
      t0 = clock64();
      for (int i = 0; i < iters; ++i) ptx::umma_bf16(tm + (uint32_t)((i % nacc) * N), da + 2 * (i & 3), db + 2 * (i & 3), idesc, i >= nacc);
      t1 = clock64();
      ptx::umma_commit(ptx::smem_addr(&bar));
    }
    __syncwarp();
    ptx::mbar_wait(ptx::smem_addr(&bar), 0);
    t2 = clock64();
    if (t0) { out[0] = t1 - t0; }
    if (threadIdx.x == 0) out[1] = t2;
    if (t0) out[2] = t0;
  }
  ptx::tc_fence_before(); __syncthreads();
  if (warp == 0) ptx::tmem_dealloc(tm, 512);
}

int main() {
  unsigned long long* d; cudaMalloc(&d, 64);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  const int iters = 2000;
  for (int N : {16, 32, 64, 128, 256})
    for (int nacc : {1, 2, 4}) {
      if (nacc * N > 512) continue;
      k<<<1, 128, 64 * 1024>>>(N, nacc, iters, d);
      unsigned long long h[3]; cudaMemcpy(h, d, 24, cudaMemcpyDeviceToHost);
      cudaError_t e = cudaDeviceSynchronize();
      printf("N=%3d nacc=%d: issue %.1f cyc/MMA, issue+drain %.1f cyc/MMA  (%s)\n", N, nacc, (double)h[0] / iters,
             (double)(h[1] - h[2]) / iters, cudaGetErrorString(e));
    }
  return 0;
}
