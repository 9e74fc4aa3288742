// mma_rate.cu -- sustained tcgen05.mma rate (clk per instruction) for SS operands, kind::f16,
// M=128, by N and by smem layout; slope between a short and a long burst removes pipeline latency.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_rate mma_rate.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "../../util_b200/csrc/ptx.cuh"
using namespace b200;

__global__ void __launch_bounds__(128, 1) k(long long* out, int N, int rg, int nmma) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) { ptx::mbar_init(ptx::smem_u32(&bar), 1); ptx::fence_barrier_init(); }
  if (warp == 0) { ptx::tmem_alloc(ptx::smem_u32(&slot), 512); ptx::tmem_relinquish(); }
  for (int i = threadIdx.x; i < 98304 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  ptx::fence_proxy_async();
  ptx::tc_fence_before(); __syncthreads(); ptx::tc_fence_after();
  const uint32_t tm = slot;
  const uint32_t idesc = ptx::umma_idesc_f16(128, N);
  const uint32_t sa = ptx::smem_u32(smem), sb = sa + 32768;
  const int KC = 10;  // K = 80
  const uint32_t lbo_a = rg ? 128 : 2048, sbo_a = rg ? KC * 128 : 128, ks_a = rg ? 256 : 4096;
  const uint32_t lbo_b = rg ? 128 : N * 16, sbo_b = sbo_a, ks_b = rg ? 256 : N * 32;
  if (warp == 0) {
    if (ptx::elect_one()) {
      const uint64_t ad = ptx::umma_desc(sa, lbo_a, sbo_a), bd = ptx::umma_desc(sb, lbo_b, sbo_b);
      long long t0 = clock64();
      for (int i = 0; i < nmma; i += 5) {
#pragma unroll
        for (int ks = 0; ks < 5; ++ks)
          ptx::umma_f16(tm + ((i / 5) & 1) * 256, ad + (uint64_t)(ks * (ks_a >> 4)), bd + (uint64_t)(ks * (ks_b >> 4)), idesc, ks > 0);
      }
      long long t1 = clock64();
      ptx::umma_commit(ptx::smem_u32(&bar));
      while (!ptx::mbar_try_wait(ptx::smem_u32(&bar), 0)) {}
      long long t2 = clock64();
      out[0] = t1 - t0; out[1] = t2 - t0;
    }
    __syncwarp();
  }
  ptx::tc_fence_before(); __syncthreads();
  if (warp == 0) ptx::tmem_dealloc(tm, 512);
}

int main() {
  long long* out; cudaMalloc(&out, 64);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 98304);
  for (int N : {64, 128, 256})
    for (int rg = 0; rg < 2; ++rg) {
      long long a[2], b[2];
      k<<<1, 128, 98304>>>(out, N, rg, 20); cudaDeviceSynchronize();
      k<<<1, 128, 98304>>>(out, N, rg, 20); cudaDeviceSynchronize(); cudaMemcpy(a, out, 16, cudaMemcpyDeviceToHost);
      k<<<1, 128, 98304>>>(out, N, rg, 420); cudaError_t e = cudaDeviceSynchronize(); cudaMemcpy(b, out, 16, cudaMemcpyDeviceToHost);
      printf("M=128 N=%3d layout=%s: %s  issue %.1f clk/MMA, sustained exec %.1f clk/MMA (ideal %d)  [20: %lld/%lld, 420: %lld/%lld]\n", N,
             rg ? "row-group" : "chunk-major", cudaGetErrorString(e), (b[0] - a[0]) / 400.0, (b[1] - a[1]) / 400.0, N / 2, a[0], a[1], b[0], b[1]);
    }
  // full chip: does the rate hold with all 148 SMs busy?
  k<<<148, 128, 98304>>>(out, 128, 0, 4000); cudaDeviceSynchronize();
  long long c[2]; cudaMemcpy(c, out, 16, cudaMemcpyDeviceToHost);
  printf("148 CTAs, N=128, 4000 MMAs: %.1f clk/MMA\n", c[1] / 4000.0);
  return 0;
}
