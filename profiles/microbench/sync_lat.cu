// sync_lat.cu -- latencies of the tcgen05 / mbarrier handshakes used by knn_tc_kernel.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o sync_lat sync_lat.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "../../util_b200/csrc/ptx.cuh"
using namespace b200;

__global__ void __launch_bounds__(128, 1) k(long long* out, int reps) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar[4];
  __shared__ uint32_t slot;
  __shared__ volatile long long t_arrive;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) { for (int i = 0; i < 4; ++i) ptx::mbar_init(ptx::smem_u32(&bar[i]), 1); ptx::fence_barrier_init(); }
  if (warp == 0) { ptx::tmem_alloc(ptx::smem_u32(&slot), 512); ptx::tmem_relinquish(); }
  for (int i = threadIdx.x; i < 65536 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  ptx::fence_proxy_async();
  ptx::tc_fence_before(); __syncthreads(); ptx::tc_fence_after();
  const uint32_t tm = slot;
  const uint32_t idesc = ptx::umma_idesc_f16(128, 128);
  const uint32_t sa = ptx::smem_u32(smem), sb = sa + 32768;
  long long acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (warp == 0 && lane == 0) {
    uint32_t ph = 0;
    for (int r = 0; r < reps; ++r) {
      // (0) commit with nothing pending: issue cost and arrival latency
      long long t0 = clock64();
      ptx::umma_commit(ptx::smem_u32(&bar[0]));
      long long t1 = clock64();
      while (!ptx::mbar_try_wait(ptx::smem_u32(&bar[0]), ph)) {}
      long long t2 = clock64();
      acc[0] += t1 - t0; acc[1] += t2 - t0;
      ph ^= 1;
      // (1) 10 MMAs (M128 N128 K16) then commit: issue cost of the MMAs, time to completion
      t0 = clock64();
      for (int i = 0; i < 10; ++i)
        ptx::umma_f16(tm, ptx::umma_desc(sa + (i % 5) * 4096, 2048, 128), ptx::umma_desc(sb + (i % 5) * 4096, 2048, 128), idesc, i > 0);
      t1 = clock64();
      ptx::umma_commit(ptx::smem_u32(&bar[0]));
      while (!ptx::mbar_try_wait(ptx::smem_u32(&bar[0]), ph)) {}
      t2 = clock64();
      acc[2] += t1 - t0; acc[3] += t2 - t0;
      ph ^= 1;
      // (2) already-complete try_wait cost
      t0 = clock64();
      bool okk = ptx::mbar_try_wait(ptx::smem_u32(&bar[0]), ph ^ 1);
      t1 = clock64();
      acc[4] += t1 - t0 + (okk ? 0 : 1000000);
    }
  }
  __syncthreads();
  // (3) plain arrive by warp 1 -> wake of warp 2 blocked in try_wait
  for (int r = 0; r < reps; ++r) {
    if (warp == 1 && lane == 0) {
      long long t = clock64(); while (clock64() - t < 3000) {}
      t_arrive = clock64();
      ptx::mbar_arrive(ptx::smem_u32(&bar[1]));
    }
    if (warp == 2 && lane == 0) {
      while (!ptx::mbar_try_wait(ptx::smem_u32(&bar[1]), r & 1)) {}
      long long t = clock64();
      acc[5] += t - t_arrive;
    }
    __syncthreads();
  }
  if (warp == 0 && lane == 0) for (int i = 0; i < 5; ++i) out[i] = acc[i];
  if (warp == 2 && lane == 0) out[5] = acc[5];
  ptx::tc_fence_before(); __syncthreads();
  if (warp == 0) ptx::tmem_dealloc(tm, 512);
}

int main() {
  long long* out; cudaMalloc(&out, 64);
  const int reps = 200;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  k<<<1, 128, 65536>>>(out, reps);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[8]; cudaMemcpy(h, out, 48, cudaMemcpyDeviceToHost);
  printf("%s\n", cudaGetErrorString(e));
  printf("commit (idle pipe): issue %.0f clk, arrive+observe %.0f clk\n", (double)h[0] / reps, (double)h[1] / reps);
  printf("10 MMAs 128x128x16: issue %.0f clk, issue->completion observed %.0f clk (ideal tensor time 640)\n", (double)h[2] / reps, (double)h[3] / reps);
  printf("try_wait on a completed phase: %.0f clk\n", (double)h[4] / reps);
  printf("mbarrier.arrive -> waiter observes: %.0f clk\n", (double)h[5] / reps);
  return 0;
}
