// tmem_ld_bw.cu -- micro-benchmark: TMEM read rate (tcgen05.ld 32x32b.x32) per SM with
// 4 and 8 epilogue warps, bare and with the min3-tree filter of knn_tc_kernel.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tmem_ld_bw tmem_ld_bw.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "../../util_b200/csrc/ptx.cuh"
using namespace b200;

template <int MODE>
__global__ void __launch_bounds__(256, 1) k(int iters, long long* cyc, float* sink) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) { ptx::tmem_alloc(ptx::smem_u32(&slot), 512); ptx::tmem_relinquish(); }
  ptx::tc_fence_before(); __syncthreads(); ptx::tc_fence_after();
  const uint32_t base = slot + ((uint32_t)((warp & 3) * 32) << 16) + (warp >> 2) * 128;
  float acc = 3.0e38f; float tau = -1.0f; int hits = 0;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll 1
    for (int c = 0; c < 4; ++c) {
      float v[32];
      ptx::tmem_ld32(base + ((it & 1) * 256) + c * 32, v);
      ptx::tmem_ld_wait();
      if (MODE == 0) {
        acc = fminf(acc, v[0]);
      } else {
        float g[4];
#pragma unroll
        for (int gi = 0; gi < 4; ++gi) {
          const float* w = v + gi * 8;
          g[gi] = fminf(fminf(fminf(w[0], w[1]), w[2]), fminf(fminf(fminf(w[3], w[4]), w[5]), fminf(w[6], w[7])));
        }
        const float m = fminf(fminf(g[0], g[1]), fminf(g[2], g[3]));
        if (__any_sync(0xffffffffu, m < tau)) hits++;
        acc = fminf(acc, m);
      }
    }
  }
  const long long t1 = clock64();
  __syncthreads();
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
  if (acc == 12345.f) sink[threadIdx.x] = acc + hits;
  ptx::tc_fence_before(); __syncthreads();
  if (warp == 0) ptx::tmem_dealloc(slot, 512);
}

int main() {
  long long* cyc; float* sink;
  cudaMalloc(&cyc, 148 * 8); cudaMalloc(&sink, 4096);
  const int iters = 20000;
  for (int mode = 0; mode < 2; ++mode)
    for (int warps : {4, 8}) {
      if (mode == 0) k<0><<<148, warps * 32>>>(iters, cyc, sink); else k<1><<<148, warps * 32>>>(iters, cyc, sink);
      cudaError_t e = cudaDeviceSynchronize();
      long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
      long long mx = 0; for (auto v : h) mx = v > mx ? v : mx;
      const double bytes = (double)iters * 4 * 32 * 4 * 32 * warps;  // per SM
      printf("mode=%d warps=%d  %s  cycles=%lld  TMEM read %.1f B/clk/SM  (%.2f clk per 128-col row-tile per warp)\n",
             mode, warps, cudaGetErrorString(e), mx, bytes / mx, (double)mx / iters);
    }
  return 0;
}
