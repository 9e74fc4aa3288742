// mma_rate2.cu -- does shared-memory traffic from other warps (mbarrier polling, LDS/STS) slow SS-operand
// tcgen05.mma (M=128 N=128 saturates the 128 B/clk smem port by itself)?  And the rate of the TS form
// (A operand in TMEM).  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_rate2 mma_rate2.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "../../util_b200/csrc/ptx.cuh"
using namespace b200;

__device__ __forceinline__ void umma_f16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n}" ::"r"(d_tmem),
      "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(acc)
      : "memory");
}

// mode 0: idle other warps; 1: 8 warps poll an mbarrier (try_wait); 2: 8 warps stream LDS/STS
__device__ int g_commit_every = 0;  // >0: a tcgen05.commit after every group of 5 MMAs (1: one barrier, 2: two)
template <int ts, int vec>
__global__ void __launch_bounds__(320, 1) k(long long* out, int N, int nmma, int mode) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar, bar2, bar3;
  __shared__ uint32_t slot;
  __shared__ volatile int done;
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  if (threadIdx.x == 0) { ptx::mbar_init(ptx::smem_u32(&bar), 1); ptx::mbar_init(ptx::smem_u32(&bar2), 1); ptx::mbar_init(ptx::smem_u32(&bar3), 1); done = 0; ptx::fence_barrier_init(); }
  if (warp == 9) { ptx::tmem_alloc(ptx::smem_u32(&slot), 512); ptx::tmem_relinquish(); }
  for (int i = threadIdx.x; i < 131072 / 4; i += 320) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  ptx::fence_proxy_async();
  ptx::tc_fence_before(); __syncthreads(); ptx::tc_fence_after();
  const uint32_t idesc = ptx::umma_idesc_f16(128, N);
  const uint32_t sa = ptx::smem_u32(smem), sb = sa + 32768;
  if (warp == 9) {
    if (ptx::elect_one()) {
      // vec: derive the operands from a value read back from shared memory, so they live in vector
      // registers and every UTCHMMA needs R2UR moves (what happens in a big warp-specialised kernel)
      const uint32_t off = vec ? slot : 0;
      const uint64_t ad = ptx::umma_desc(sa + off, 2048, 128), bd = ptx::umma_desc(sb + off, N * 16, 128);
      long long t0 = clock64();
      for (int i = 0; i < nmma; i += 5) {
#pragma unroll
        for (int ks = 0; ks < 5; ++ks) {
          if (ts) umma_f16_ts(off + ((i / 5) & 1) * 192, off + 384 + ks * 8, bd + (uint64_t)(ks * (N * 2)), idesc, ks > 0);
          else ptx::umma_f16(off + ((i / 5) & 1) * 256, ad + (uint64_t)(ks * 256), bd + (uint64_t)(ks * (N * 2)), idesc, ks > 0);
        }
        if (g_commit_every >= 1) ptx::umma_commit(ptx::smem_u32(&bar2));
        if (g_commit_every >= 2) ptx::umma_commit(ptx::smem_u32(&bar3));
      }
      long long t1 = clock64();
      ptx::umma_commit(ptx::smem_u32(&bar));
      while (!ptx::mbar_try_wait(ptx::smem_u32(&bar), 0)) {}
      long long t2 = clock64();
      out[0] = t1 - t0; out[1] = t2 - t0;
      done = 1;
    }
    __syncwarp();
  } else if (warp < 8) {
    float acc = 0;
    if (mode == 1) { while (!done) { if (ptx::mbar_try_wait(ptx::smem_u32(&bar2), 0)) break; } }
    if (mode == 2) {
      float* f = reinterpret_cast<float*>(smem + 98304);
      while (!done) { for (int i = 0; i < 16; ++i) { acc += f[(threadIdx.x + i * 256) & 8191]; f[(threadIdx.x + i * 256 + 7) & 8191] = acc; } }
    }
    if (acc == 1.2345f) out[3] = 1;
  }
  ptx::tc_fence_before(); __syncthreads();
  if (warp == 9) ptx::tmem_dealloc(0, 512);
}

template <int ts, int vec>
void run(long long* out) {
  auto kern = k<ts, vec>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 131072);
  const char* names[] = {"others idle", "8 warps polling try_wait", "8 warps LDS/STS"};
  for (int N : {64, 96, 128, 256})
    for (int mode = 0; mode < 3; mode += 2) {
      if (ts && N == 256) continue;
      long long a[2], b[2];
      kern<<<1, 320, 131072>>>(out, N, 20, mode); cudaDeviceSynchronize();
      kern<<<1, 320, 131072>>>(out, N, 20, mode); cudaDeviceSynchronize(); cudaMemcpy(a, out, 16, cudaMemcpyDeviceToHost);
      kern<<<1, 320, 131072>>>(out, N, 820, mode); cudaError_t e = cudaDeviceSynchronize(); cudaMemcpy(b, out, 16, cudaMemcpyDeviceToHost);
      printf("%s %s N=%3d %-26s: %s  sustained %.1f clk/MMA (ideal %d)\n", ts ? "TS(A in TMEM)" : "SS(A in smem)", vec ? "vector-reg operands " : "uniform-reg operands", N, names[mode],
             cudaGetErrorString(e), (b[1] - a[1]) / 800.0, N / 2);
    }
}
int main(int argc, char** argv) {
  long long* out; cudaMalloc(&out, 64);
  if (argc > 1) {  // commit study
    for (int ce = 0; ce <= 2; ++ce) {
      cudaMemcpyToSymbol(g_commit_every, &ce, 4);
      printf("--- commits per group of 5 MMAs: %d\n", ce);
      run<0, 0>(out); run<1, 0>(out);
    }
    return 0;
  }
  run<0, 0>(out); run<0, 1>(out); run<1, 0>(out); run<1, 1>(out);
  return 0;
}
