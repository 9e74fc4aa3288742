// microbench.cu -- measured denominators for the rooflines bench.py reports (debug header).
// MEASURED_PEAKS.json holds the HBM and dense bf16 figures; the fp64 CUDA-core rate that bounds the
// LSHEP / ShepMod dense passes is not in it, so it is measured here: independent DFMA
// chains, enough of them per thread to cover the pipe latency, every SM filled.
#include "common.cuh"
#include "../../include/b200approx_debug.h"

namespace b200 {
namespace {

__global__ void __launch_bounds__(256) dfma_kernel(double* out, int iters, double seed) {
  double a[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = seed + threadIdx.x * 1e-9 + i;
  const double m = 1.0000001, c = 1e-9;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = fma(a[i], m, c);
  }
  double s = 0.0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += a[i];
  if (s == 12345.678) out[0] = s;  // keeps the chains alive
}

__global__ void __launch_bounds__(256) ffma_kernel(float* out, int iters, float seed) {
  float a[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = seed + threadIdx.x * 1e-6f + i;
  const float m = 1.0001f, c = 1e-6f;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = fmaf(a[i], m, c);
  }
  float s = 0.0f;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += a[i];
  if (s == 12345.678f) out[0] = s;
}

}  // namespace
}  // namespace b200

extern "C" int b200_debug_fma_peak(int32_t device, double* fp64_tflops, double* fp32_tflops) {
  using namespace b200;
  B200_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  B200_CUDA(cudaGetDeviceProperties(&prop, device));
  double* buf = nullptr;
  B200_CUDA(cudaMalloc(&buf, 64));
  cudaEvent_t e0, e1;
  B200_CUDA(cudaEventCreate(&e0));
  B200_CUDA(cudaEventCreate(&e1));
  const int blocks = prop.multiProcessorCount * 8, threads = 256;
  double best[2] = {0.0, 0.0};
  for (int which = 0; which < 2; ++which) {
    const int iters = which == 0 ? 4000 : 40000;
    for (int rep = 0; rep < 4; ++rep) {
      cudaEventRecord(e0);
      if (which == 0) dfma_kernel<<<blocks, threads>>>(buf, iters, 1.0 + rep);
      else ffma_kernel<<<blocks, threads>>>(reinterpret_cast<float*>(buf), iters, 1.0f + rep);
      cudaEventRecord(e1);
      B200_CUDA(cudaEventSynchronize(e1));
      float ms = 0;
      cudaEventElapsedTime(&ms, e0, e1);
      const double flop = 2.0 * 8.0 * iters * (double)blocks * threads;
      if (rep > 0) best[which] = std::max(best[which], flop / (ms * 1e-3) / 1e12);
    }
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(buf);
  if (fp64_tflops) *fp64_tflops = best[0];
  if (fp32_tflops) *fp32_tflops = best[1];
  return 0;
}
