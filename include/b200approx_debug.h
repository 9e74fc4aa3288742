/* b200approx_debug.h -- test and measurement hooks of libb200approx.so.  Not part of the
 * drop-in boundary (include/b200approx.h): nothing here has a counterpart in the reference,
 * and a port of util.approximate never calls it.  Used by tests/test_gpu_tensor.py. */
#ifndef B200APPROX_DEBUG_H
#define B200APPROX_DEBUG_H
#include "b200approx.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Run the tensor path on q <= 4096 queries and return the raw stage-1 values (q, cols) with
 * cols = 128*ceil(n/128), in fitted-row order: dots[p][j] = -2 s^2 (z_p - mu).(y_j - mu) as the
 * tensor cores accumulated it from the fp16 operand images, s = stat "tc_scale"; the padding
 * columns j >= n are 0. */
int b200_tc_debug_scores(b200_model* m, const double* z, int64_t q, int32_t k, float* scores,
                         int64_t cols);

/* Measured FMA rates of the CUDA cores (independent chains on every SM), the roofline denominators
 * bench.py uses for the dense fp64 / fp32 scans: TFLOP/s, 2 flops per FMA. */
int b200_debug_fma_peak(int32_t device, double* fp64_tflops, double* fp32_tflops);

#ifdef __cplusplus
}
#endif
#endif
