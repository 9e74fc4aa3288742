// common.cuh -- shared declarations of libb200approx (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <math_constants.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <string>
#include <vector>
#include <algorithm>
#include <cmath>

#include "../../include/b200approx.h"

namespace b200 {

void set_error(const char* fmt, ...);

#define B200_CUDA(call)                                                              \
  do {                                                                               \
    cudaError_t e_ = (call);                                                         \
    if (e_ != cudaSuccess) {                                                         \
      b200::set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
      return 2;                                                                      \
    }                                                                                \
  } while (0)

#define B200_TRY(call)        \
  do {                        \
    int rc_ = (call);         \
    if (rc_) return rc_;      \
  } while (0)

// Grow-only device buffer (library-owned scratch reused across calls).
struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  int reserve(size_t bytes) {
    if (bytes <= cap) return 0;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    size_t want = bytes + bytes / 4 + 256;
    cudaError_t e = cudaMalloc(&p, want);
    if (e != cudaSuccess) {
      set_error("cudaMalloc(%zu) -> %s", want, cudaGetErrorString(e));
      return 2;
    }
    cap = want;
    return 0;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
  template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

// Tensor-path operand images of the fitted points (knn_tc.cu).
struct TcData {
  bool ready = false;
  int K = 0;            // padded contraction length (multiple of 16)
  int64_t tiles = 0;    // 128-point tiles
  double scale = 1.0;   // power of two applied after centring
  double ymax = 0.0;    // max over points of max(||yc||, ||fp16(yc)||) (scaled units)
  double cmax = 0.0;    // max |coordinate| of the raw data
  DevBuf mu;            // d doubles: per-dimension centre
  DevBuf img;           // tiles * 128 * K halfs, UMMA K-major no-swizzle tile images
  DevBuf nlo;           // tiles * 4 floats: per 32-position chunk a lower bound of the squared centred norms
  DevBuf perm;          // n int32: image position -> fitted row (positions ordered by centred norm)
  DevBuf xs;            // (n,d) fp64: the fitted rows in image order, gathered by the re-rank (absent for huge sets)
};

struct Stats {
  double uncertified = 0, ms_stage1 = 0, ms_rerank = 0, ms_weights = 0, ms_exact = 0,
         launches = 0, tc_queries = 0, exact_queries = 0, last_cap = 0, ms_prep = 0, last_keep = 0, last_slots = 0, last_halves = 0;
};

}  // namespace b200

struct b200_model {
  int device = 0;
  int64_t n = 0;
  int d = 0;
  int sm_count = 148;
  double* x = nullptr;  // (n,d) row-major fp64 on the device
  b200::TcData tc;
  b200::TcData tc_sm;  // images of the ShepMod radius filter (ordered and bounded by s^2 rw^2 - ||y||^2)
  // ShepMod / LSHEP fitted state
  b200::DevBuf rw, lf, la;  // rw (n); lshep f (r,n), a (r,n,d)
  int lshep_r = 0;
  b200::DevBuf x32;         // fitted rows as floats, rows padded with zeros to 4*ceil(d/4) (fp32 prefilters; lazy)
  double x_absmax = -1.0;   // max |coordinate| of the fitted rows (lazy; bound for the fp32 prefilters)
  b200::DevBuf vor_g;       // stand-alone Voronoi: n x n inner products of the fitted rows (+ overflow flag)
  bool vor_ready = false;
  bool shepmod_ready = false, lshep_ready = false;
  // cached response values for fused weighted sums
  b200::DevBuf y;
  int y_r = 0;  // columns of the response values currently on the device (0: none)
  // scratch
  b200::DevBuf zq, idx32, d2, wts, cnt, err, out, part_d2, part_idx, cand, cand_meta, zimg,
      flags, misc, ids64;
  // options
  int knn_path = 0;
  int margin = -1;  // extra candidates kept by the tensor filter (-1: max(4, min(k/2, 16)))
  int tc_stats = 0;   // 1: run the instrumented stage-1 kernel and accumulate its counters in tc_stat_vals
  double tc_stat_vals[16] = {0};
  int tc_parts = 1;   // bulk copies per data tile
  int tc_slots = 0;   // force the depth of the data-tile ring (0: automatic)
  int tc_halves = 0;  // force 128*halves query rows per CTA (0: automatic)
  int tc_split = 0;   // internal: the data stream of this call is split over this many concurrent launches
  cudaStream_t aux_stream[4] = {nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t aux_event[4] = {nullptr, nullptr, nullptr, nullptr};
  int tc_mode = 0;  // ceilings: 1 = no selection, 2 = no epilogue work at all (outputs meaningless; stage 1 timed only)
  int no_tail_split = 0;         // 1: the last, partly filled round of query blocks is not split over idle SMs
  int force_split_fallback = 0;  // 1: uncertified queries go to the split brute-force scan instead of the rescan
  int64_t rows_per_call = 1 << 20;
  b200::Stats st;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev2 = nullptr, ev3 = nullptr;
  cudaStream_t copy_stream = nullptr;  // host->device query chunks overlap the previous chunk's kernels
  cudaEvent_t ev_copy = nullptr;
  // deferred fallback (b200_nn_predict's chunked batches): uncertified queries are only noted, as indices
  // relative to the whole batch, and resolved together after the last chunk
  bool defer_fallback = false;
  int64_t defer_base = 0;
  std::vector<int32_t> deferred;
  float* dbg_scores = nullptr;  // tests only: raw tensor-path scores are dumped here
};

namespace b200 {

// api.cu: picks the tensor or the exact kNN path
int knn_dispatch(b200_model* m, const double* z_dev, int64_t q, int k, int64_t self_offset,
                 int32_t* idx_dev, double* d2_dev, cudaStream_t st);
// knn_exact.cu
int knn_exact(b200_model* m, const double* z_dev, int64_t q, int k, int64_t self_offset,
              int32_t* idx_dev, double* d2_dev, const int32_t* qsel, int64_t nsel,
              cudaStream_t st);
int knn_rescan(b200_model* m, const double* z_dev, int64_t q, int k, int64_t self_offset, int32_t* idx_dev,
               double* d2_dev, const int32_t* fail_list, int nf, cudaStream_t st);
int merge_topk(const double* d2_parts, const int32_t* idx_parts, int parts, int64_t q, int k,
               double* d2_out, int32_t* idx_out, cudaStream_t st);
// knn_tc.cu
bool tc_supported(const b200_model* m, int k);
bool shepmod_filter_supported(const b200_model* m);
int shepmod_filter(b200_model* m, const double* z_dev, int64_t q, int32_t* cand_rows, int32_t* cand_cnt,
                   cudaStream_t st);
int knn_tc(b200_model* m, const double* z_dev, int64_t q, int k, int64_t self_offset,
           int32_t* idx_dev, double* d2_dev, cudaStream_t st);
// weights.cu
int nn_weights(b200_model* m, const double* z_dev, int64_t q, int k, int method,
               const int32_t* idx_dev, const double* d2_dev, const double* y_dev, int r,
               double* wts_dev, int32_t* cnt_dev, int32_t* err_dev, double* out_dev, cudaStream_t st);
// shepard_all.cu
int shepard_all(b200_model* m, const double* z_dev, int64_t q, const double* y_dev, int r, double* wts_dev,
                double* out_dev, cudaStream_t st);
// voronoi_all.cu
int voronoi_all(b200_model* m, const double* z_dev, int64_t q, int cap, const double* y_dev, int r,
                int32_t* ids_dev, double* wts_dev, int32_t* cnt_dev, int32_t* err_dev, double* out_dev,
                cudaStream_t st);
int eval_mesh(const double* ip_dev, int n, const double* ap_dev, int64_t q, double* mesh_dev, cudaStream_t st);
// shepmod.cu
int data_absmax(b200_model* m, cudaStream_t st, double* out);  // max |x_ij|, cached in the model
int data_f32(b200_model* m, cudaStream_t st, const float** out);  // padded float copy of the fitted rows, cached
int shepmod_fit(b200_model* m, cudaStream_t st, int* ier);
int shepmod_predict(b200_model* m, const double* z_dev, int64_t q, int cap, int32_t* ids_dev,
                    double* wts_dev, int32_t* cnt_dev, int32_t* ier_dev, const double* y_dev, int r,
                    double* out_dev, cudaStream_t st);
// lshep.cu
int lshep_fit(b200_model* m, const double* f_dev, int r, cudaStream_t st, int* ier, int mode = 0);
int lshep_predict(b200_model* m, double radius, const double* z_dev, int64_t q, double* out_dev,
                  int32_t* ier_dev, cudaStream_t st);

__host__ __device__ inline int64_t imin64(int64_t a, int64_t b) { return a < b ? a : b; }

// fp32 prefilter of the dense fp64 scans (ShepMod / LSHEP evaluation): a node certainly lies at
// distance >= L from the query when the fp32 squared distance `af` of the fp32-rounded coordinates
// satisfies af >= pre_thr(L).  With u = 2^-24, cm >= every |coordinate| involved, the fp32 difference
// vector t obeys ||t|| <= dist (1+u) + g0, g0 = 2.0001 u sqrt(d) cm, and the fma chain gives
// af <= ||t||^2 (1 + d u); so dist >= L follows from af >= (L (1+u) + g0)^2 (1 + d u).  The float
// evaluation below rounds every step upwards with room to spare.
__device__ __forceinline__ float pre_gap(int d, double cm) {  // g0, rounded up
  return (float)(2.0001 * 5.9604644775390625e-08 * sqrt((double)d) * cm) * 1.000001f;
}
__device__ __forceinline__ float pre_thr(float L_up, float g0, float eps1) {  // L_up >= L (1+u)
  const float lim = (L_up + g0) * 1.0000005f;
  return lim * lim * eps1;
}
__device__ __forceinline__ float pre_eps1(int d) { return 1.0f + (float)(d + 16) * 2.384185791015625e-07f; }
// Tile of `tc` fitted rows -> shared memory: doubles [tc][d] from x, floats [tc][dsf] from the padded
// float image (dsf = 4*ceil(d/4)); both are contiguous ranges, so plain linear copies.
__device__ __forceinline__ void load_tile_f64_f32(const double* __restrict__ src, const float* __restrict__ src32,
                                                  int tc, int d, int dsf, double* __restrict__ xs,
                                                  float* __restrict__ xf, int t, int nt) {
  for (int e = t; e < tc * d; e += nt) xs[e] = src[e];
  const float4* s4 = reinterpret_cast<const float4*>(src32);
  float4* d4 = reinterpret_cast<float4*>(xf);
  for (int e = t; e < tc * (dsf >> 2); e += nt) d4[e] = s4[e];
}
__device__ __forceinline__ float pre_up(double L) {  // float >= L (1 + 2^-24), finite or +inf
  return (float)L * 1.0000005f;
}

// Exact fp64 arithmetic with the reference's roundings: no FMA contraction.
__device__ __forceinline__ double sqdist_seq(const double* __restrict__ a,
                                             const double* __restrict__ b, int d) {
  double s = 0.0;
  for (int j = 0; j < d; ++j) {
    double t = __dsub_rn(a[j], b[j]);
    s = __dadd_rn(s, __dmul_rn(t, t));
  }
  return s;
}

// Same arithmetic (dimension order, separate roundings), 16-byte loads, sixteen (then eight) loads in flight:
// for gather-bound callers whose lanes each walk a different row.
// BPLAIN: b is read with ordinary loads (a row staged in shared memory) instead of __ldg.
template <bool BPLAIN = false>
__device__ __forceinline__ double sqdist_seq_wide(const double* __restrict__ a, const double* __restrict__ b, int d) {
  double s = 0.0;
  int j = 0;
  if ((d & 1) == 0 && ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b)) & 15) == 0) {
    const double2* a2 = reinterpret_cast<const double2*>(a);
    const double2* b2 = reinterpret_cast<const double2*>(b);
    const int h = d >> 1;
    int i = 0;
    for (; i + 16 <= h; i += 16) {  // sixteen 16-byte loads of the gathered row in flight
      double2 va[16], vb[16];
#pragma unroll
      for (int e = 0; e < 16; ++e) va[e] = __ldg(a2 + i + e);
#pragma unroll
      for (int e = 0; e < 16; ++e) vb[e] = BPLAIN ? b2[i + e] : __ldg(b2 + i + e);
#pragma unroll
      for (int e = 0; e < 16; ++e) {
        const double t0 = __dsub_rn(va[e].x, vb[e].x);
        s = __dadd_rn(s, __dmul_rn(t0, t0));
        const double t1 = __dsub_rn(va[e].y, vb[e].y);
        s = __dadd_rn(s, __dmul_rn(t1, t1));
      }
    }
    for (; i + 8 <= h; i += 8) {
      double2 va[8], vb[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) { va[e] = __ldg(a2 + i + e); vb[e] = BPLAIN ? b2[i + e] : __ldg(b2 + i + e); }
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const double t0 = __dsub_rn(va[e].x, vb[e].x);
        s = __dadd_rn(s, __dmul_rn(t0, t0));
        const double t1 = __dsub_rn(va[e].y, vb[e].y);
        s = __dadd_rn(s, __dmul_rn(t1, t1));
      }
    }
    j = i * 2;
  }
  for (; j < d; ++j) {
    const double t = __dsub_rn(a[j], b[j]);
    s = __dadd_rn(s, __dmul_rn(t, t));
  }
  return s;
}

}  // namespace b200
