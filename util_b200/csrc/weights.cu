// weights.cu -- sub-model weights on the k nearest neighbours and the weighted sums.
//
// Replaces, for the whole query batch, the per-query body of
// NearestNeighbor._predict (nearest_neighbor/__init__.py:78-83: construct
// self.method(), fit it on the k neighbours, predict the query) and
// WeightedApproximator.predict's sum(self.y[i]*w) (base.py:159-169):
//   Average  nearest_neighbor/__init__.py:7-14
//   Shepard  WEIGHT, shepard/weight.f90:1-34
//   Voronoi  PREDICT, voronoi/voronoi.f90:57-155 on the k neighbours
// These are gather/stream kernels bound by HBM: each query reads its k fitted
// rows once.  fp64 arithmetic keeps the reference's operand grouping and
// dimension-order sums without FMA contraction.
#include <math_constants.h>

#include "common.cuh"

namespace b200 {

namespace {

// One warp per query; lane i owns neighbour i (+32, ...).  KL = ceil(k/32) <= 4.
// Distances come from the kNN stage (d2 is bit-identical to the reference's
// sum of squared differences), so no fitted row is read again here: per query
// this stage moves k*(8+4+8) bytes in and k*8 (+8r) bytes out.
__global__ void __launch_bounds__(256)
avg_shepard_kernel(const double* __restrict__ d2, int64_t q,
                   int k, int method, const int32_t* __restrict__ idx,
                   const double* __restrict__ y, int r, double* __restrict__ wts,
                   int32_t* __restrict__ cnt, int32_t* __restrict__ err, double* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int64_t p = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (p >= q) return;
  const int32_t* id = idx + p * k;
  constexpr int KL = 4;
  double w[KL];
  int nb[KL];
  if (method == B200_AVERAGE) {
    const double a = 1.0 / (double)k;  // nearest_neighbor/__init__.py:12
#pragma unroll
    for (int u = 0; u < KL; ++u) {
      const int i = lane + 32 * u;
      nb[u] = i < k ? id[i] : -1;
      w[u] = a;
    }
  } else {
    double dist[KL];
    double mn = CUDART_INF, mx = 0.0;
#pragma unroll
    for (int u = 0; u < KL; ++u) {
      const int i = lane + 32 * u;
      nb[u] = i < k ? id[i] : -1;
      dist[u] = 0.0;
      if (i < k) {
        dist[u] = sqrt(d2[p * k + i]);  // weight.f90:13-15 (the kNN stage's exact d2, same arithmetic)
        mn = fmin(mn, dist[u]);
        mx = fmax(mx, dist[u]);
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o));
      mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    const double eps = 1.4901161193847656e-08;  // SQRT(EPSILON(1d0)), weight.f90:17
#pragma unroll
    for (int u = 0; u < KL; ++u) {
      if (mx < eps) w[u] = 1.0;                                  // :22-24
      else if (mn < eps) w[u] = dist[u] < eps ? 1.0 : 0.0;        // :25-28
      else w[u] = __ddiv_rn(1.0, __dmul_rn(dist[u], dist[u]));   // :29-31
    }
    // SUM(WTS) in index order (:34): serial over k through shuffles
    double s = 0.0;
#pragma unroll
    for (int u = 0; u < KL; ++u) {
      const int lim = min(32, k - 32 * u);
      for (int l = 0; l < lim; ++l) s = __dadd_rn(s, __shfl_sync(0xffffffffu, w[u], l));
    }
#pragma unroll
    for (int u = 0; u < KL; ++u) w[u] = __ddiv_rn(w[u], s);
  }
  if (wts) {
#pragma unroll
    for (int u = 0; u < KL; ++u) {
      const int i = lane + 32 * u;
      if (i < k) wts[p * k + i] = w[u];
    }
  }
  if (lane == 0) {
    if (cnt) cnt[p] = k;
    if (err) err[p] = 0;
  }
  if (out) {
    for (int c = 0; c < r; ++c) {
      double yv[KL];
#pragma unroll
      for (int u = 0; u < KL; ++u) yv[u] = nb[u] >= 0 ? __dmul_rn(y[(int64_t)nb[u] * r + c], w[u]) : 0.0;
      double s = 0.0;  // base.py:169: left to right from 0
#pragma unroll
      for (int u = 0; u < KL; ++u) {
        const int lim = min(32, k - 32 * u);
        for (int l = 0; l < lim; ++l) s = __dadd_rn(s, __shfl_sync(0xffffffffu, yv[u], l));
      }
      if (lane == 0) out[p * r + c] = s;
    }
  }
}

// Same arithmetic, one THREAD per query, for k <= KMAX <= 32: the k distances and weights of a query
// live in registers, so the index-order sums are plain loops instead of 2k shuffles by a half-empty
// warp (k = 16: 1.0 ms -> see profiles/r01 for 10^6 queries).  A warp reads 32 consecutive rows of
// d2 / idx, i.e. contiguous memory; only the y gather is scattered.
template <int KMAX>
__global__ void __launch_bounds__(128)
avg_shepard_thread_kernel(const double* __restrict__ d2, int64_t q, int k, int method,
                          const int32_t* __restrict__ idx, const double* __restrict__ y, int r,
                          double* __restrict__ wts, int32_t* __restrict__ cnt, int32_t* __restrict__ err,
                          double* __restrict__ out) {
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= q) return;
  double w[KMAX];
  if (method == B200_AVERAGE) {
    const double a = 1.0 / (double)k;  // nearest_neighbor/__init__.py:12
#pragma unroll
    for (int i = 0; i < KMAX; ++i) w[i] = a;
  } else {
    double mn = CUDART_INF, mx = 0.0;
#pragma unroll
    for (int i = 0; i < KMAX; ++i) {
      w[i] = 0.0;
      if (i < k) {
        w[i] = sqrt(d2[p * k + i]);  // weight.f90:13-15 (the kNN stage's exact d2, same arithmetic)
        mn = fmin(mn, w[i]);
        mx = fmax(mx, w[i]);
      }
    }
    const double eps = 1.4901161193847656e-08;  // SQRT(EPSILON(1d0)), weight.f90:17
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < KMAX; ++i) {
      if (i < k) {
        if (mx < eps) w[i] = 1.0;                               // :22-24
        else if (mn < eps) w[i] = w[i] < eps ? 1.0 : 0.0;        // :25-28
        else w[i] = __ddiv_rn(1.0, __dmul_rn(w[i], w[i]));      // :29-31
        s = __dadd_rn(s, w[i]);                                 // SUM(WTS) in index order (:34)
      }
    }
#pragma unroll
    for (int i = 0; i < KMAX; ++i) w[i] = __ddiv_rn(w[i], s);
  }
  if (wts) {
#pragma unroll
    for (int i = 0; i < KMAX; ++i)
      if (i < k) wts[p * k + i] = w[i];
  }
  if (cnt) cnt[p] = k;
  if (err) err[p] = 0;
  if (out) {
    int nb[KMAX];
#pragma unroll
    for (int i = 0; i < KMAX; ++i) nb[i] = i < k ? idx[p * k + i] : 0;
    for (int c = 0; c < r; ++c) {
      double s = 0.0;  // base.py:169: left to right from 0
#pragma unroll
      for (int i = 0; i < KMAX; ++i)
        if (i < k) s = __dadd_rn(s, nb[i] >= 0 ? __dmul_rn(y[(int64_t)nb[i] * r + c], w[i]) : CUDART_NAN);
      out[p * r + c] = s;
    }
  }
}

// Voronoi weights on the k neighbours: one block per query.  Shared memory holds
// the k rows, their Gram matrix G (the reference's DOTS cache, voronoi.f90:105-129),
// p.x_i and the per-centre maximum ratio.  Evaluated in the closed form the
// sorted / SKIP / CYCLE control flow prunes towards:
//   w_c = max(0, 1 - max_o max(0, num/den)),
//   num = (p.x_o + x_c.x_c) - (p.x_c + x_c.x_o), den = x_o.x_o + x_c.x_c - 2 x_c.x_o
// with the operand grouping of voronoi.f90:131-134.  ERROR=2 is raised when any
// pair among the k neighbours has |den| <= sqrt(eps) (:137,145-148).
__global__ void __launch_bounds__(256)
voronoi_kernel(const double* __restrict__ x, int d, const double* __restrict__ z, int64_t q, int k,
               const int32_t* __restrict__ idx, const double* __restrict__ y, int r,
               double* __restrict__ wts, int32_t* __restrict__ cnt, int32_t* __restrict__ err,
               double* __restrict__ out, int32_t* __restrict__ ids_out) {
  extern __shared__ double sm[];
  const int dp = d | 1;                 // odd row stride: conflict-free column walks
  double* P = sm;                       // [k][dp]
  double* G = P + (size_t)k * dp;       // [k][k+1]
  double* pdot = G + (size_t)k * (k + 1);
  double* w = pdot + k;
  double* pts = w + k;                  // [d] query
  __shared__ int s_err;
  const int64_t p = blockIdx.x;
  const int t = threadIdx.x, nt = blockDim.x;
  const int32_t* id = idx + p * k;
  if (t == 0) s_err = 0;
  for (int j = t; j < d; j += nt) pts[j] = z[p * d + j];
  for (int e = t; e < k * d; e += nt) {
    const int i = e / d, j = e - i * d;
    // a neighbour slot can only be empty (-1) for a query with non-finite coordinates: the result is then NaN
    P[(size_t)i * dp + j] = id[i] >= 0 ? x[(int64_t)id[i] * d + j] : CUDART_NAN;
  }
  __syncthreads();
  const double HUGEV = 1.7976931348623157e308;
  for (int e = t; e < k * (k + 1) / 2 + k; e += nt) {
    if (e < k) {  // p . x_i  (SUM(EVAL_PT * POINTS(:,I)), :131)
      double s = 0.0;
      for (int j = 0; j < d; ++j) s = __dadd_rn(s, __dmul_rn(pts[j], P[(size_t)e * dp + j]));
      pdot[e] = s;
      double dd = 0.0;  // :83-91 overflow check on the squared distance
      for (int j = 0; j < d; ++j) {
        double df = __dsub_rn(pts[j], P[(size_t)e * dp + j]);
        dd = __dadd_rn(dd, __dmul_rn(df, df));
      }
      if (dd > HUGEV) atomicMax(&s_err, 1);
    } else {
      // unrank e-k -> (a<=b) over the upper triangle: rows a' < a hold a(2k-a+1)/2 entries
      const int f0 = e - k;
      int a = (int)(((float)(2 * k + 1) - sqrtf((float)((2 * k + 1) * (2 * k + 1) - 8 * f0))) * 0.5f);
      a = a < 0 ? 0 : (a > k - 1 ? k - 1 : a);
      while (a > 0 && a * (2 * k - a + 1) / 2 > f0) --a;          // float rounding: at most one step either way
      while ((a + 1) * (2 * k - a) / 2 <= f0) ++a;
      const int b = a + (f0 - a * (2 * k - a + 1) / 2);
      double s = 0.0;
      for (int j = 0; j < d; ++j) s = __dadd_rn(s, __dmul_rn(P[(size_t)a * dp + j], P[(size_t)b * dp + j]));
      G[(size_t)a * (k + 1) + b] = s;
      G[(size_t)b * (k + 1) + a] = s;
      if (s > HUGEV) atomicMax(&s_err, 1);
    }
  }
  __syncthreads();
  const double MINV = 1.4901161193847656e-08;
  // the (centre, other) pairs are spread over all threads: `parts` threads share a centre, each takes every
  // parts-th other point; the maximum ratio is order independent, so the split changes no bit
  const int parts = nt >= 2 * k ? nt / k : 1;
  double* part_mr = pts + d;  // [parts][k]
  for (int e = t; e < k * parts; e += nt) {
    const int c = e % k, pt = e / k;
    double mr = 0.0;
    const double gcc = G[(size_t)c * (k + 1) + c], pc = pdot[c];
    for (int o = pt; o < k; o += parts) {
      if (o == c) continue;
      const double gco = G[(size_t)c * (k + 1) + o];
      const double num = __dsub_rn(__dadd_rn(pdot[o], gcc), __dadd_rn(pc, gco));                 // :131-132
      const double den = __dsub_rn(__dadd_rn(G[(size_t)o * (k + 1) + o], gcc), __dmul_rn(2.0, gco));  // :134
      if (fabs(den) > MINV) {
        const double ratio = __ddiv_rn(num, den);
        if (ratio > mr) mr = ratio;
      } else {
        atomicMax(&s_err, 2);
      }
    }
    part_mr[(size_t)pt * k + c] = mr;
  }
  __syncthreads();
  for (int c = t; c < k; c += nt) {
    double mr = part_mr[c];
    for (int pt = 1; pt < parts; ++pt) mr = fmax(mr, part_mr[(size_t)pt * k + c]);
    w[c] = fmax(__dsub_rn(1.0, mr), 0.0);  // :151
  }
  __syncthreads();
  // normalisation and output.  The sums keep the reference's order (one thread each); the divisions, the
  // compaction of the supported neighbours and the stores are spread over the block.
  __shared__ double s_sum;
  __shared__ int s_m;
  int* slot = reinterpret_cast<int*>(pdot);  // [k] output position of neighbour i, -1: no support (pdot is done with)
  if (t == 0) {
    double s = 0.0;
    for (int i = 0; i < k; ++i) s = __dadd_rn(s, w[i]);  // :154
    s_sum = s;
  }
  __syncthreads();
  const int e = s_err;
  for (int i = t; i < k; i += nt) w[i] = __ddiv_rn(w[i], s_sum);
  __syncthreads();
  if (t == 0) {
    int m = 0;
    for (int i = 0; i < k; ++i) slot[i] = (e == 0 && w[i] > 0) ? m++ : -1;  // voronoi/__init__.py:58 support > 0
    s_m = m;
  }
  __syncthreads();
  const int m = s_m;
  for (int i = t; i < k; i += nt) {
    if (slot[i] >= 0) {
      if (wts) wts[p * k + slot[i]] = w[i];
      if (ids_out) ids_out[p * k + slot[i]] = id[i];
    }
    if (i >= m) {
      if (wts) wts[p * k + i] = 0.0;
      if (ids_out) ids_out[p * k + i] = -1;
    }
  }
  const int rr = r < 8 ? r : 8;
  if (out && t < rr) {
    double acc = 0.0;
    for (int i = 0; i < k; ++i)
      if (slot[i] >= 0) acc = __dadd_rn(acc, id[i] >= 0 ? __dmul_rn(y[(int64_t)id[i] * r + t], w[i]) : CUDART_NAN);
    out[p * r + t] = acc;
  }
  if (t == 0) {
    if (cnt) cnt[p] = m;
    if (err) err[p] = e;
  }
}

// Remaining response columns (r > 8) for Voronoi: plain weighted sums from ids/wts.
__global__ void wsum_kernel(const int32_t* __restrict__ ids, const double* __restrict__ wts,
                            const int32_t* __restrict__ cnt, int64_t q, int stride,
                            const double* __restrict__ y, int r, int c0, double* __restrict__ out) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int rc = r - c0;
  if (e >= q * rc) return;
  const int64_t p = e / rc;
  const int c = c0 + (int)(e - p * rc);
  double s = 0.0;
  for (int i = 0; i < cnt[p]; ++i) {
    const int id = ids[p * stride + i];
    s = __dadd_rn(s, id >= 0 ? __dmul_rn(y[(int64_t)id * r + c], wts[p * stride + i]) : CUDART_NAN);
  }
  out[p * r + c] = s;
}

}  // namespace

// idx_dev (q,k) int32 neighbours, ascending distance.  For Voronoi the kept
// neighbours' global ids are compacted in place into m->idx32-compatible layout
// through ids_out = idx_dev is NOT done; callers read m->misc (q,k) instead.
int nn_weights(b200_model* m, const double* z_dev, int64_t q, int k, int method,
               const int32_t* idx_dev, const double* d2_dev, const double* y_dev, int r, double* wts_dev,
               int32_t* cnt_dev, int32_t* err_dev, double* out_dev, cudaStream_t st) {
  if (q == 0) return 0;
  if (method == B200_AVERAGE || method == B200_SHEPARD) {
    if (k > 128) {
      set_error("nn_weights: k=%d > 128 unsupported for Average/Shepard weighting", k);
      return 1;
    }
    const int64_t threads = q * 32;
    if (method == B200_SHEPARD && !d2_dev) {
      set_error("nn_weights: Shepard weighting needs the kNN squared distances");
      return 1;
    }
    double* outp = y_dev ? out_dev : nullptr;
    const unsigned tb = (unsigned)((q + 127) / 128);
    if (k <= 8 && q >= 4096)
      avg_shepard_thread_kernel<8><<<tb, 128, 0, st>>>(d2_dev, q, k, method, idx_dev, y_dev, r, wts_dev, cnt_dev, err_dev, outp);
    else if (k <= 16 && q >= 4096)
      avg_shepard_thread_kernel<16><<<tb, 128, 0, st>>>(d2_dev, q, k, method, idx_dev, y_dev, r, wts_dev, cnt_dev, err_dev, outp);
    else if (k <= 32 && q >= 4096)
      avg_shepard_thread_kernel<32><<<tb, 128, 0, st>>>(d2_dev, q, k, method, idx_dev, y_dev, r, wts_dev, cnt_dev, err_dev, outp);
    else
      avg_shepard_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(d2_dev, q, k, method, idx_dev, y_dev, r, wts_dev,
                                                                             cnt_dev, err_dev, outp);
    B200_CUDA(cudaGetLastError());
    m->st.launches += 1;
    return 0;
  }
  if (method == B200_VORONOI) {
    const int d = m->d;
    const int vthreads = k > 32 ? 256 : 128;
    const size_t smem = ((size_t)k * (d | 1) + (size_t)k * (k + 1) + 2 * (size_t)k + d + vthreads) * 8;  // + the partial maxima
    if (smem > 200 * 1024) {
      set_error("nn_weights: Voronoi on k=%d neighbours in %d-D needs %zu B shared memory", k, d, smem);
      return 1;
    }
    B200_TRY(m->misc.reserve((size_t)q * k * sizeof(int32_t)));
    B200_TRY(m->wts.reserve((size_t)q * k * sizeof(double)));
    B200_TRY(m->cnt.reserve((size_t)q * sizeof(int32_t)));
    double* wbuf = wts_dev ? wts_dev : m->wts.as<double>();
    int32_t* cbuf = cnt_dev ? cnt_dev : m->cnt.as<int32_t>();
    B200_CUDA(cudaFuncSetAttribute(voronoi_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    voronoi_kernel<<<(unsigned)q, vthreads, smem, st>>>(m->x, d, z_dev, q, k, idx_dev, y_dev, r, wbuf, cbuf,
                                                   err_dev, y_dev ? out_dev : nullptr,
                                                   m->misc.as<int32_t>());
    B200_CUDA(cudaGetLastError());
    m->st.launches += 1;
    if (y_dev && out_dev && r > 8) {
      const int64_t tot = q * (r - 8);
      wsum_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(m->misc.as<int32_t>(), wbuf, cbuf, q, k,
                                                                  y_dev, r, 8, out_dev);
      B200_CUDA(cudaGetLastError());
      m->st.launches += 1;
    }
    return 0;
  }
  set_error("nn_weights: unknown method %d", method);
  return 1;
}

}  // namespace b200
