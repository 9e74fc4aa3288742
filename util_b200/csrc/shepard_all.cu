// shepard_all.cu -- stand-alone Shepard(): inverse-distance-squared weights over ALL fitted points.
//
// Replaces the per-query loop of Shepard._predict (shepard/__init__.py:35-47), one call of
// WEIGHT (shepard/weight.f90:1-34) per query, for the whole batch:
//   d_i = NORM2(p - x_i);  all d < EPS -> equal weights;  some d < EPS -> equal share among those;
//   else w_i = 1/d_i^2;  w /= SUM(w)                                   (EPS = sqrt(epsilon))
// The reference returns the dense (q, n) weight matrix.  Here pass 1 (one thread per query,
// fitted rows streamed through shared memory in index order) reduces each query to its
// normaliser and, when y is given, the weighted sum; pass 2 writes the dense weights only if the
// caller asks for them.  fp64 CUDA-core work, 3d+~25 operations per pair.
#include "common.cuh"

namespace b200 {

namespace {

constexpr int SA_QB = 128;
constexpr double SA_EPS = 1.4901161193847656e-08;  // SQRT(EPSILON(1d0)), weight.f90:17

struct SaNorm {
  double s;      // SUM of raw weights (1/d^2, or the count of d < EPS)
  int32_t mode;  // 0: inverse-distance, 1: exact/near hits share equally, 2: everything coincides
};

template <int DMAX>
__global__ void __launch_bounds__(SA_QB)
shepard_all_kernel(const double* __restrict__ x, int64_t n, int d, const double* __restrict__ z, int64_t q,
                   int tile_pts, const double* __restrict__ y, int r, SaNorm* __restrict__ norm,
                   double* __restrict__ out) {
  extern __shared__ double sm[];
  double* xs = sm;                          // [tile][d]
  double* ys = xs + (size_t)tile_pts * d;   // [tile]  (first response column)
  const int t = threadIdx.x;
  const int64_t p = (int64_t)blockIdx.x * SA_QB + t;
  const bool live = p < q;
  double zq[DMAX];
#pragma unroll
  for (int j = 0; j < DMAX; ++j) zq[j] = (live && j < d) ? z[p * d + j] : 0.0;
  double sw = 0.0, swy = 0.0, ne = 0.0, sey = 0.0, mx = 0.0, sy = 0.0;
  for (int64_t base = 0; base < n; base += tile_pts) {
    const int tc = (int)imin64(tile_pts, n - base);
    __syncthreads();
    for (int e = t; e < tc * d; e += SA_QB) xs[e] = x[base * d + e];
    if (y)
      for (int e = t; e < tc; e += SA_QB) ys[e] = y[(base + e) * r];
    __syncthreads();
    if (!live) continue;
    for (int i = 0; i < tc; ++i) {
      const double* xr = xs + (size_t)i * d;
      double a = 0.0;
#pragma unroll
      for (int j = 0; j < DMAX; ++j)
        if (j < d) {
          const double df = __dsub_rn(xr[j], zq[j]);  // PTS(:,I) - PT(:), weight.f90:14
          a = __dadd_rn(a, __dmul_rn(df, df));
        }
      const double dist = sqrt(a);
      const double yi = y ? ys[i] : 0.0;
      mx = fmax(mx, dist);
      sy = __dadd_rn(sy, yi);
      if (dist < SA_EPS) {  // :25-28
        ne += 1.0;
        sey = __dadd_rn(sey, yi);
      } else {
        const double w = __ddiv_rn(1.0, __dmul_rn(dist, dist));  // :31
        sw = __dadd_rn(sw, w);
        swy = __dadd_rn(swy, __dmul_rn(yi, w));
      }
    }
  }
  if (!live) return;
  SaNorm nm;
  double val;
  if (mx < SA_EPS) {  // :22-24
    nm.mode = 2;
    nm.s = (double)n;
    val = sy / (double)n;
  } else if (ne > 0.0) {
    nm.mode = 1;
    nm.s = ne;
    val = sey / ne;
  } else {
    nm.mode = 0;
    nm.s = sw;
    val = swy / sw;  // sum(y_i * (w_i / S)) up to the order of the roundings
  }
  norm[p] = nm;
  if (out) out[p * r] = val;
}

// dense weights (q, n): w_i / S with the branch decided in pass 1
__global__ void shepard_all_weights_kernel(const double* __restrict__ x, int64_t n, int d,
                                           const double* __restrict__ z, int64_t q,
                                           const SaNorm* __restrict__ norm, double* __restrict__ wts) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= q * n) return;
  const int64_t p = e / n, i = e - p * n;
  const SaNorm nm = norm[p];
  const double dist = sqrt(sqdist_seq(x + i * d, z + p * d, d));
  double w;
  if (nm.mode == 2) w = 1.0;
  else if (nm.mode == 1) w = dist < SA_EPS ? 1.0 : 0.0;
  else w = __ddiv_rn(1.0, __dmul_rn(dist, dist));
  wts[e] = __ddiv_rn(w, nm.s);  // :34
}

// remaining response columns from the dense definition (r > 1 is rare)
__global__ void shepard_all_cols_kernel(const double* __restrict__ x, int64_t n, int d,
                                        const double* __restrict__ z, int64_t q,
                                        const SaNorm* __restrict__ norm, const double* __restrict__ y, int r,
                                        double* __restrict__ out) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= q * (r - 1)) return;
  const int64_t p = e / (r - 1);
  const int c = 1 + (int)(e - p * (r - 1));
  const SaNorm nm = norm[p];
  double acc = 0.0;
  for (int64_t i = 0; i < n; ++i) {
    const double dist = sqrt(sqdist_seq(x + i * d, z + p * d, d));
    double w;
    if (nm.mode == 2) w = 1.0;
    else if (nm.mode == 1) w = dist < SA_EPS ? 1.0 : 0.0;
    else w = __ddiv_rn(1.0, __dmul_rn(dist, dist));
    acc = __dadd_rn(acc, __dmul_rn(y[i * r + c], __ddiv_rn(w, nm.s)));
  }
  out[p * r + c] = acc;
}

}  // namespace

int shepard_all(b200_model* m, const double* z_dev, int64_t q, const double* y_dev, int r, double* wts_dev,
                double* out_dev, cudaStream_t st) {
  if (q == 0) return 0;
  const int d = m->d;
  if (d > 64) {
    set_error("shepard_all: dimension %d > 64 unsupported", d);
    return 1;
  }
  B200_TRY(m->cand_meta.reserve((size_t)q * sizeof(SaNorm) + 64));
  SaNorm* norm = m->cand_meta.as<SaNorm>();
  int tile_pts = (int)((96 * 1024) / ((size_t)(d + 1) * 8));
  if (tile_pts > 512) tile_pts = 512;
  const size_t smem = (size_t)tile_pts * (d + 1) * 8;
  const unsigned gb = (unsigned)((q + SA_QB - 1) / SA_QB);
#define SA_LAUNCH(DM)                                                                              \
  do {                                                                                             \
    auto kern = shepard_all_kernel<DM>;                                                            \
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    kern<<<gb, SA_QB, smem, st>>>(m->x, m->n, d, z_dev, q, tile_pts, y_dev, r > 0 ? r : 1, norm,    \
                                  y_dev ? out_dev : nullptr);                                      \
  } while (0)
  if (d <= 2) SA_LAUNCH(2);
  else if (d <= 4) SA_LAUNCH(4);
  else if (d <= 8) SA_LAUNCH(8);
  else if (d <= 16) SA_LAUNCH(16);
  else if (d <= 32) SA_LAUNCH(32);
  else SA_LAUNCH(64);
#undef SA_LAUNCH
  B200_CUDA(cudaGetLastError());
  m->st.launches += 1;
  if (y_dev && r > 1) {
    const int64_t tot = q * (r - 1);
    shepard_all_cols_kernel<<<(unsigned)((tot + 127) / 128), 128, 0, st>>>(m->x, m->n, d, z_dev, q, norm, y_dev, r,
                                                                          out_dev);
    B200_CUDA(cudaGetLastError());
    m->st.launches += 1;
  }
  if (wts_dev) {
    const int64_t tot = q * m->n;
    shepard_all_weights_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(m->x, m->n, d, z_dev, q, norm, wts_dev);
    B200_CUDA(cudaGetLastError());
    m->st.launches += 1;
  }
  return 0;
}

}  // namespace b200
