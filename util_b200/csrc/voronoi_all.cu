// voronoi_all.cu -- stand-alone Voronoi(): cell weights over ALL fitted points.
//
// Replaces the per-query loop of Voronoi._predict (voronoi/__init__.py:41-62), one call of PREDICT
// (voronoi/voronoi.f90:57-155) per query against the n x n cache of inner products DOTS that _fit
// allocates (:32-38, MAKE_HUGE f90:50-55) and PREDICT fills lazily (:105-129).  Here the Gram matrix is
// computed once per model by a kernel and kept in HBM (n <= 4096: 128 MiB), and a block per query
// evaluates the closed form the reference's sorted / SKIP / CYCLE control flow prunes towards
// (SURVEY.md 3.4):
//   w_c = max(0, 1 - max_o max(0, num/den)),
//   num = (p.x_o + x_c.x_c) - (p.x_c + x_c.x_o),  den = x_o.x_o + x_c.x_c - 2 x_c.x_o
// with the operand grouping of voronoi.f90:131-134, then w /= SUM(w) (:154).  ERROR 1: a squared
// value overflows (:83-91); ERROR 2: |den| <= sqrt(eps) for some pair, i.e. duplicated fit points
// (:137,145-148).  fp64 CUDA-core work, n^2 divisions per query; the Gram rows stream from L2.
#include "common.cuh"

namespace b200 {

namespace {

constexpr double VA_HUGE = 1.7976931348623157e308;
constexpr double VA_MIN = 1.4901161193847656e-08;  // SQRT(EPSILON(1d0))

__global__ void voronoi_gram_kernel(const double* __restrict__ x, int n, int d, double* __restrict__ g,
                                    int* __restrict__ flag) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= (int64_t)n * n) return;
  const int a = (int)(e / n), b = (int)(e - (int64_t)a * n);
  double s = 0.0;
  for (int j = 0; j < d; ++j) s = __dadd_rn(s, __dmul_rn(x[(int64_t)a * d + j], x[(int64_t)b * d + j]));  // :120-128
  g[e] = s;
  if (s > VA_HUGE) atomicMax(flag, 1);
}

__global__ void __launch_bounds__(256)
voronoi_all_kernel(const double* __restrict__ x, int n, int d, const double* __restrict__ g,
                   const int* __restrict__ gflag, const double* __restrict__ z, int64_t q, int cap,
                   const double* __restrict__ y, int r, int32_t* __restrict__ ids, double* __restrict__ wts,
                   int32_t* __restrict__ cnt, int32_t* __restrict__ err, double* __restrict__ out) {
  extern __shared__ double sm[];
  double* pdot = sm;       // [n]  p . x_i
  double* w = pdot + n;    // [n]
  double* pts = w + n;     // [d]
  __shared__ int s_err;
  __shared__ double s_sum;
  const int64_t p = blockIdx.x;
  const int t = threadIdx.x, nt = blockDim.x;
  if (t == 0) s_err = *gflag;
  for (int j = t; j < d; j += nt) pts[j] = z[p * d + j];
  __syncthreads();
  for (int i = t; i < n; i += nt) {
    const double* xi = x + (int64_t)i * d;
    double s = 0.0, dd = 0.0;
    for (int j = 0; j < d; ++j) {
      s = __dadd_rn(s, __dmul_rn(pts[j], xi[j]));  // SUM(EVAL_PT * POINTS(:,I)), :131
      const double df = __dsub_rn(pts[j], xi[j]);  // :83-91 overflow check on the squared distance
      dd = __dadd_rn(dd, __dmul_rn(df, df));
    }
    pdot[i] = s;
    if (dd > VA_HUGE) atomicMax(&s_err, 1);
  }
  __syncthreads();
  // one WARP per centre c: lanes stride over the other points o, so a Gram row is read coalesced
  const int lane = t & 31, wid = t >> 5, nw = nt >> 5;
  for (int c = wid; c < n; c += nw) {
    const double* gc = g + (int64_t)c * n;
    const double gcc = gc[c], pc = pdot[c];
    double mr = 0.0;
    bool dup = false;
    for (int o = lane; o < n; o += 32) {
      if (o == c) continue;
      const double gco = gc[o];
      const double num = __dsub_rn(__dadd_rn(pdot[o], gcc), __dadd_rn(pc, gco));                      // :131-132
      const double den = __dsub_rn(__dadd_rn(g[(int64_t)o * n + o], gcc), __dmul_rn(2.0, gco));      // :134
      if (fabs(den) > VA_MIN) {
        const double ratio = __ddiv_rn(num, den);
        if (ratio > mr) mr = ratio;
      } else {
        dup = true;
      }
    }
    for (int s = 16; s > 0; s >>= 1) mr = fmax(mr, __shfl_xor_sync(0xffffffffu, mr, s));
    if (__any_sync(0xffffffffu, dup) && lane == 0) atomicMax(&s_err, 2);
    if (lane == 0) w[c] = fmax(__dsub_rn(1.0, mr), 0.0);  // :151
  }
  __syncthreads();
  if (t == 0) {
    double s = 0.0;
    for (int i = 0; i < n; ++i) s = __dadd_rn(s, w[i]);  // :154, index order
    s_sum = s;
    const int e = s_err;
    int m = 0;
    double acc[8];
    const int rr = r < 8 ? r : 8;
    for (int c = 0; c < rr; ++c) acc[c] = 0.0;
    for (int i = 0; i < n; ++i) {
      const double wi = __ddiv_rn(w[i], s);
      if (e == 0 && wi > 0) {  // voronoi/__init__.py:58 support > 0, ascending point index
        if (m < cap) {
          if (wts) wts[p * cap + m] = wi;
          if (ids) ids[p * cap + m] = i;
        }
        if (out)
          for (int c = 0; c < rr; ++c) acc[c] = __dadd_rn(acc[c], __dmul_rn(y[(int64_t)i * r + c], wi));  // base.py:169
        ++m;
      }
    }
    cnt[p] = m;
    err[p] = e;
    if (out)
      for (int c = 0; c < rr; ++c) out[p * r + c] = acc[c];
  }
  __syncthreads();
  // response columns beyond the first eight: one thread per column, same left-to-right sum
  if (out && r > 8 && s_err == 0) {
    const double s = s_sum;
    for (int c = 8 + t; c < r; c += nt) {
      double a = 0.0;
      for (int i = 0; i < n; ++i) {
        const double wi = __ddiv_rn(w[i], s);
        if (wi > 0) a = __dadd_rn(a, __dmul_rn(y[(int64_t)i * r + c], wi));
      }
      out[p * r + c] = a;
    }
  } else if (out && r > 8) {
    for (int c = 8 + t; c < r; c += nt) out[p * r + c] = 0.0;
  }
}

// The same evaluation for any n: p.x_i and w live in a per-block slice of global scratch instead of
// shared memory, the blocks are persistent (a few per SM, each walks its share of the queries), and the
// inner products x_c.x_o either stream from the n x n cache in HBM (20 GB at n = 50 000 -- what made the
// reference's stand-alone Voronoi impractical fits a B200) or, when no cache fits, are recomputed on the
// fly with the same dimension-order sums.  The non-zero weights are compacted in index order before the
// serial normalisation, so its length is the support size, not n (adding a zero weight changes no sum).
__global__ void __launch_bounds__(256)
voronoi_all_big_kernel(const double* __restrict__ x, int n, int d, const double* __restrict__ g,
                       const double* __restrict__ diag, const int* __restrict__ gflag, const double* __restrict__ z,
                       int64_t q, int cap, const double* __restrict__ y, int r, double* __restrict__ scratch,
                       int32_t* __restrict__ ids, double* __restrict__ wts, int32_t* __restrict__ cnt,
                       int32_t* __restrict__ err, double* __restrict__ out) {
  extern __shared__ double sm[];
  double* pts = sm;  // [d]
  __shared__ int s_err, s_total;
  __shared__ double s_sum;
  __shared__ int s_cnt[256];
  double* pdot = scratch + (size_t)blockIdx.x * 3 * n;  // [n]  p . x_i
  double* w = pdot + n;                                 // [n]
  int* order = reinterpret_cast<int*>(w + n);           // [n] indices of the non-zero weights, ascending
  const int t = threadIdx.x, nt = blockDim.x;
  const int lane = t & 31, wid = t >> 5, nw = nt >> 5;
  for (int64_t p = blockIdx.x; p < q; p += gridDim.x) {
    __syncthreads();
    if (t == 0) s_err = *gflag;
    for (int j = t; j < d; j += nt) pts[j] = z[p * d + j];
    __syncthreads();
    for (int i = t; i < n; i += nt) {
      const double* xi = x + (int64_t)i * d;
      double s = 0.0, dd = 0.0;
      for (int j = 0; j < d; ++j) {
        s = __dadd_rn(s, __dmul_rn(pts[j], xi[j]));  // :131
        const double df = __dsub_rn(pts[j], xi[j]);  // :83-91
        dd = __dadd_rn(dd, __dmul_rn(df, df));
      }
      pdot[i] = s;
      if (dd > VA_HUGE) atomicMax(&s_err, 1);
    }
    __syncthreads();
    for (int c = wid; c < n; c += nw) {
      const double* gc = g ? g + (int64_t)c * n : nullptr;
      const double* xc = x + (int64_t)c * d;
      const double gcc = diag[c], pc = pdot[c];
      double mr = 0.0;
      bool dup = false;
      for (int o = lane; o < n; o += 32) {
        if (o == c) continue;
        double gco;
        if (gc) {
          gco = gc[o];
        } else {
          const double* xo = x + (int64_t)o * d;
          gco = 0.0;
          for (int j = 0; j < d; ++j) gco = __dadd_rn(gco, __dmul_rn(xc[j], xo[j]));  // :120-128
        }
        const double num = __dsub_rn(__dadd_rn(pdot[o], gcc), __dadd_rn(pc, gco));   // :131-132
        const double den = __dsub_rn(__dadd_rn(diag[o], gcc), __dmul_rn(2.0, gco));  // :134
        if (fabs(den) > VA_MIN) {
          const double ratio = __ddiv_rn(num, den);
          if (ratio > mr) mr = ratio;
        } else {
          dup = true;
        }
      }
      for (int s = 16; s > 0; s >>= 1) mr = fmax(mr, __shfl_xor_sync(0xffffffffu, mr, s));
      if (__any_sync(0xffffffffu, dup) && lane == 0) atomicMax(&s_err, 2);
      if (lane == 0) w[c] = fmax(__dsub_rn(1.0, mr), 0.0);  // :151
    }
    __syncthreads();
    // ordered compaction of the non-zero weights: thread t owns the contiguous segment [t*seg, (t+1)*seg)
    const int seg = (n + nt - 1) / nt, lo = t * seg, hi = min(n, lo + seg);
    int mine = 0;
    for (int i = lo; i < hi; ++i) mine += w[i] > 0.0;
    s_cnt[t] = mine;
    __syncthreads();
    if (t == 0) {
      int run = 0;
      for (int i = 0; i < nt; ++i) {
        const int c = s_cnt[i];
        s_cnt[i] = run;
        run += c;
      }
      s_total = run;
    }
    __syncthreads();
    {
      int at = s_cnt[t];
      for (int i = lo; i < hi; ++i)
        if (w[i] > 0.0) order[at++] = i;
    }
    __syncthreads();
    const int S = s_total;
    if (t == 0) {
      double s = 0.0;
      for (int i = 0; i < S; ++i) s = __dadd_rn(s, w[order[i]]);  // :154, index order (zero weights add nothing)
      s_sum = s;
      const int e = s_err;
      int m = 0;
      double acc[8];
      const int rr = r < 8 ? r : 8;
      for (int c = 0; c < rr; ++c) acc[c] = 0.0;
      for (int k = 0; k < S; ++k) {
        const int i = order[k];
        const double wi = __ddiv_rn(w[i], s);
        if (e == 0 && wi > 0) {  // voronoi/__init__.py:58 support > 0, ascending point index
          if (m < cap) {
            if (wts) wts[p * cap + m] = wi;
            if (ids) ids[p * cap + m] = i;
          }
          if (out)
            for (int c = 0; c < rr; ++c) acc[c] = __dadd_rn(acc[c], __dmul_rn(y[(int64_t)i * r + c], wi));  // base.py:169
          ++m;
        }
      }
      cnt[p] = m;
      err[p] = e;
      if (out)
        for (int c = 0; c < rr; ++c) out[p * r + c] = acc[c];
    }
    __syncthreads();
    if (out && r > 8) {  // response columns beyond the first eight: one thread per column, same order
      const double s = s_sum;
      const bool ok = s_err == 0;
      for (int c = 8 + t; c < r; c += nt) {
        double a = 0.0;
        for (int k = 0; ok && k < S; ++k) {
          const int i = order[k];
          const double wi = __ddiv_rn(w[i], s);
          if (wi > 0) a = __dadd_rn(a, __dmul_rn(y[(int64_t)i * r + c], wi));
        }
        out[p * r + c] = a;
      }
    }
  }
}

__global__ void voronoi_diag_kernel(const double* __restrict__ x, int n, int d, double* __restrict__ diag,
                                    int* __restrict__ flag) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double s = 0.0;
  for (int j = 0; j < d; ++j) s = __dadd_rn(s, __dmul_rn(x[(int64_t)i * d + j], x[(int64_t)i * d + j]));  // :105-110
  diag[i] = s;
  if (s > VA_HUGE) atomicMax(flag, 1);
}

// EVAL_MESH (voronoi/voronoi.f90:5-44): the naive full evaluation from precomputed inner products,
// column-major as the Fortran sees them: ip (n,n), ap (q,n), mesh (q,n).  Block per approximation
// point, warp per interpolation point xi, lanes over the cell centres c; operand order of :26-29,
// |denominator| <= EPSILON skipped (:30), row sum in index order (:42).
__global__ void __launch_bounds__(256)
eval_mesh_kernel(const double* __restrict__ ip, int n, const double* __restrict__ ap, int64_t q,
                 double* __restrict__ mesh) {
  extern __shared__ double sm[];
  double* a = sm;        // [n] app_int_prods(y, :)
  double* w = a + n;     // [n]
  __shared__ double s_sum;
  const int64_t y = blockIdx.x;
  const int t = threadIdx.x, nt = blockDim.x;
  for (int c = t; c < n; c += nt) a[c] = ap[y + (int64_t)c * q];
  __syncthreads();
  const int lane = t & 31, wid = t >> 5, nw = nt >> 5;
  for (int xi = wid; xi < n; xi += nw) {
    const double gxx = ip[xi + (int64_t)xi * n], axi = a[xi];
    double mr = 0.0;
    for (int c = lane; c < n; c += 32) {
      if (c == xi) continue;
      const double gxc = ip[xi + (int64_t)c * n];
      const double num = __dadd_rn(__dsub_rn(__dsub_rn(a[c], axi), gxc), gxx);
      const double den = __dadd_rn(__dsub_rn(__dsub_rn(ip[c + (int64_t)c * n], ip[c + (int64_t)xi * n]), gxc), gxx);
      if (fabs(den) > 2.220446049250313e-16) {
        const double ratio = __ddiv_rn(num, den);
        if (ratio > mr) mr = ratio;
      }
    }
    for (int s = 16; s > 0; s >>= 1) mr = fmax(mr, __shfl_xor_sync(0xffffffffu, mr, s));
    if (lane == 0) w[xi] = fmax(__dsub_rn(1.0, mr), 0.0);
  }
  __syncthreads();
  if (t == 0) {
    double s = 0.0;
    for (int i = 0; i < n; ++i) s = __dadd_rn(s, w[i]);
    s_sum = s;
  }
  __syncthreads();
  const double s = s_sum;
  for (int xi = t; xi < n; xi += nt) mesh[y + (int64_t)xi * q] = __ddiv_rn(w[xi], s);
}

}  // namespace

int eval_mesh(const double* ip_dev, int n, const double* ap_dev, int64_t q, double* mesh_dev, cudaStream_t st) {
  if (q == 0 || n == 0) return 0;
  const size_t smem = (size_t)2 * n * 8;
  if (smem > 200 * 1024) {
    set_error("eval_mesh: n=%d interpolation points exceed the shared-memory row buffers", n);
    return 1;
  }
  B200_CUDA(cudaFuncSetAttribute(eval_mesh_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  eval_mesh_kernel<<<(unsigned)q, 256, smem, st>>>(ip_dev, n, ap_dev, q, mesh_dev);
  B200_CUDA(cudaGetLastError());
  return 0;
}

int voronoi_all(b200_model* m, const double* z_dev, int64_t q, int cap, const double* y_dev, int r,
                int32_t* ids_dev, double* wts_dev, int32_t* cnt_dev, int32_t* err_dev, double* out_dev,
                cudaStream_t st) {
  if (q == 0) return 0;
  const int64_t n = m->n;
  const int d = m->d;
  if (n > 4096) {
    // any n: Gram cache in HBM while it fits (<= 40 GB), else inner products on the fly
    const bool cache = (double)n * (double)n * 8.0 <= 40e9;
    if (!m->vor_ready) {
      B200_TRY(m->vor_g.reserve((cache ? (size_t)n * n * 8 : 0) + (size_t)n * 8 + 64));
      double* diag = reinterpret_cast<double*>(m->vor_g.as<char>() + (cache ? (size_t)n * n * 8 : 0));
      int* flag = reinterpret_cast<int*>(diag + n);
      B200_CUDA(cudaMemsetAsync(flag, 0, 4, st));
      if (cache) {
        const int64_t tot = n * n;
        voronoi_gram_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(m->x, (int)n, d, m->vor_g.as<double>(), flag);
      }
      voronoi_diag_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(m->x, (int)n, d, diag, flag);
      B200_CUDA(cudaGetLastError());
      m->st.launches += 2;
      m->vor_ready = true;
    }
    double* diag = reinterpret_cast<double*>(m->vor_g.as<char>() + (cache ? (size_t)n * n * 8 : 0));
    const int* flag = reinterpret_cast<const int*>(diag + n);
    const int blocks = (int)std::min<int64_t>(q, (int64_t)m->sm_count * 4);
    B200_TRY(m->misc.reserve((size_t)blocks * 3 * n * 8 + 64));
    voronoi_all_big_kernel<<<blocks, 256, (size_t)d * 8, st>>>(m->x, (int)n, d, cache ? m->vor_g.as<double>() : nullptr, diag,
                                                              flag, z_dev, q, cap, y_dev, r > 0 ? r : 1,
                                                              m->misc.as<double>(), ids_dev, wts_dev, cnt_dev, err_dev,
                                                              y_dev ? out_dev : nullptr);
    B200_CUDA(cudaGetLastError());
    m->st.launches += 1;
    return 0;
  }
  if (!m->vor_ready) {
    B200_TRY(m->vor_g.reserve((size_t)n * n * 8 + 64));
    int* gflag = reinterpret_cast<int*>(m->vor_g.as<char>() + (size_t)n * n * 8);
    B200_CUDA(cudaMemsetAsync(gflag, 0, 4, st));
    const int64_t tot = n * n;
    voronoi_gram_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(m->x, (int)n, d, m->vor_g.as<double>(), gflag);
    B200_CUDA(cudaGetLastError());
    m->st.launches += 1;
    m->vor_ready = true;
  }
  const int* flag = reinterpret_cast<const int*>(m->vor_g.as<char>() + (size_t)n * n * 8);
  const size_t smem = ((size_t)2 * n + d) * 8;
  B200_CUDA(cudaFuncSetAttribute(voronoi_all_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  voronoi_all_kernel<<<(unsigned)q, 256, smem, st>>>(m->x, (int)n, d, m->vor_g.as<double>(), flag, z_dev, q, cap, y_dev,
                                                     r > 0 ? r : 1, ids_dev, wts_dev, cnt_dev, err_dev,
                                                     y_dev ? out_dev : nullptr);
  B200_CUDA(cudaGetLastError());
  m->st.launches += 1;
  return 0;
}

}  // namespace b200
