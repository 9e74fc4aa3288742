// lshep.cu -- linear Shepard (LSHEP) fit and predict.
//
// Fit  = LSHEP, non-robust branch (linear_shepard.f95:91-134,177; the Python
//        wrapper never passes RLSHEP, linear_shepard/__init__.py:36,53):
//          1. self-excluding kNN with NP = min(N-1,(3M+1)/2) neighbours (:98-115)
//          2. per node a weighted NP x M least-squares problem (:116-134).  The
//             reference calls LAPACK DGELSS (SVD, minimum norm); here it is a
//             batched Householder QR with column pivoting, one thread per node,
//             the systems stored [element][node] in scratch so loads coalesce.
//             Full-rank systems agree with DGELSS to rounding.  A node whose pivots
//             fall (near) the rank threshold is solved again the way DGELSS does it:
//             a one-sided Jacobi SVD (one warp per such node), singular values
//             <= NP*eps*sigma_max dropped, minimum-norm solution, IER=2 when the
//             rank is short (:132-135, delaunay/lapack.f:2390-2420).
//          3. DIAM = max pairwise squared distance (:114) only enters through
//             RW = min(sqrt(DIAM)/2, RW) (:177).  A lower bound L <= DIAM from two
//             farthest-point sweeps proves the clamp inactive when
//             max RW <= sqrt(L)/2; otherwise the exact O(N^2) maximum is taken.
// Eval = LSHEPVAL over the query batch (linear_shepard.f95:365-472): one thread
//        per query, nodes stream through shared memory in index order.  With the
//        wrapper's radius=2 thousands of nodes contribute per query, so this is
//        a dense fp64 CUDA-core pass: sum(W), sum(W*nodal value), the weighted
//        slope; a second pass applies the slope-angle test (:460-469).
#include <math_constants.h>

#include "common.cuh"

namespace b200 {

namespace {

// ---------------------------------------------------------------- fit kernels
__global__ void lshep_build_kernel(const double* __restrict__ x, int64_t n, int d, int np,
                                   const int32_t* __restrict__ idx, const double* __restrict__ d2,
                                   const double* __restrict__ f, double* __restrict__ W,
                                   double* __restrict__ rw) {
  // W: [(np*(d+1))][n]; element (i, j) of node k at W[(j*np + i)*n + k], j = d is the rhs
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const double rk = sqrt(d2[k * np + np - 1]);  // :116,:121
  rw[k] = rk;
  const double rp = __dmul_rn(1.1, rk);  // :125
  const double fk = f[k];
  for (int i = 0; i < np; ++i) {
    const int nb = idx[k * np + i];
    const double dist = sqrt(d2[k * np + i]);
    const double om = __ddiv_rn(__dsub_rn(rp, dist), __dmul_rn(rp, dist));  // :126
    for (int j = 0; j < d; ++j)
      W[((size_t)j * np + i) * n + k] = __dmul_rn(om, __dsub_rn(x[(int64_t)nb * d + j], x[k * d + j]));  // :118-120,:129-131
    W[((size_t)d * np + i) * n + k] = __dmul_rn(om, __dsub_rn(f[nb], fk));  // :117,:128
  }
}

// Householder QR with column pivoting, one thread per node, in place on W.
// a (n,d) gets the solution; rank (n) the numerical rank.
// Nodes whose smallest pivot is within 1e-7 of the first one are appended to `sus` and solved
// again by lshep_svd_kernel (which also decides their rank the way DGELSS does).
template <int DMAX>
__global__ void lshep_qr_kernel(double* __restrict__ W, int64_t n, int d, int np,
                                double* __restrict__ a, int32_t* __restrict__ rank_out,
                                int32_t* __restrict__ sus, int32_t* __restrict__ sus_cnt) {
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
#define A_(i, j) W[((size_t)(j) * np + (i)) * n + k]
  int perm[DMAX];
  double cn[DMAX];
  for (int j = 0; j < d; ++j) {
    perm[j] = j;
    double s = 0.0;
    for (int i = 0; i < np; ++i) s += A_(i, j) * A_(i, j);
    cn[j] = s;
  }
  int rank = d;
  double r00 = 0.0, rmin = CUDART_INF;
  const double tol = (double)np * 2.220446049250313e-16;  // RCOND, :100
  for (int c = 0; c < d; ++c) {
    // pivot: remaining column with the largest norm (recomputed for accuracy)
    int pv = c;
    double best = -1.0;
    for (int j = c; j < d; ++j) {
      double s = 0.0;
      for (int i = c; i < np; ++i) s += A_(i, j) * A_(i, j);
      cn[j] = s;
      if (s > best) { best = s; pv = j; }
    }
    if (pv != c) {
      for (int i = 0; i < np; ++i) { const double tmp = A_(i, c); A_(i, c) = A_(i, pv); A_(i, pv) = tmp; }
      const int tp = perm[c]; perm[c] = perm[pv]; perm[pv] = tp;
    }
    const double nrm = sqrt(best);
    if (c == 0) r00 = nrm;
    rmin = fmin(rmin, nrm);
    if (!(nrm > tol * r00) || nrm == 0.0) { rank = c; break; }
    const double a0 = A_(c, c);
    const double alpha = a0 >= 0 ? -nrm : nrm;
    const double v0 = a0 - alpha;
    // v = (v0, A(c+1.., c)); beta = 2 / (v.v)
    double vv = v0 * v0;
    for (int i = c + 1; i < np; ++i) vv += A_(i, c) * A_(i, c);
    const double beta = vv > 0 ? 2.0 / vv : 0.0;
    for (int j = c + 1; j <= d; ++j) {  // includes the rhs column j = d
      double s = v0 * A_(c, j);
      for (int i = c + 1; i < np; ++i) s += A_(i, c) * A_(i, j);
      s *= beta;
      A_(c, j) -= s * v0;
      for (int i = c + 1; i < np; ++i) A_(i, j) -= s * A_(i, c);
    }
    A_(c, c) = alpha;
  }
  // back substitution on the leading rank x rank block
  double sol[DMAX];
  for (int j = 0; j < d; ++j) sol[j] = 0.0;
  for (int c = rank - 1; c >= 0; --c) {
    double s = A_(c, d);
    for (int j = c + 1; j < rank; ++j) s -= A_(c, j) * sol[j];
    sol[c] = s / A_(c, c);
  }
  for (int j = 0; j < d; ++j) a[k * d + perm[j]] = sol[j];  // :134
  if (rank_out) rank_out[k] = rank;
  if (rank < d || !(rmin > 1e-7 * r00)) sus[atomicAdd(sus_cnt, 1)] = (int32_t)k;
#undef A_
}

// Minimum-norm least squares the way DGELSS defines it, by one warp on a system in shared memory:
// A = [d+1][np] (column j at A + j*np, column d the right-hand side; destroyed), V = [d][d] work,
// wk = [2d] work.  One-sided Jacobi rotations orthogonalise the columns (A V = U S); then
// x = sum_j (a_j . b / s_j^2) v_j over the singular values s_j > NP*eps*s_max (RCOND = NP*eps,
// linear_shepard.f95:100; the cut of delaunay/lapack.f:2390-2400).  sol[d] receives x; returns the rank.
__device__ int warp_gelss(double* A, double* V, double* wk, double* sol, int np, int d, int lane) {
  for (int e = lane; e < d * d; e += 32) V[e] = (e / d == e % d) ? 1.0 : 0.0;
  __syncwarp();
  for (int sweep = 0; sweep < 40; ++sweep) {
    int rotated = 0;
    for (int p = 0; p + 1 < d; ++p)
      for (int q = p + 1; q < d; ++q) {
        double al = 0.0, be = 0.0, ga = 0.0;
        for (int i = lane; i < np; i += 32) {
          const double ap = A[(size_t)p * np + i], aq = A[(size_t)q * np + i];
          al = fma(ap, ap, al);
          be = fma(aq, aq, be);
          ga = fma(ap, aq, ga);
        }
        for (int o = 16; o > 0; o >>= 1) {
          al += __shfl_xor_sync(0xffffffffu, al, o);
          be += __shfl_xor_sync(0xffffffffu, be, o);
          ga += __shfl_xor_sync(0xffffffffu, ga, o);
        }
        if (!(fabs(ga) > 1e-15 * sqrt(al * be)) || ga == 0.0) continue;  // already orthogonal (warp-uniform)
        ++rotated;
        const double zeta = (be - al) / (2.0 * ga);
        const double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
        const double c = 1.0 / sqrt(1.0 + t * t), sn = c * t;
        for (int i = lane; i < np; i += 32) {
          const double ap = A[(size_t)p * np + i], aq = A[(size_t)q * np + i];
          A[(size_t)p * np + i] = c * ap - sn * aq;
          A[(size_t)q * np + i] = sn * ap + c * aq;
        }
        for (int i = lane; i < d; i += 32) {
          const double vp = V[(size_t)p * d + i], vq = V[(size_t)q * d + i];
          V[(size_t)p * d + i] = c * vp - sn * vq;
          V[(size_t)q * d + i] = sn * vp + c * vq;
        }
        __syncwarp();
      }
    if (!rotated) break;
  }
  double* coef = wk;       // [d] a_j . b
  double* sig2 = wk + d;   // [d] s_j^2
  for (int j = 0; j < d; ++j) {
    double s2 = 0.0, ab = 0.0;
    for (int i = lane; i < np; i += 32) {
      const double aj = A[(size_t)j * np + i];
      s2 = fma(aj, aj, s2);
      ab = fma(aj, A[(size_t)d * np + i], ab);
    }
    for (int o = 16; o > 0; o >>= 1) {
      s2 += __shfl_xor_sync(0xffffffffu, s2, o);
      ab += __shfl_xor_sync(0xffffffffu, ab, o);
    }
    if (lane == 0) {
      sig2[j] = s2;
      coef[j] = ab;
    }
  }
  __syncwarp();
  double smax2 = 0.0;
  for (int j = 0; j < d; ++j) smax2 = fmax(smax2, sig2[j]);
  const double rcond = (double)np * 2.220446049250313e-16;
  const double thr = fmax(rcond * sqrt(smax2), 2.2250738585072014e-308);  // max(RCOND*S(1), SFMIN)
  int rank = 0;
  for (int j = 0; j < d; ++j) rank += sqrt(sig2[j]) > thr ? 1 : 0;
  for (int i = lane; i < d; i += 32) {
    double xi = 0.0;
    for (int j = 0; j < d; ++j)
      if (sqrt(sig2[j]) > thr) xi = fma(coef[j] / sig2[j], V[(size_t)j * d + i], xi);
    sol[i] = xi;
  }
  __syncwarp();
  return rank;
}

// DGELSS semantics for the nodes listed in `sus`: one warp per node.  The node's weighted system is
// rebuilt in shared memory and solved by warp_gelss.  Rank short of d raises the fit's IER to 2 (:135).
__global__ void __launch_bounds__(32)
lshep_svd_kernel(const double* __restrict__ x, int64_t n, int d, int np, const int32_t* __restrict__ idx,
                 const double* __restrict__ d2, const double* __restrict__ f, const int32_t* __restrict__ sus,
                 double* __restrict__ a, int32_t* __restrict__ flag) {
  extern __shared__ double sm[];
  double* A = sm;                         // [d+1][np]
  double* V = sm + (size_t)(d + 1) * np;  // [d][d]
  double* wk = V + (size_t)d * d;         // [2d]
  double* sol = wk + 2 * d;               // [d]
  const int lane = threadIdx.x;
  const int64_t k = sus[blockIdx.x];
  const double rk = sqrt(d2[k * np + np - 1]);
  const double rp = __dmul_rn(1.1, rk);
  const double fk = f[k];
  for (int i = lane; i < np; i += 32) {  // the same arithmetic as lshep_build_kernel
    const int nb = idx[k * np + i];
    const double dist = sqrt(d2[k * np + i]);
    const double om = __ddiv_rn(__dsub_rn(rp, dist), __dmul_rn(rp, dist));
    for (int j = 0; j < d; ++j) A[(size_t)j * np + i] = __dmul_rn(om, __dsub_rn(x[(int64_t)nb * d + j], x[k * d + j]));
    A[(size_t)d * np + i] = __dmul_rn(om, __dsub_rn(f[nb], fk));
  }
  __syncwarp();
  const int rank = warp_gelss(A, V, wk, sol, np, d, lane);
  for (int i = lane; i < d; i += 32) a[k * d + i] = sol[i];
  if (lane == 0 && rank < d) atomicMax(flag, 2);  // :135
}

// ---- robust fit (LSHEP with RLSHEP present, linear_shepard.f95:136-175; IRLS :181-282, MAD :284-308,
// QUANTILE :310-335).  One warp per node, everything in shared memory.
// k-th smallest (1-based) of v[0..n): the value with #(< it) < k <= #(<= it) -- what QUANTILE's
// recursion returns.
__device__ double warp_kth(const double* v, int n, int k, int lane) {
  double found = 0.0;
  int have = 0;
  for (int i = lane; i < n; i += 32) {
    const double x = v[i];
    int lt = 0, le = 0;
    for (int j = 0; j < n; ++j) {
      lt += v[j] < x;
      le += v[j] <= x;
    }
    if (lt < k && k <= le) {
      found = x;
      have = 1;
    }
  }
  // every lane that found it holds the same value: take the lowest such lane
  const unsigned m = __ballot_sync(0xffffffffu, have);
  return __shfl_sync(0xffffffffu, found, __ffs(m) - 1);
}
__device__ double warp_mad(const double* res, double* tmp, int np, int lane) {
  const double factor = __ddiv_rn(1.0, 0.6745);  // :297
  const int nmid = np / 2;
  double med;
  if (np % 2 == 0) med = __ddiv_rn(__dadd_rn(warp_kth(res, np, nmid, lane), warp_kth(res, np, nmid + 1, lane)), 2.0);
  else med = warp_kth(res, np, nmid + 1, lane);
  for (int i = lane; i < np; i += 32) tmp[i] = fabs(__dsub_rn(res[i], med));
  __syncwarp();
  double r;
  if (np % 2 == 0)
    r = __ddiv_rn(__dmul_rn(factor, __dadd_rn(warp_kth(tmp, np, nmid, lane), warp_kth(tmp, np, nmid + 1, lane))), 2.0);
  else
    r = __dmul_rn(factor, warp_kth(tmp, np, nmid + 1, lane));
  __syncwarp();
  return r;
}
// residuals RES = ALPHA_IN . sol - BETA_IN (:137, :272), row by row in dimension order
__device__ void warp_residuals(const double* Ain, const double* bin, const double* sol, double* res, int np, int d,
                               int lane) {
  for (int i = lane; i < np; i += 32) {
    double v = 0.0;
    for (int j = 0; j < d; ++j) v = __dadd_rn(v, __dmul_rn(Ain[(size_t)j * np + i], sol[j]));
    res[i] = __dsub_rn(v, bin[i]);
  }
  __syncwarp();
}
__device__ double warp_bisq_objective(const double* res, double s, int np) {  // :243, :275 (sequential SUM)
  double e = 0.0;
  for (int j = 0; j < np; ++j) {
    const double t = __ddiv_rn(res[j], s);
    const double w = fmax(0.0, __dsub_rn(1.0, __dmul_rn(t, t)));
    e = __dadd_rn(e, __dsub_rn(1.0, __dmul_rn(__dmul_rn(w, w), w)));
  }
  return e;
}
// IRLS (:181-282): returns IER 0 / 1 / 2; sol and res updated, omega = final weights, *s scaled (:237).
__device__ int warp_irls(double* A, double* V, double* wk, double* sol, const double* Ain, const double* bin,
                         double* res, double* omega, double* s, int func, int np, int d, int lane) {
  *s = __dmul_rn(func == 1 ? 1.0 : 3.0, *s);  // :231, :237
  double err_old = 0.0;
  if (func == 2) err_old = warp_bisq_objective(res, *s, np);
  for (int it = 0; it < 5; ++it) {
    for (int j = lane; j < np; j += 32) {
      double w;
      if (func == 1) w = res[j] == 0.0 ? 1.0 : sqrt(fmin(1.0, __ddiv_rn(*s, fabs(res[j]))));  // :249-256
      else {
        const double t = __ddiv_rn(res[j], *s);
        w = fmax(0.0, __dsub_rn(1.0, __dmul_rn(t, t)));  // :258-259
      }
      omega[j] = w;
      for (int c = 0; c < d; ++c) A[(size_t)c * np + j] = __dmul_rn(w, Ain[(size_t)c * np + j]);  // :262-265
      A[(size_t)d * np + j] = __dmul_rn(w, bin[j]);
    }
    __syncwarp();
    const int rank = warp_gelss(A, V, wk, sol, np, d, lane);  // :266-267
    if (rank < d) return 1;                                   // :268-271
    warp_residuals(Ain, bin, sol, res, np, d, lane);          // :272
  }
  if (func == 2 && warp_bisq_objective(res, *s, np) > err_old) return 2;  // :274-280
  return 0;
}

__global__ void __launch_bounds__(32)
lshep_robust_kernel(const double* __restrict__ x, int64_t n, int d, int np, const int32_t* __restrict__ idx,
                    const double* __restrict__ d2, const double* __restrict__ f, double* __restrict__ a,
                    double* __restrict__ rw, unsigned long long* __restrict__ last_ier) {
  extern __shared__ double sm[];
  double* A = sm;                              // [d+1][np] weighted system (work)
  double* V = A + (size_t)(d + 1) * np;        // [d][d]
  double* wk = V + (size_t)d * d;              // [2d]
  double* sol = wk + 2 * d;                    // [d]
  double* Ain = sol + d;                       // [d][np] ALPHA_IN
  double* bin = Ain + (size_t)d * np;          // [np] BETA_IN
  double* res = bin + np;                      // [np]
  double* omega = res + np;                    // [np]
  double* dist = omega + np;                   // [np]
  double* tmp = dist + np;                     // [np]
  const int lane = threadIdx.x;
  const int64_t k = blockIdx.x;
  const double fk = f[k];
  for (int i = lane; i < np; i += 32) {
    const int nb = idx[k * np + i];
    dist[i] = sqrt(d2[k * np + i]);                                                    // :116
    bin[i] = __dsub_rn(f[nb], fk);                                                     // :117
    for (int j = 0; j < d; ++j) {
      const double v = __dsub_rn(x[(int64_t)nb * d + j], x[k * d + j]);                // :118-120
      Ain[(size_t)j * np + i] = v;
      A[(size_t)j * np + i] = v;                                                       // OMEGA = 1 (:122-123)
    }
    A[(size_t)d * np + i] = bin[i];
    omega[i] = 1.0;
  }
  __syncwarp();
  double rk = dist[np - 1];                                                            // :121
  int rank = warp_gelss(A, V, wk, sol, np, d, lane);                                   // :132
  for (int j = lane; j < d; j += 32) a[k * d + j] = sol[j];                            // :134
  // IER is a plain scalar the node loop overwrites (:135, :145, :157): the code set by the highest node wins
  if (lane == 0 && rank < d) atomicMax(last_ier, ((unsigned long long)(k + 1) << 2) | 2ull);
  warp_residuals(Ain, bin, sol, res, np, d, lane);                                     // :137
  double mar = warp_mad(res, tmp, np, lane);                                           // :138
  bool failed = false;
  if (mar > 2.220446049250313e-16) {                                                   // :139
    int e1 = warp_irls(A, V, wk, sol, Ain, bin, res, omega, &mar, 1, np, d, lane);     // :140-143
    if (e1 != 0) {
      failed = true;                                                                   // :144-146
    } else {
      for (int j = lane; j < d; j += 32) a[k * d + j] = sol[j];                        // :148
      mar = warp_mad(res, tmp, np, lane);                                              // :150
      if (mar > 2.220446049250313e-16) {                                               // :151
        e1 = warp_irls(A, V, wk, sol, Ain, bin, res, omega, &mar, 2, np, d, lane);     // :152-155
        if (e1 != 0) failed = true;                                                    // :156-158
        else
          for (int j = lane; j < d; j += 32) a[k * d + j] = sol[j];                    // :160
      }
    }
  }
  if (failed) {
    if (lane == 0) atomicMax(last_ier, ((unsigned long long)(k + 1) << 2) | 3ull);
  } else {
    for (int i = 0; i < np; ++i) {                                                     // :164-174
      if (omega[i] < 0.8) {
        rk = i == 0 ? __ddiv_rn(dist[0], 2.0) : __ddiv_rn(__dadd_rn(dist[i - 1], dist[i]), 2.0);
        break;
      }
    }
  }
  if (lane == 0) rw[k] = rk;
}

// ---- RIPPLE (linear_shepard.f95:474-665): one block per node.
// MAD as RIPPLE calls it (:584): order statistics at NP/2 (+1) of an array of `len` = M+1 values.
__device__ double warp_mad_np(const double* res, double* tmp, int len, int np, int lane) {
  const double factor = __ddiv_rn(1.0, 0.6745);
  const int nmid = np / 2;
  double med;
  if (np % 2 == 0) med = __ddiv_rn(__dadd_rn(warp_kth(res, len, nmid, lane), warp_kth(res, len, nmid + 1, lane)), 2.0);
  else med = warp_kth(res, len, nmid + 1, lane);
  for (int i = lane; i < len; i += 32) tmp[i] = fabs(__dsub_rn(res[i], med));
  __syncwarp();
  double r;
  if (np % 2 == 0)
    r = __ddiv_rn(__dmul_rn(factor, __dadd_rn(warp_kth(tmp, len, nmid, lane), warp_kth(tmp, len, nmid + 1, lane))), 2.0);
  else
    r = __dmul_rn(factor, warp_kth(tmp, len, nmid + 1, lane));
  __syncwarp();
  return r;
}

constexpr int RP_MMAX = 16;            // dimension limit of the RIPPLE kernel (thread-private (M+1)x(M+1) systems)
constexpr int RP_THREADS = 128;

// S (n, np2): the np2 nearest other nodes of every node, ascending (the self-kNN's output).
__global__ void __launch_bounds__(RP_THREADS)
ripple_kernel(const double* __restrict__ x, int64_t n, int d, int np0, int np2, const int32_t* __restrict__ S,
              const double* __restrict__ f, double* __restrict__ a, double* __restrict__ rw,
              unsigned long long* __restrict__ last_ier) {
  extern __shared__ double sm[];
  const int mc = d + 3, m1 = d + 1, npmax = np0 + d + 1;
  double* A = sm;                                 // [d+1][npmax] weighted system (work of warp_gelss)
  double* V = A + (size_t)(d + 1) * npmax;        // [d][d]
  double* wk = V + (size_t)d * d;                 // [2d]
  double* sol = wk + 2 * d;                       // [d]
  double* Ain = sol + d;                          // [d][npmax]
  double* bin = Ain + (size_t)d * npmax;          // [npmax]
  double* res = bin + npmax;                      // [npmax]
  double* omega = res + npmax;                    // [npmax]
  double* tmp = omega + npmax;                    // [npmax]
  double* dist = tmp + npmax;                     // [m1]
  double* rbest = dist + m1;                      // [RP_THREADS] best residual of each thread
  double* csol = rbest + RP_THREADS;              // [d] the winning start solution
  int* D = reinterpret_cast<int*>(csol + d);      // [np0][mc] chains
  int* tbest = D + np0 * mc;                      // [RP_THREADS] its task number
  int* keep = tbest + RP_THREADS;                 // [m1] points of the winning subset (then sorted by distance)
  int* keep2 = keep + m1;                         // [m1]
  __shared__ int s_np, s_fail;
  const int t = threadIdx.x, lane = t & 31;
  const int64_t k = blockIdx.x;
  const int32_t* Sk = S + k * np2;
  // ---- chains through the neighbour lists (:540-548)
  if (t < np0) {
    int* ch = D + t * mc;
    ch[0] = Sk[t];
    for (int j = 1; j < mc; ++j) {
      const int32_t* Sp = S + (int64_t)ch[j - 1] * np2;
      for (int l = 0; l <= j; ++l) {  // L = 1..J: the last admissible entry stays
        const int cand = Sp[l];
        bool fresh = cand != (int)k;
        for (int u = 0; u < j && fresh; ++u) fresh = ch[u] != cand;
        if (fresh) ch[j] = cand;
      }
    }
  }
  __syncthreads();
  // ---- every chain, every pair of left-out points (never the first): (M+1) x M least squares by
  // Householder QR in thread-private memory; the residual is the last component of Q^T b (:550-575)
  const int per_chain = (mc - 2) * (mc - 1) / 2;  // pairs (L, I), 2 <= L < I <= M+3
  const int ntask = np0 * per_chain;
  double my_res = CUDART_INF;
  int my_task = 0x7fffffff;
  double my_sol[RP_MMAX];
  const double fk = f[k];
  for (int task = t; task < ntask; task += RP_THREADS) {
    const int j = task / per_chain;
    int rem = task - j * per_chain, I = 2, L = 1;
    while (rem >= I - 1) {  // I = 3.. (0-based 2..), L = 2..I-1 (0-based 1..I-1)
      rem -= I - 1;
      ++I;
    }
    L = 1 + rem;
    double T[(RP_MMAX + 1) * (RP_MMAX + 1)];  // rows r, columns 0..d-1 and the right-hand side in column d
    const int* ch = D + j * mc;
    int r = 0;
    for (int u = 0; u < mc; ++u) {
      if (u == L || u == I) continue;
      const int p = ch[u];
      for (int c = 0; c < d; ++c) T[r * m1 + c] = __dsub_rn(x[(int64_t)p * d + c], x[k * d + c]);
      T[r * m1 + d] = __dsub_rn(f[p], fk);
      ++r;
    }
    bool singular = false;
    double rmax = 0.0;
    for (int c = 0; c < d; ++c) {
      double nrm2 = 0.0;
      for (int i = c; i < m1; ++i) nrm2 = fma(T[i * m1 + c], T[i * m1 + c], nrm2);
      const double nrm = sqrt(nrm2);
      rmax = fmax(rmax, nrm);
      if (!(nrm > (double)m1 * 2.220446049250313e-16 * rmax) || nrm == 0.0) {
        singular = true;  // DGELSS would drop a singular value here; such subsets are not taken as starts
        break;
      }
      const double a0 = T[c * m1 + c];
      const double alpha = a0 >= 0 ? -nrm : nrm;
      const double v0 = a0 - alpha;
      double vv = v0 * v0;
      for (int i = c + 1; i < m1; ++i) vv = fma(T[i * m1 + c], T[i * m1 + c], vv);
      const double beta = vv > 0 ? 2.0 / vv : 0.0;
      for (int cc = c + 1; cc <= d; ++cc) {
        double sdot = v0 * T[c * m1 + cc];
        for (int i = c + 1; i < m1; ++i) sdot = fma(T[i * m1 + c], T[i * m1 + cc], sdot);
        sdot *= beta;
        T[c * m1 + cc] -= sdot * v0;
        for (int i = c + 1; i < m1; ++i) T[i * m1 + cc] -= sdot * T[i * m1 + c];
      }
      T[c * m1 + c] = alpha;
    }
    if (singular) continue;
    const double rr = fabs(T[d * m1 + d]);
    if (rr < my_res) {  // tasks arrive in increasing order: strict < keeps the first minimum (:567)
      my_res = rr;
      my_task = task;
      for (int c = d - 1; c >= 0; --c) {
        double sdot = T[c * m1 + d];
        for (int cc = c + 1; cc < d; ++cc) sdot -= T[c * m1 + cc] * my_sol[cc];
        my_sol[c] = sdot / T[c * m1 + c];
      }
    }
  }
  rbest[t] = my_res;
  tbest[t] = my_task;
  __syncthreads();
  for (int o = RP_THREADS / 2; o > 0; o >>= 1) {
    if (t < o) {
      const double r2 = rbest[t + o];
      const int t2 = tbest[t + o];
      if (r2 < rbest[t] || (r2 == rbest[t] && t2 < tbest[t])) {
        rbest[t] = r2;
        tbest[t] = t2;
      }
    }
    __syncthreads();
  }
  const int win = tbest[0];
  if (win == my_task && win != 0x7fffffff) {
    for (int c = 0; c < d; ++c) csol[c] = my_sol[c];
    const int j = win / per_chain;
    int rem = win - j * per_chain, I = 2;
    while (rem >= I - 1) {
      rem -= I - 1;
      ++I;
    }
    const int L = 1 + rem;
    int r = 0;
    for (int u = 0; u < mc; ++u)
      if (u != L && u != I) keep[r++] = D[j * mc + u];
  }
  if (t == 0) {
    s_fail = win == 0x7fffffff;
    s_np = np0;
  }
  __syncthreads();
  if (t >= 32) return;  // the rest is one warp's work
  if (s_fail) {         // every subset singular: no start (the reference would carry a truncated-SVD solution)
    if (lane == 0) {
      atomicMax(last_ier, ((unsigned long long)(k + 1) << 2) | 2ull);
      rw[k] = sqrt(sqdist_seq(x + (int64_t)Sk[np0 - 1] * d, x + k * d, d));
      for (int c = 0; c < d; ++c) a[k * d + c] = 0.0;
    }
    return;
  }
  // ---- residuals of the winning subset and the scale MAR (:577-584)
  for (int i = lane; i < m1; i += 32) {
    const int p = keep[i];
    bin[i] = __dsub_rn(f[p], fk);
    double s2 = 0.0;
    for (int c = 0; c < d; ++c) {
      const double v = __dsub_rn(x[(int64_t)p * d + c], x[k * d + c]);
      Ain[(size_t)c * npmax + i] = v;
      s2 = __dadd_rn(s2, __dmul_rn(v, v));
    }
    dist[i] = s2;
  }
  __syncwarp();
  for (int i = lane; i < m1; i += 32) {
    double v = 0.0;
    for (int c = 0; c < d; ++c) v = __dadd_rn(v, __dmul_rn(Ain[(size_t)c * npmax + i], csol[c]));
    res[i] = __dsub_rn(v, bin[i]);
  }
  __syncwarp();
  double mar = fmax(warp_mad_np(res, tmp, m1, np0, lane), 1.0e-6);
  // ---- the NP nearest nodes, plus the subset's points beyond the NP-th neighbour (:585-598)
  for (int i = lane; i < np0; i += 32) {
    const int p = Sk[i];
    bin[i] = __dsub_rn(f[p], fk);
    for (int c = 0; c < d; ++c) Ain[(size_t)c * npmax + i] = __dsub_rn(x[(int64_t)p * d + c], x[k * d + c]);
  }
  __syncwarp();
  double rwk = 0.0;
  for (int c = 0; c < d; ++c) rwk = __dadd_rn(rwk, __dmul_rn(Ain[(size_t)c * npmax + np0 - 1], Ain[(size_t)c * npmax + np0 - 1]));
  // stable sort of the subset by distance (SORT, :590): rank by counting
  for (int i = lane; i < m1; i += 32) {
    int rk = 0;
    for (int j = 0; j < m1; ++j) rk += (dist[j] < dist[i] || (dist[j] == dist[i] && j < i)) ? 1 : 0;
    keep2[rk] = keep[i];
    tmp[rk] = dist[i];
  }
  __syncwarp();
  if (lane == 0) {
    int np = np0;
    for (int i = 0; i < m1; ++i) {
      const int p = keep2[i];
      bool among = false;
      for (int u = 0; u < np0; ++u) among = among || Sk[u] == p;
      if (tmp[i] > rwk || (tmp[i] == rwk && !among)) {
        bin[np] = __dsub_rn(f[p], fk);
        for (int c = 0; c < d; ++c) Ain[(size_t)c * npmax + np] = __dsub_rn(x[(int64_t)p * d + c], x[k * d + c]);
        ++np;
      }
    }
    s_np = np;
  }
  __syncwarp();
  const int np = s_np;
  double rk_out = sqrt(rwk);  // :599
  // the helpers index columns with the row count as leading dimension: compact [d][npmax] -> [d][np]
  if (np != npmax) {
    for (int c = 1; c < d; ++c) {
      double v[2];
      int cntv = 0;
      for (int i = lane; i < np; i += 32) v[cntv++] = Ain[(size_t)c * npmax + i];
      __syncwarp();
      cntv = 0;
      for (int i = lane; i < np; i += 32) Ain[(size_t)c * np + i] = v[cntv++];
      __syncwarp();
    }
  }
  for (int c = lane; c < d; c += 32) sol[c] = csol[c];
  __syncwarp();
  warp_residuals(Ain, bin, sol, res, np, d, lane);  // :600
  for (int j = lane; j < d; j += 32) a[k * d + j] = csol[j];
  int e1 = warp_irls(A, V, wk, sol, Ain, bin, res, omega, &mar, 1, np, d, lane);  // :601-604
  bool failed = e1 != 0;
  if (!failed) {
    for (int j = lane; j < d; j += 32) a[k * d + j] = sol[j];                     // :609
    e1 = warp_irls(A, V, wk, sol, Ain, bin, res, omega, &mar, 2, np, d, lane);    // :611-614
    failed = e1 != 0;
    if (!failed)
      for (int j = lane; j < d; j += 32) a[k * d + j] = sol[j];                   // :619
  }
  if (failed) {
    if (lane == 0) atomicMax(last_ier, ((unsigned long long)(k + 1) << 2) | 2ull);  // :605-607, :615-617
  } else {
    for (int i = 0; i < np; ++i) {                                                // :621-634
      if (omega[i] < 0.8) {
        double l1 = 0.0, l0 = 0.0;
        for (int c = 0; c < d; ++c) {
          l1 = __dadd_rn(l1, __dmul_rn(Ain[(size_t)c * np + i], Ain[(size_t)c * np + i]));
          if (i > 0) l0 = __dadd_rn(l0, __dmul_rn(Ain[(size_t)c * np + i - 1], Ain[(size_t)c * np + i - 1]));
        }
        rk_out = i == 0 ? __ddiv_rn(sqrt(l1), 2.0) : __ddiv_rn(__dadd_rn(sqrt(l0), sqrt(l1)), 2.0);
        break;
      }
    }
  }
  if (lane == 0) rw[k] = rk_out;
}

// squared distance of every node to one anchor row; block max -> atomic max (as ordered uint64)
__global__ void far_from_kernel(const double* __restrict__ x, int64_t n, int d,
                                const double* __restrict__ anchor, unsigned long long* __restrict__ best) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  double v = -1.0;
  if (i < n) v = sqdist_seq(x + i * d, anchor, d);
  // pack (d2 bits, index) so the max also yields the farthest row; d2 >= 0 so bits order as uint
  unsigned long long key = i < n ? ((unsigned long long)__double_as_longlong(v) & ~0xFFFFFULL) : 0ULL;
  unsigned long long mx = key;
  for (int o = 16; o > 0; o >>= 1) {
    unsigned long long t = __shfl_xor_sync(0xffffffffu, mx, o);
    mx = t > mx ? t : mx;
  }
  if ((threadIdx.x & 31) == 0) atomicMax(best, mx);
  (void)key;
}
__global__ void far_pick_kernel(const double* __restrict__ x, int64_t n, int d,
                                const double* __restrict__ anchor, const unsigned long long* best,
                                int* __restrict__ pick) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double v = sqdist_seq(x + i * d, anchor, d);
  if (((unsigned long long)__double_as_longlong(v) & ~0xFFFFFULL) == *best) atomicMin(pick, (int)i);
}
__global__ void col_mean_kernel(const double* __restrict__ x, int64_t n, int d, double* __restrict__ mu) {
  // one block per dimension
  __shared__ double sh[256];
  const int j = blockIdx.x;
  double s = 0.0;
  for (int64_t i = threadIdx.x; i < n; i += blockDim.x) s += x[i * d + j];
  sh[threadIdx.x] = s;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) mu[j] = sh[0] / (double)n;
}
__global__ void max_reduce_kernel(const double* __restrict__ v, int64_t n, unsigned long long* __restrict__ best) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  unsigned long long mx = i < n ? (unsigned long long)__double_as_longlong(v[i]) : 0ULL;
  for (int o = 16; o > 0; o >>= 1) {
    unsigned long long t = __shfl_xor_sync(0xffffffffu, mx, o);
    mx = t > mx ? t : mx;
  }
  if ((threadIdx.x & 31) == 0) atomicMax(best, mx);
}
// exact DIAM: thread per row i, all rows j stream through shared memory
__global__ void __launch_bounds__(128)
diam_exact_kernel(const double* __restrict__ x, int64_t n, int d, int tile_pts,
                  unsigned long long* __restrict__ best) {
  extern __shared__ double sm[];
  double* zs = sm;
  double* xs = sm + (size_t)d * 128;
  const int t = threadIdx.x;
  const int64_t p = (int64_t)blockIdx.x * 128 + t;
  for (int e = t; e < d * 128; e += 128) {
    const int qq = e / d, j = e - qq * d;
    const int64_t pp = (int64_t)blockIdx.x * 128 + qq;
    zs[(size_t)j * 128 + qq] = pp < n ? x[pp * d + j] : 0.0;
  }
  double mx = 0.0;
  for (int64_t base = 0; base < n; base += tile_pts) {
    const int tc = (int)imin64(tile_pts, n - base);
    __syncthreads();
    for (int e = t; e < tc * d; e += 128) xs[e] = x[base * d + e];
    __syncthreads();
    if (p < n)
      for (int i = 0; i < tc; ++i) {
        double a = 0.0;
        for (int j = 0; j < d; ++j) {
          const double df = __dsub_rn(xs[(size_t)i * d + j], zs[(size_t)j * 128 + t]);
          a = __dadd_rn(a, __dmul_rn(df, df));
        }
        mx = fmax(mx, a);
      }
  }
  unsigned long long key = (unsigned long long)__double_as_longlong(mx);
  for (int o = 16; o > 0; o >>= 1) {
    unsigned long long tt = __shfl_xor_sync(0xffffffffu, key, o);
    key = tt > key ? tt : key;
  }
  if ((t & 31) == 0) atomicMax(best, key);
}
__global__ void clamp_rw_kernel(double* __restrict__ rw, int64_t n, double half) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) rw[i] = fmin(half, rw[i]);  // :177
}
// unit normals [A(:,K); -1] / ||.|| used by the slope-angle test (:463-464)
__global__ void unit_slope_kernel(const double* __restrict__ a, int64_t n, int d, double* __restrict__ ah) {
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  double s = 0.0;
  for (int j = 0; j < d; ++j) s = __dadd_rn(s, __dmul_rn(a[k * d + j], a[k * d + j]));
  s = sqrt(__dadd_rn(s, 1.0));
  for (int j = 0; j < d; ++j) ah[k * (d + 1) + j] = __ddiv_rn(a[k * d + j], s);
  ah[k * (d + 1) + d] = __ddiv_rn(-1.0, s);
}

// --------------------------------------------------------------- eval kernel
constexpr int LS_QB = 128;

template <int DMAX>
__global__ void __launch_bounds__(LS_QB)
lshep_eval_kernel(const double* __restrict__ x, const float* __restrict__ x32, const double* __restrict__ f,
                  const double* __restrict__ a,
                  const double* __restrict__ ah, const double* __restrict__ rw, double radius, int64_t n,
                  int d, const double* __restrict__ z, int64_t q, int tile_pts, double* __restrict__ out,
                  int32_t* __restrict__ ier, int ostride, double x_absmax) {
  extern __shared__ double sm[];
  double* xs = sm;                          // [tile][d]
  double* as = xs + (size_t)tile_pts * d;   // [tile][d]   gradients, later unit normals [tile][d+1]
  double* fs = as + (size_t)tile_pts * (d + 1);
  double* rs = fs + tile_pts;
  // squared-radius band: outside [lo, hi) the test sqrt(acc) < rk is decided without the square root
  // (it is correctly rounded and monotone; the band is ~100 ulp wide), inside it falls back to sqrt
  double* r2lo = rs + tile_pts;
  double* r2hi = r2lo + tile_pts;
  // fp32 prefilter (common.cuh): float copies of the tile, rows padded to a multiple of 4
  constexpr int D4 = (DMAX + 3) / 4;
  const int dsf = ((d + 3) / 4) * 4;
  float* xf = reinterpret_cast<float*>(r2hi + tile_pts);  // [tile_pts][dsf]
  float* rkf = xf + (size_t)tile_pts * dsf;               // [tile_pts] prefilter threshold of the node's radius
  unsigned char* qn = reinterpret_cast<unsigned char*>(rkf + tile_pts);  // [tile_pts][LS_QB] per-thread node queues
  const int t = threadIdx.x;
  const int64_t p = (int64_t)blockIdx.x * LS_QB + t;
  const bool live = p < q;
  double zq[DMAX], slope[DMAX + 1];
  float zf[D4 * 4];
  double cm = x_absmax;
#pragma unroll
  for (int j = 0; j < DMAX; ++j) {
    zq[j] = (live && j < d) ? z[p * d + j] : 0.0;
    cm = fmax(cm, fabs(zq[j]));
    slope[j] = 0.0;
  }
#pragma unroll
  for (int j = 0; j < D4 * 4; ++j) zf[j] = j < DMAX ? (float)zq[j] : 0.0f;
  {  // one rounding gap for the whole block (the largest coordinate of its queries and of the data), so
     // that a node's threshold is the same for every thread and is prepared once per tile
    __shared__ double s_cm[32];
    double wm = cm;
    for (int o = 16; o > 0; o >>= 1) wm = fmax(wm, __shfl_xor_sync(0xffffffffu, wm, o));
    if ((threadIdx.x & 31) == 0) s_cm[threadIdx.x >> 5] = wm;
    __syncthreads();
    for (int w2 = 0; w2 < (int)(blockDim.x >> 5); ++w2) cm = fmax(cm, s_cm[w2]);
  }
  const float g0 = pre_gap(d, cm), eps1 = pre_eps1(d);
  float thr_fb = CUDART_INF_F;  // while no contributor is known: nodes nearer than 1/temp still matter
  slope[DMAX] = 0.0;
  double D[DMAX + 1];
  int DI[DMAX + 1];
#pragma unroll 1
  for (int j = 0; j <= DMAX; ++j) { D[j] = 0.0; DI[j] = 0; }
  int J = 0, c = 0, hit = -1;
  double temp = 0.0, sw = 0.0, sv = 0.0;
  const int m1 = d + 1;

  for (int64_t base = 0; base < n; base += tile_pts) {
    const int tc = (int)imin64(tile_pts, n - base);
    __syncthreads();
    load_tile_f64_f32(x + base * d, x32 + base * dsf, tc, d, dsf, xs, xf, t, LS_QB);
    for (int e = t; e < tc * d; e += LS_QB) as[e] = a[base * d + e];
    for (int e = t; e < tc; e += LS_QB) {
      const double rk = __dmul_rn(rw[base + e], radius);
      fs[e] = f[base + e];
      rs[e] = rk;
      r2lo[e] = rk * rk * (1.0 - 1e-14);
      r2hi[e] = rk * rk * (1.0 + 1e-14);
      rkf[e] = pre_thr(pre_up(rk), g0, eps1);
    }
    __syncthreads();
    if (!live || hit >= 0) continue;
    // Phase A: nodes four at a time through the fp32 prefilter (see shepmod_eval_kernel); the nodes it
    // cannot rule out are queued per thread in shared memory, in index order.  With `radius` = 2 several
    // per cent of all nodes contribute to a query: taken inline, each lane would reach its contributors at
    // different nodes and the warp would run the fp64 path at ~10 % lane use.  Phase B below runs it on
    // every lane's own queue instead.  Thresholds are those at the start of the tile (only ever looser).
    int cq = 0;
    for (int i0 = 0; i0 < tc; i0 += 4) {
      float af[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
      for (int j4 = 0; j4 < D4; ++j4)
        if (j4 * 4 < d) {
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const float4 v = reinterpret_cast<const float4*>(xf + (size_t)(i0 + u) * dsf)[j4];
            const float t0 = zf[j4 * 4] - v.x, t1 = zf[j4 * 4 + 1] - v.y, t2 = zf[j4 * 4 + 2] - v.z,
                        t3 = zf[j4 * 4 + 3] - v.w;
            af[u] = fmaf(t0, t0, af[u]);
            af[u] = fmaf(t1, t1, af[u]);
            af[u] = fmaf(t2, t2, af[u]);
            af[u] = fmaf(t3, t3, af[u]);
          }
        }
      unsigned need = 0;
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        float thr = rkf[i0 + u];
        if (c == 0) thr = fmaxf(thr, thr_fb);
        if (i0 + u < tc && !(af[u] >= thr)) need |= 1u << u;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (need & (1u << u)) qn[(size_t)(cq++) * LS_QB + t] = (unsigned char)(i0 + u);
    }
    // Phase B: exact fp64 path over my queue
    for (int e2 = 0; e2 < cq; ++e2) {
      {
      const int i = qn[(size_t)e2 * LS_QB + t];
      const double* xr = xs + (size_t)i * d;
      double df[DMAX];
      double acc = 0.0;
#pragma unroll
      for (int j = 0; j < DMAX; ++j)
        if (j < d) {
          df[j] = __dsub_rn(zq[j], xr[j]);
          acc = __dadd_rn(acc, __dmul_rn(df[j], df[j]));
        }
      const double rk = rs[i];
      const bool inside = acc < r2lo[i] ? true : (acc >= r2hi[i] ? false : sqrt(acc) < rk);  // DIST < RW(K), :427
      if (inside) {
        const double dist = sqrt(acc);  // :426
        if (__dsub_rn(rk, dist) == rk) { hit = (int)(base + i); break; }  // :428-431
        const double u = __ddiv_rn(__dsub_rn(rk, dist), __dmul_rn(rk, dist));
        const double w = __dmul_rn(u, u);  // :433
        const double* ar = as + (size_t)i * d;
        double nv = 0.0;  // :435
#pragma unroll
        for (int j = 0; j < DMAX; ++j)
          if (j < d) {
            nv = __dadd_rn(nv, __dmul_rn(ar[j], df[j]));
            slope[j] = fma(ar[j], w, slope[j]);
          }
        nv = __dadd_rn(nv, fs[i]);
        slope[DMAX] -= w;
        sw = __dadd_rn(sw, w);
        sv = fma(w, nv, sv);
        ++c;
      } else if (c == 0) {
        const double w = __ddiv_rn(1.0, sqrt(acc));  // :439-445
        if (w > temp) {
          D[J] = w;  // rare bookkeeping (only until the first contributor): local memory is fine
          DI[J] = (int)(base + i);
          int jm = 0;
          double dm = D[0];
#pragma unroll 1
          for (int u2 = 1; u2 < m1; ++u2)
            if (D[u2] < dm) { dm = D[u2]; jm = u2; }
          J = jm;
          temp = dm;
          thr_fb = temp > 0.0 ? pre_thr(pre_up(1.0 / temp), g0, eps1) : CUDART_INF_F;
        }
      }
      }
    }
  }
  int e = 0;
  double val = 0.0;
  bool need2 = false;
  if (live) {
    if (hit >= 0) {
      val = f[hit];
    } else if (c == 0) {  // :450-454
      e = 1;
      double sd = 0.0;
#pragma unroll 1
      for (int u2 = 0; u2 < m1; ++u2) { D[u2] = __dmul_rn(D[u2], D[u2]); sd = __dadd_rn(sd, D[u2]); }
#pragma unroll 1
      for (int u2 = 0; u2 < m1; ++u2) val = __dadd_rn(val, __dmul_rn(__ddiv_rn(D[u2], sd), f[DI[u2]]));
    } else {
      val = sv / sw;  // :456-457 up to summation order
      need2 = true;
      double nn = slope[DMAX] * slope[DMAX];
#pragma unroll
      for (int j = 0; j < DMAX; ++j) nn = fma(slope[j], slope[j], nn);
      nn = sqrt(nn);
#pragma unroll
      for (int j = 0; j <= DMAX; ++j) slope[j] = slope[j] / nn;  // :462 (scale of SW cancels)
    }
  }
  // second pass: IER = 2 iff some contributing normal has cosine < 0.9 with the mean (:460-469)
  for (int64_t base = 0; base < n; base += tile_pts) {
    const int tc = (int)imin64(tile_pts, n - base);
    if (__syncthreads_and(!need2 || e == 2)) break;  // every query of the block is decided
    load_tile_f64_f32(x + base * d, x32 + base * dsf, tc, d, dsf, xs, xf, t, LS_QB);
    for (int e2 = t; e2 < tc * (d + 1); e2 += LS_QB) as[e2] = ah[base * (d + 1) + e2];
    for (int e2 = t; e2 < tc; e2 += LS_QB) {
      const double rk = __dmul_rn(rw[base + e2], radius);
      rs[e2] = rk;
      r2lo[e2] = rk * rk * (1.0 - 1e-14);
      r2hi[e2] = rk * rk * (1.0 + 1e-14);
      rkf[e2] = pre_thr(pre_up(rk), g0, eps1);
    }
    __syncthreads();
    if (!need2 || e == 2) continue;
    for (int i0 = 0; i0 < tc && e != 2; i0 += 4) {
      float af[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
      for (int j4 = 0; j4 < D4; ++j4)
        if (j4 * 4 < d) {
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const float4 v = reinterpret_cast<const float4*>(xf + (size_t)(i0 + u) * dsf)[j4];
            const float t0 = zf[j4 * 4] - v.x, t1 = zf[j4 * 4 + 1] - v.y, t2 = zf[j4 * 4 + 2] - v.z,
                        t3 = zf[j4 * 4 + 3] - v.w;
            af[u] = fmaf(t0, t0, af[u]);
            af[u] = fmaf(t1, t1, af[u]);
            af[u] = fmaf(t2, t2, af[u]);
            af[u] = fmaf(t3, t3, af[u]);
          }
        }
      unsigned need = 0;
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (i0 + u < tc && !(af[u] >= rkf[i0 + u])) need |= 1u << u;
      while (need) {
      const int i = i0 + __ffs(need) - 1;
      need &= need - 1;
      const double* xr = xs + (size_t)i * d;
      double acc = 0.0;
#pragma unroll
      for (int j = 0; j < DMAX; ++j)
        if (j < d) {
          const double dfj = __dsub_rn(zq[j], xr[j]);
          acc = __dadd_rn(acc, __dmul_rn(dfj, dfj));
        }
      if (acc < r2lo[i] ? true : (acc >= r2hi[i] ? false : sqrt(acc) < rs[i])) {
        const double* hr = as + (size_t)i * (d + 1);
        double dp = slope[DMAX] * hr[d];
#pragma unroll
        for (int j = 0; j < DMAX; ++j)
          if (j < d) dp = fma(slope[j], hr[j], dp);
        if (dp < 0.9) { e = 2; break; }
      }
      }
    }
  }
  if (live) {
    out[p * ostride] = val;
    ier[p * ostride] = e;
  }
}

}  // namespace

int lshep_fit(b200_model* m, const double* f_dev, int r, cudaStream_t st, int* ier, int mode) {
  // mode 0: LSHEP; 1: LSHEP with RLSHEP present (robust); 2: RIPPLE
  const bool robust = mode == 1, ripple = mode == 2;
  const int64_t n = m->n;
  const int d = m->d;
  *ier = 0;
  if (d < 1 || n <= d + (ripple ? 3 : 1)) {  // :91-94, :520-523
    *ier = 1;
    return 0;
  }
  if (ripple && d > RP_MMAX) {
    set_error("ripple: dimension %d > %d unsupported", d, RP_MMAX);
    return 1;
  }
  if (d > 64) {
    set_error("lshep_fit: dimension %d > 64 unsupported", d);
    return 1;
  }
  const int np0 = (int)imin64(n - 1, (3 * d + 1) / 2);  // :98
  const int np = ripple ? (int)imin64(n - 1, (3 * d + 5) / 2) : np0;  // RIPPLE's neighbour lists are longer (:525)
  B200_TRY(m->idx32.reserve((size_t)n * np * 4));
  B200_TRY(m->d2.reserve((size_t)n * np * 8));
  B200_TRY(m->rw.reserve((size_t)n * 8));
  B200_TRY(m->lf.reserve((size_t)r * n * 8));
  B200_TRY(m->la.reserve((size_t)r * n * (2 * d + 1) * 8));  // a (r,n,d) then unit normals (r,n,d+1)
  B200_TRY(m->misc.reserve((size_t)np * (d + 1) * n * 8 + 64));
  B200_TRY(m->flags.reserve(256));
  B200_TRY(m->cnt.reserve((size_t)n * 4 + 64));  // nodes to be solved again by the SVD kernel
  B200_CUDA(cudaMemcpyAsync(m->lf.p, f_dev, (size_t)r * n * 8, cudaMemcpyDeviceToDevice, st));
  B200_TRY(knn_dispatch(m, m->x, n, np, 0, m->idx32.as<int32_t>(), m->d2.as<double>(), st));
  B200_CUDA(cudaMemsetAsync(m->flags.p, 0, 256, st));
  int32_t* flag = m->flags.as<int32_t>();
  const unsigned gb = (unsigned)((n + 127) / 128);
  unsigned long long* last_ier = reinterpret_cast<unsigned long long*>(m->flags.as<char>() + 32);
  for (int c = 0; c < r; ++c) {
    double* ac = m->la.as<double>() + (size_t)c * n * d;
    if (ripple) {
      const int npmax = np0 + d + 1, m1 = d + 1, mc = d + 3;
      const size_t smem = ((size_t)(2 * d + 1) * npmax + (size_t)d * d + 4 * (size_t)d + 4 * (size_t)npmax + m1 + RP_THREADS) * 8 +
                          ((size_t)np0 * mc + RP_THREADS + 2 * m1) * 4 + 64;
      B200_CUDA(cudaFuncSetAttribute(ripple_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      ripple_kernel<<<(unsigned)n, RP_THREADS, smem, st>>>(m->x, n, d, np0, np, m->idx32.as<int32_t>(),
                                                          m->lf.as<double>() + (size_t)c * n, ac, m->rw.as<double>(), last_ier);
      unit_slope_kernel<<<gb, 128, 0, st>>>(ac, n, d, m->la.as<double>() + (size_t)r * n * d + (size_t)c * n * (d + 1));
      B200_CUDA(cudaGetLastError());
      m->st.launches += 2;
      continue;
    }
    if (robust) {
      // RW depends on the response through the final IRLS weights; with several columns the last one stays,
      // as in the wrapper's per-column loop (linear_shepard/__init__.py:47-55 keeps one rw per column: the
      // caller reads it back after each column through b200_lshep_fit_robust's rw argument)
      const size_t smem = ((size_t)(2 * d + 1) * np + (size_t)d * d + 3 * (size_t)d + 5 * (size_t)np) * 8;
      B200_CUDA(cudaFuncSetAttribute(lshep_robust_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      lshep_robust_kernel<<<(unsigned)n, 32, smem, st>>>(m->x, n, d, np, m->idx32.as<int32_t>(), m->d2.as<double>(),
                                                         m->lf.as<double>() + (size_t)c * n, ac, m->rw.as<double>(),
                                                         last_ier);
      unit_slope_kernel<<<gb, 128, 0, st>>>(ac, n, d, m->la.as<double>() + (size_t)r * n * d + (size_t)c * n * (d + 1));
      B200_CUDA(cudaGetLastError());
      m->st.launches += 2;
      continue;
    }
    lshep_build_kernel<<<gb, 128, 0, st>>>(m->x, n, d, np, m->idx32.as<int32_t>(), m->d2.as<double>(),
                                           m->lf.as<double>() + (size_t)c * n, m->misc.as<double>(),
                                           m->rw.as<double>());
    int32_t* sus = m->cnt.as<int32_t>();
    int32_t* sus_cnt = flag + 4;
    B200_CUDA(cudaMemsetAsync(sus_cnt, 0, 4, st));
    if (d <= 16) lshep_qr_kernel<16><<<gb, 128, 0, st>>>(m->misc.as<double>(), n, d, np, ac, nullptr, sus, sus_cnt);
    else lshep_qr_kernel<64><<<gb, 128, 0, st>>>(m->misc.as<double>(), n, d, np, ac, nullptr, sus, sus_cnt);
    int h_sus = 0;
    B200_CUDA(cudaMemcpyAsync(&h_sus, sus_cnt, 4, cudaMemcpyDeviceToHost, st));
    B200_CUDA(cudaStreamSynchronize(st));
    if (h_sus > 0) {
      const size_t smem = ((size_t)(d + 1) * np + (size_t)d * d + 3 * (size_t)d) * 8;
      B200_CUDA(cudaFuncSetAttribute(lshep_svd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      lshep_svd_kernel<<<(unsigned)h_sus, 32, smem, st>>>(m->x, n, d, np, m->idx32.as<int32_t>(), m->d2.as<double>(),
                                                          m->lf.as<double>() + (size_t)c * n, sus, ac, flag);
      m->st.launches += 1;
    }
    unit_slope_kernel<<<gb, 128, 0, st>>>(ac, n, d, m->la.as<double>() + (size_t)r * n * d + (size_t)c * n * (d + 1));
    B200_CUDA(cudaGetLastError());
    m->st.launches += 3;
  }
  // ---- DIAM clamp (:114,:177)
  unsigned long long* best = reinterpret_cast<unsigned long long*>(m->flags.as<char>() + 64);
  int* pick = reinterpret_cast<int*>(m->flags.as<char>() + 128);
  double* mu = reinterpret_cast<double*>(m->misc.p);  // scratch free again (QR done, stream ordered)
  unsigned long long h_best[3];
  col_mean_kernel<<<d, 256, 0, st>>>(m->x, n, d, mu);
  const double* anchor = mu;
  for (int sweep = 0; sweep < 2; ++sweep) {
    B200_CUDA(cudaMemsetAsync(best + sweep, 0, 8, st));
    int big = 0x7fffffff;
    B200_CUDA(cudaMemcpyAsync(pick, &big, 4, cudaMemcpyHostToDevice, st));
    far_from_kernel<<<gb, 128, 0, st>>>(m->x, n, d, anchor, best + sweep);
    far_pick_kernel<<<gb, 128, 0, st>>>(m->x, n, d, anchor, best + sweep, pick);
    int h_pick = 0;
    B200_CUDA(cudaMemcpyAsync(&h_pick, pick, 4, cudaMemcpyDeviceToHost, st));
    B200_CUDA(cudaStreamSynchronize(st));
    anchor = m->x + (int64_t)h_pick * d;
    m->st.launches += 2;
  }
  B200_CUDA(cudaMemsetAsync(best + 2, 0, 8, st));
  max_reduce_kernel<<<gb, 128, 0, st>>>(m->rw.as<double>(), n, best + 2);
  B200_CUDA(cudaMemcpyAsync(h_best, best, 24, cudaMemcpyDeviceToHost, st));
  int h_flag = 0;
  B200_CUDA(cudaMemcpyAsync(&h_flag, flag, 4, cudaMemcpyDeviceToHost, st));
  B200_CUDA(cudaStreamSynchronize(st));
  double lower, maxrw;
  memcpy(&lower, &h_best[1], 8);  // truncated key: a valid lower bound of the true d2
  memcpy(&maxrw, &h_best[2], 8);
  if (!(maxrw <= sqrt(lower) / 2.0)) {
    // the clamp may bind: take the exact diameter
    const size_t zb = (size_t)d * 128 * 8;
    int tile_pts = (int)((96 * 1024 - zb) / ((size_t)d * 8));
    if (tile_pts > 256) tile_pts = 256;
    const size_t smem = zb + (size_t)tile_pts * d * 8;
    B200_CUDA(cudaFuncSetAttribute(diam_exact_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    B200_CUDA(cudaMemsetAsync(best, 0, 8, st));
    diam_exact_kernel<<<gb, 128, smem, st>>>(m->x, n, d, tile_pts, best);
    B200_CUDA(cudaMemcpyAsync(h_best, best, 8, cudaMemcpyDeviceToHost, st));
    B200_CUDA(cudaStreamSynchronize(st));
    double diam;
    memcpy(&diam, &h_best[0], 8);
    clamp_rw_kernel<<<gb, 128, 0, st>>>(m->rw.as<double>(), n, sqrt(diam) / 2.0);
    B200_CUDA(cudaGetLastError());
    m->st.launches += 2;
  }
  if (robust || ripple) {
    unsigned long long h_last = 0;
    B200_CUDA(cudaMemcpy(&h_last, last_ier, 8, cudaMemcpyDeviceToHost));
    h_flag = (int)(h_last & 3ull);
  }
  *ier = h_flag;
  m->lshep_r = r;
  m->lshep_ready = true;
  return 0;
}

int lshep_predict(b200_model* m, double radius, const double* z_dev, int64_t q, double* out_dev,
                  int32_t* ier_dev, cudaStream_t st) {
  if (!m->lshep_ready) {
    set_error("lshep_predict: model not fitted (call b200_lshep_fit)");
    return 1;
  }
  if (q == 0) return 0;
  const int64_t n = m->n;
  const int d = m->d, r = m->lshep_r;
  const int dsf = ((d + 3) / 4) * 4;
  const size_t per_pt = (size_t)(2 * d + 5) * 8 + (size_t)(dsf + 1) * 4 + LS_QB;  // + one queue byte per thread
  int tile_pts = (int)((56 * 1024) / per_pt);  // small tiles: several blocks per SM hide the scan's latencies
  if (tile_pts > 256) tile_pts = 256;  // node numbers inside a tile fit the queues' bytes
  tile_pts &= ~3;
  if (tile_pts < 4) tile_pts = 4;
  const size_t smem = (size_t)tile_pts * per_pt;
  double absmax = 0.0;
  B200_TRY(data_absmax(m, st, &absmax));
  const float* x32 = nullptr;
  B200_TRY(data_f32(m, st, &x32));
  const unsigned gb = (unsigned)((q + LS_QB - 1) / LS_QB);
  for (int c = 0; c < r; ++c) {
    const double* fc = m->lf.as<double>() + (size_t)c * n;
    const double* ac = m->la.as<double>() + (size_t)c * n * d;
    const double* hc = m->la.as<double>() + (size_t)r * n * d + (size_t)c * n * (d + 1);
#define LS_LAUNCH(DM)                                                                                 \
  do {                                                                                                \
    auto kern = lshep_eval_kernel<DM>;                                                                \
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));    \
    kern<<<gb, LS_QB, smem, st>>>(m->x, x32, fc, ac, hc, m->rw.as<double>(), radius, n, d, z_dev, q,   \
                                  tile_pts, out_dev + c, ier_dev + c, r, absmax);                     \
  } while (0)
    if (d <= 2) LS_LAUNCH(2);
    else if (d <= 4) LS_LAUNCH(4);
    else if (d <= 6) LS_LAUNCH(6);
    else if (d <= 8) LS_LAUNCH(8);
    else if (d <= 10) LS_LAUNCH(10);
    else if (d <= 12) LS_LAUNCH(12);
    else if (d <= 16) LS_LAUNCH(16);
    else if (d <= 24) LS_LAUNCH(24);
    else if (d <= 32) LS_LAUNCH(32);
    else LS_LAUNCH(64);
#undef LS_LAUNCH
    B200_CUDA(cudaGetLastError());
    m->st.launches += 1;
  }
  return 0;
}

}  // namespace b200
