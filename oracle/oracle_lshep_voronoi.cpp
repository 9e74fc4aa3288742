// oracle_lshep_voronoi.cpp -- CPU restatement of LSHEP / LSHEPVAL
// (linear_shepard/linear_shepard.f95), Voronoi PREDICT (voronoi/voronoi.f90) and
// the NearestNeighbor sub-model loop (nearest_neighbor/__init__.py:69-83).
// TEST INFRASTRUCTURE ONLY -- see oracle.cpp for who may use oracle/.
// Parity for these Fortran routines is UNPINNED (no gfortran in this image);
// LSHEP's least squares goes through the real LAPACK DGELSS, handed in as a
// function pointer taken from scipy.linalg.cython_lapack (oracle/harness.py).

#include "oracle_common.h"

extern "C" {

int orc_shepmod_val(const double* xp, int d, int64_t n, const double* x, const double* rw,
                    double* wts);
void orc_shepard_weight(const double* pts, int64_t n, int d, const double* pt, double* wts);

typedef void (*dgelss_fn)(int* m, int* n, int* nrhs, double* a, int* lda, double* b, int* ldb,
                          double* s, double* rcond, int* rank, double* work, int* lwork,
                          int* info);

// ---------------------------------------------------------------------------
// LSHEP, non-robust branch (linear_shepard.f95:23-179 minus :136-175, which the
// Python wrapper can never reach).  x (n,d) row-major, f (n).
// a (n,d) row-major: a[k*d + j] = A(j+1, k+1).  rw (n) WITHOUT the wrapper's
// radius multiplier.  Returns IER (0, 1 too few points, 2 some system rank
// deficient); rank (n) optional per-node DGELSS rank.
// ---------------------------------------------------------------------------
int orc_lshep_fit(const double* x, int64_t n, int d, const double* f, double* a, double* rw,
                  int32_t* rank_out, dgelss_fn dgelss) {
  if (d < 1 || n <= d + 1) return 1;                       // :91-94
  const int np_ = (int)std::min<int64_t>(n - 1, (3 * d + 1) / 2);  // :98
  const double rcond = np_ * DBL_EPSILON;                   // :100
  int ier = 0;
  double diam = 0.0;
#pragma omp parallel
  {
    std::vector<Cand> c(np_);
    std::vector<double> alpha((size_t)np_ * d), beta(np_), sigma(d), dist(np_);
    int lwork = 3 * d + 2 * np_ + 64 * (np_ + d);  // >= the Fortran 3M+2NP (:102); extra is harmless
    std::vector<double> work(lwork);
    double ldiam = 0.0;
    int lier = 0;
#pragma omp for schedule(dynamic, 16)
    for (int64_t k = 0; k < n; ++k) {
      double mx;
      knn_one(x, n, d, x + k * d, np_, k, c.data(), &mx);  // :105-115 (stable SORT: ties -> lower index)
      ldiam = std::max(ldiam, mx);                          // :114
      for (int i = 0; i < np_; ++i) dist[i] = std::sqrt(c[i].d2);  // :116
      rw[k] = dist[np_ - 1];                                // :121
      const double rp = 1.1 * rw[k];                        // :125
      for (int i = 0; i < np_; ++i) {
        const double om = (rp - dist[i]) / (rp * dist[i]);  // :126
        beta[i] = om * (f[c[i].i] - f[k]);                  // :117,:128
        for (int j = 0; j < d; ++j)                         // :118-120,:129-131 (column-major NP x M)
          alpha[(size_t)j * np_ + i] = om * (x[c[i].i * d + j] - x[k * d + j]);
      }
      int m_ = np_, n_ = d, nrhs = 1, lda = np_, ldb = np_, rank = 0, info = 0;
      double rc = rcond;
      dgelss(&m_, &n_, &nrhs, alpha.data(), &lda, beta.data(), &ldb, sigma.data(), &rc, &rank,
             work.data(), &lwork, &info);                   // :132
      for (int j = 0; j < d; ++j) a[k * d + j] = beta[j];   // :134
      if (rank < d) lier = 2;                               // :135
      if (rank_out) rank_out[k] = rank;
    }
#pragma omp critical
    {
      diam = std::max(diam, ldiam);
      if (lier) ier = lier;
    }
  }
  const double half = std::sqrt(diam) / 2.0;  // :177
  for (int64_t k = 0; k < n; ++k) rw[k] = std::min(half, rw[k]);
  return ier;
}

// ---------------------------------------------------------------------------
// LSHEPVAL (linear_shepard.f95:365-472), one query.  a (n,d) row-major as above.
// ---------------------------------------------------------------------------
double orc_lshep_val(const double* xp, int d, int64_t n, const double* x, const double* f,
                     const double* a, const double* rw, int32_t* ier_out) {
  const int m1 = d + 1;
  std::vector<double> D(m1, 0.0);
  std::vector<int64_t> ID(m1, 0);
  std::vector<double> sw, swc;
  std::vector<int64_t> sk;  // node index of each contributor (its slope is [A(:,K); -1])
  int j = 0;
  double temp = 0.0;
  int ier = 0;
  for (int64_t k = 0; k < n; ++k) {
    const double* xk = x + k * d;
    double dist = std::sqrt(sqdist(xp, xk, d));  // :426
    if (dist < rw[k]) {
      if (rw[k] - dist == rw[k]) {  // :428-431
        *ier_out = 0;
        return f[k];
      }
      double t = (rw[k] - dist) / (rw[k] * dist);  // :433
      sw.push_back(t * t);
      double s = 0.0;  // :435 DOT_PRODUCT(A(:,K), XP - X(:,K)) + F(K)
      for (int i = 0; i < d; ++i) s += a[k * d + i] * (xp[i] - xk[i]);
      swc.push_back(s + f[k]);
      sk.push_back(k);
    } else {
      double w = 1.0 / dist;  // :439-445
      if (w > temp) {
        D[j] = w;
        ID[j] = k;
        j = (int)(std::min_element(D.begin(), D.end()) - D.begin());
        temp = D[j];
      }
    }
  }
  const size_t I = sw.size();
  if (I == 0) {  // :450-454
    double s = 0.0;
    for (int i = 0; i < m1; ++i) {
      D[i] = D[i] * D[i];
    }
    for (int i = 0; i < m1; ++i) s += D[i];
    double v = 0.0;
    for (int i = 0; i < m1; ++i) v += (D[i] / s) * f[ID[i]];
    *ier_out = 1;
    return v;
  }
  double s = 0.0;  // :456
  for (size_t i = 0; i < I; ++i) s += sw[i];
  for (size_t i = 0; i < I; ++i) sw[i] = sw[i] / s;
  double val = 0.0;  // :457
  for (size_t i = 0; i < I; ++i) val += sw[i] * swc[i];
  // :460-469 slope-angle check
  std::vector<double> slope(m1, 0.0);
  for (size_t i = 0; i < I; ++i) {  // MATMUL(SSLOPE, SW): accumulate over contributors
    for (int r = 0; r < d; ++r) slope[r] += a[sk[i] * d + r] * sw[i];
    slope[d] += -1.0 * sw[i];
  }
  double nn = 0.0;
  for (int r = 0; r < m1; ++r) nn += slope[r] * slope[r];
  nn = std::sqrt(nn);
  for (int r = 0; r < m1; ++r) slope[r] = slope[r] / nn;
  for (size_t i = 0; i < I; ++i) {
    double n2 = 0.0;
    for (int r = 0; r < d; ++r) n2 += a[sk[i] * d + r] * a[sk[i] * d + r];
    n2 += 1.0;
    n2 = std::sqrt(n2);
    double dp = 0.0;
    for (int r = 0; r < d; ++r) dp += slope[r] * (a[sk[i] * d + r] / n2);
    dp += slope[d] * (-1.0 / n2);
    if (dp < 0.9) {
      ier = 2;
      break;
    }
  }
  *ier_out = ier;
  return val;
}

int orc_lshep_predict(const double* x, int64_t n, int d, const double* f, const double* a,
                      const double* rw, const double* z, int64_t q, double* out, int32_t* ier) {
#pragma omp parallel for schedule(dynamic, 4)
  for (int64_t p = 0; p < q; ++p) out[p] = orc_lshep_val(z + p * d, d, n, x, f, a, rw, ier + p);
  return 0;
}

// ---------------------------------------------------------------------------
// Voronoi PREDICT (voronoi.f90:57-155), one query, literal control flow.
// pts (n,d) row-major.  The DOTS cache is rebuilt per call (results do not
// depend on it).  QSORTC is restated as a stable sort by (distance, index).
// Returns ERROR (0 ok, 1 overflow, 2 duplicate points).
// ---------------------------------------------------------------------------
int orc_voronoi_predict(const double* pts, int64_t n, int d, const double* pt, double* w) {
  const double MAXV = DBL_MAX, MINV = std::sqrt(DBL_EPSILON);
  std::vector<double> dist(n), pdot(n), self(n);
  std::vector<int64_t> order(n);
  std::vector<char> skip(n, 0);
  for (int64_t i = 0; i < n; ++i) {
    w[i] = 0.0;
    double s = 0.0;  // :83-86 SUM((EVAL_PT - POINTS(:,I))**2)
    for (int j = 0; j < d; ++j) {
      double t = pt[j] - pts[i * d + j];
      s += t * t;
    }
    dist[i] = s;
    if (dist[i] > MAXV) return 1;  // :88-91
    double sp = 0.0;               // SUM(EVAL_PT * POINTS(:,I)), recomputed at every use in the source
    for (int j = 0; j < d; ++j) sp += pt[j] * pts[i * d + j];
    pdot[i] = sp;
    self[i] = dot(pts + i * d, pts + i * d, d);  // DOTS(I,I) (:105-112)
    if (self[i] > MAXV) return 1;
  }
  std::iota(order.begin(), order.end(), 0);
  std::stable_sort(order.begin(), order.end(),
                   [&](int64_t a, int64_t b) { return dist[a] < dist[b]; });  // :93
  for (int64_t ic_ = 0; ic_ < n; ++ic_) {                                     // :95
    const int64_t ic = order[ic_];
    if (skip[ic]) continue;
    double max_ratio = 0.0;
    bool zeroed = false;
    for (int64_t io_ = 0; io_ < n; ++io_) {  // :100
      const int64_t io = order[io_];
      if (ic == io) continue;
      const double dco = dot(pts + ic * d, pts + io * d, d);  // DOTS(IC,IO) (:121-129)
      if (dco > MAXV) return 1;
      const double c_to_pt = pdot[io] + self[ic] - (pdot[ic] + dco);  // :131-132
      const double c_to_o = self[io] + self[ic] - 2 * dco;            // :134
      if (std::fabs(c_to_o) > MINV) {                                 // :137
        const double ratio = c_to_pt / c_to_o;
        if (ratio > max_ratio) max_ratio = ratio;
        else if (ratio < 0.0) skip[io] = 1;
        else if (ratio >= 1.0) {
          w[ic] = 0.0;
          zeroed = true;
          break;
        }
      } else {
        return 2;  // :145-148
      }
    }
    if (!zeroed) w[ic] = std::max(1.0 - max_ratio, 0.0);  // :151
  }
  double s = 0.0;  // :154
  for (int64_t i = 0; i < n; ++i) s += w[i];
  for (int64_t i = 0; i < n; ++i) w[i] = w[i] / s;
  return 0;
}

// ---------------------------------------------------------------------------
// NearestNeighbor._predict with a sub-model (nearest_neighbor/__init__.py:69-83).
// method: 0 Average, 1 Shepard, 2 Voronoi, 3 ShepMod.  Output ragged rows padded
// to k: ids (q,k) global indices = closest[guess_pts], wts (q,k), cnt (q),
// err (q) sub-model error code (Voronoi ERROR / ShepMod: 0|1 IER; -1 too few).
// ---------------------------------------------------------------------------
int orc_nn_predict(const double* x, int64_t n, int d, const double* z, int64_t q, int k,
                   int method, int64_t* ids, double* wts, int32_t* cnt, int32_t* err) {
  if (k < 1 || k > n) return 1;
#pragma omp parallel
  {
    std::vector<Cand> c(k);
    std::vector<double> loc((size_t)k * d), lw(k), lrw(k);
#pragma omp for schedule(dynamic, 8)
    for (int64_t p = 0; p < q; ++p) {
      const double* pt = z + p * d;
      knn_one(x, n, d, pt, k, -1, c.data());  // :75-76
      for (int i = 0; i < k; ++i) std::memcpy(&loc[(size_t)i * d], x + c[i].i * d, sizeof(double) * d);  // :79
      int e = 0, m = 0;
      if (method == 0) {  // Average :7-14
        for (int i = 0; i < k; ++i) lw[i] = 1.0 / k;
      } else if (method == 1) {  // Shepard: all k, no filtering (shepard/__init__.py:35-47)
        orc_shepard_weight(loc.data(), k, d, pt, lw.data());
      } else if (method == 2) {  // Voronoi (voronoi/__init__.py:52-61)
        e = orc_voronoi_predict(loc.data(), k, d, pt, lw.data());
      } else {  // ShepMod fit on the k neighbours then predict (modified_shepard/__init__.py:28-58)
        if (k <= d + 1) e = -1;
        else {
          for (int i = 0; i < k; ++i) {
            std::vector<Cand> cc(d + 1);
            knn_one(loc.data(), k, d, &loc[(size_t)i * d], d + 1, i, cc.data());
            lrw[i] = std::sqrt(cc[d].d2);
          }
          e = orc_shepmod_val(pt, d, k, loc.data(), lrw.data(), lw.data());
        }
      }
      for (int i = 0; i < k; ++i) {
        bool keep = (method <= 1) ? true : (e >= 0 && !(method == 2 && e != 0) && lw[i] > 0);
        if (keep) {
          ids[p * k + m] = c[i].i;  // closest[guess_pts] :82
          wts[p * k + m] = lw[i];
          ++m;
        }
      }
      for (int i = m; i < k; ++i) {
        ids[p * k + i] = -1;
        wts[p * k + i] = 0.0;
      }
      cnt[p] = m;
      err[p] = e;
    }
  }
  return 0;
}

// ---------------------------------------------------------------------------
// EVAL_MESH (voronoi/voronoi.f90:5-44): the naive full evaluation from precomputed
// inner products.  Arrays are column-major as the Fortran sees them: int_prods (n,n),
// app_int_prods (q,n), mesh (q,n).  Operand order of :26-29; ratios with
// |denominator| <= EPSILON are skipped (:30); SUM over the row in index order (:42).
// ---------------------------------------------------------------------------
void orc_eval_mesh(const double* ip, int64_t n, const double* ap, int64_t q, double* mesh) {
  for (int64_t y = 0; y < q; ++y) {
    for (int64_t xi = 0; xi < n; ++xi) {
      double max_ratio = 0.0;
      for (int64_t c = 0; c < n; ++c) {
        if (c == xi) continue;
        const double num = ap[y + c * q] - ap[y + xi * q] - ip[xi + c * n] + ip[xi + xi * n];
        const double den = ip[c + c * n] - ip[c + xi * n] - ip[xi + c * n] + ip[xi + xi * n];
        if (std::fabs(den) > DBL_EPSILON) {
          const double ratio = num / den;
          if (ratio > max_ratio) max_ratio = ratio;
        }
      }
      mesh[y + xi * q] = std::max(1.0 - max_ratio, 0.0);
    }
    double s = 0.0;
    for (int64_t xi = 0; xi < n; ++xi) s += mesh[y + xi * q];
    for (int64_t xi = 0; xi < n; ++xi) mesh[y + xi * q] /= s;
  }
}

// WeightedApproximator.predict's sum(self.y[i]*w ...) (base.py:169): left to
// right from integer 0, one multiply and one add rounding per term.
void orc_weighted_sum(const int64_t* ids, const double* wts, const int32_t* cnt, int64_t q,
                      int stride, const double* y, double* out) {
  for (int64_t p = 0; p < q; ++p) {
    double s = 0.0;
    for (int i = 0; i < cnt[p]; ++i) s += y[ids[p * stride + i]] * wts[p * stride + i];
    out[p] = s;
  }
}

}  // extern "C"
