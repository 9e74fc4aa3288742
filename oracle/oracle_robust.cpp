// oracle_robust.cpp -- CPU restatement of the robust branch of LSHEP and of its helpers
// IRLS, MAD and QUANTILE (linear_shepard/linear_shepard.f95:136-175, 181-335).
// TEST INFRASTRUCTURE ONLY -- see oracle.cpp for who may use oracle/.
// Parity UNPINNED: the Fortran cannot be compiled in this image and the reference ships no
// vectors for it; pinned only through the invariants the source implies (tests/test_oracle.py:
// QUANTILE = k-th order statistic, MAD against numpy's median, IRLS weights in [0,1], a
// noise-free linear response reproduced exactly).  Every least-squares solve is the real LAPACK
// DGELSS, handed in as a function pointer (oracle/harness.py).

#include "oracle_common.h"

extern "C" {

typedef void (*dgelss_fn)(int* m, int* n, int* nrhs, double* a, int* lda, double* b, int* ldb,
                          double* s, double* rcond, int* rank, double* work, int* lwork,
                          int* info);

// QUANTILE (f95:310-335): the K-th smallest element (1-based), by the source's recursion on
// the elements below / above A(K).
double orc_quantile(const double* a, int n, int k) {
  std::vector<double> cur(a, a + n);
  for (;;) {
    const double ak = cur[k - 1];                                     // :324
    int below = 0, above = 0;
    for (double v : cur) {
      below += v < ak;
      above += v > ak;
    }
    if (below >= k) {                                                 // :326-327
      std::vector<double> nx;
      for (double v : cur)
        if (v < ak) nx.push_back(v);
      cur.swap(nx);
    } else {
      const int j = above + k - (int)cur.size();                      // :329
      if (j > 0) {                                                    // :330-331
        std::vector<double> nx;
        for (double v : cur)
          if (v > ak) nx.push_back(v);
        cur.swap(nx);
        k = j;
      } else {
        return ak;                                                    // :333
      }
    }
  }
}

// MAD (f95:284-308): median absolute deviation / 0.6745.
double orc_mad(const double* res, int np_) {
  const double factor = 1.0 / 0.6745;                                 // :297
  const int nmid = np_ / 2;                                           // :298
  std::vector<double> dev(np_);
  if (np_ % 2 == 0) {                                                 // :299-302
    const double med = (orc_quantile(res, np_, nmid) + orc_quantile(res, np_, nmid + 1)) / 2.0;
    for (int i = 0; i < np_; ++i) dev[i] = std::fabs(res[i] - med);
    return factor * (orc_quantile(dev.data(), np_, nmid) + orc_quantile(dev.data(), np_, nmid + 1)) / 2.0;
  }
  const double med = orc_quantile(res, np_, nmid + 1);                // :304-305
  for (int i = 0; i < np_; ++i) dev[i] = std::fabs(res[i] - med);
  return factor * orc_quantile(dev.data(), np_, nmid + 1);
}

// IRLS (f95:181-282).  alpha_in (np x d, column-major), beta_in (np), res (np) in/out, s in/out
// (the source scales its INTENT(INOUT) argument by the tuning constant, :237).  Outputs omega (np)
// and sol (d).  Returns IER 0 / 1 rank deficient / 2 bisquare did not converge / 3 bad FUNC.
int orc_irls(int d, int np_, double* s, int iter, int func, double* omega, double* sol, const double* beta_in,
             const double* alpha_in, double* res, dgelss_fn dgelss) {
  if (func != 1 && func != 2) return 3;                               // :233-236
  const double tuning[2] = {1.0, 3.0};                                // :231
  *s = tuning[func - 1] * *s;                                         // :237
  double rcond = np_ * DBL_EPSILON;                                   // :239
  int lwork = 3 * d + 2 * np_ + 64 * (np_ + d);
  std::vector<double> work(lwork), alpha((size_t)np_ * d), beta(np_), sigma(d);
  auto objective = [&]() {                                            // :243, :275
    double e = 0.0;
    for (int j = 0; j < np_; ++j) {
      const double t = res[j] / *s;
      const double w = std::max(0.0, 1.0 - t * t);
      e += 1.0 - w * w * w;
    }
    return e;
  };
  double err_old = 0.0;
  if (func == 2) err_old = objective();
  for (int it = 0; it < iter; ++it) {                                 // :245
    for (int j = 0; j < np_; ++j) {
      if (func == 1) {                                                // :249-256 Huber
        omega[j] = res[j] == 0.0 ? 1.0 : std::sqrt(std::min(1.0, *s / std::fabs(res[j])));
      } else {                                                        // :258-259 bisquare
        const double t = res[j] / *s;
        omega[j] = std::max(0.0, 1.0 - t * t);
      }
    }
    for (int j = 0; j < np_; ++j) {                                   // :262-265
      beta[j] = omega[j] * beta_in[j];
      for (int c = 0; c < d; ++c) alpha[(size_t)c * np_ + j] = omega[j] * alpha_in[(size_t)c * np_ + j];
    }
    int m_ = np_, n_ = d, nrhs = 1, lda = np_, ldb = np_, rank = 0, info = 0;
    dgelss(&m_, &n_, &nrhs, alpha.data(), &lda, beta.data(), &ldb, sigma.data(), &rcond, &rank, work.data(),
           &lwork, &info);                                            // :266-267
    for (int c = 0; c < d; ++c) sol[c] = beta[c];
    if (rank < d) return 1;                                           // :268-271
    for (int j = 0; j < np_; ++j) {                                   // :272 MATMUL(ALPHA_IN, BETA) - BETA_IN
      double v = 0.0;
      for (int c = 0; c < d; ++c) v += alpha_in[(size_t)c * np_ + j] * sol[c];
      res[j] = v - beta_in[j];
    }
  }
  if (func == 2 && objective() > err_old) return 2;                   // :274-280
  return 0;
}

// LSHEP with RLSHEP present (f95:91-179): unit weights for the first fit, then up to two IRLS
// stages, RW cut at the first neighbour whose final weight is below 0.8.  x (n,d) row-major,
// a (n,d) row-major, rw (n) before the wrapper's radius multiplier, omega_out (n, np) optional
// final weights.  Returns IER 0/1/2/3.
int orc_lshep_fit_robust(const double* x, int64_t n, int d, const double* f, double* a, double* rw,
                         double* omega_out, dgelss_fn dgelss) {
  if (d < 1 || n <= d + 1) return 1;                                  // :91-94
  const int np_ = (int)std::min<int64_t>(n - 1, (3 * d + 1) / 2);     // :98
  int ier = 0;
  int64_t ier_k = -1;
  double diam = 0.0;
#pragma omp parallel
  {
    std::vector<Cand> c(np_);
    std::vector<double> alpha_in((size_t)np_ * d), alpha((size_t)np_ * d), beta_in(np_), beta(np_), sigma(d),
        dist(np_), omega(np_), res(np_), sol(d);
    int lwork = 3 * d + 2 * np_ + 64 * (np_ + d);
    std::vector<double> work(lwork);
    double ldiam = 0.0;
    int lier = 0;
    int64_t lier_k = -1;  // IER is a scalar the serial node loop overwrites: the highest node that set it wins
#pragma omp for schedule(dynamic, 16)
    for (int64_t k = 0; k < n; ++k) {
      double mx;
      knn_one(x, n, d, x + k * d, np_, k, c.data(), &mx);             // :105-115
      ldiam = std::max(ldiam, mx);
      for (int i = 0; i < np_; ++i) {
        dist[i] = std::sqrt(c[i].d2);                                 // :116
        beta_in[i] = f[c[i].i] - f[k];                                // :117
        for (int j = 0; j < d; ++j) alpha_in[(size_t)j * np_ + i] = x[c[i].i * d + j] - x[k * d + j];  // :118-120
        omega[i] = 1.0;                                               // :122-123
      }
      rw[k] = dist[np_ - 1];                                          // :121
      alpha = alpha_in;                                               // :128-131 with OMEGA = 1
      beta = beta_in;
      int m_ = np_, n_ = d, nrhs = 1, lda = np_, ldb = np_, rank = 0, info = 0;
      double rc = np_ * DBL_EPSILON;                                  // :100
      dgelss(&m_, &n_, &nrhs, alpha.data(), &lda, beta.data(), &ldb, sigma.data(), &rc, &rank, work.data(),
             &lwork, &info);                                          // :132
      for (int j = 0; j < d; ++j) a[k * d + j] = beta[j];             // :134
      if (rank < d && k > lier_k) { lier = 2; lier_k = k; }           // :135
      for (int i = 0; i < np_; ++i) {                                 // :137
        double v = 0.0;
        for (int j = 0; j < d; ++j) v += alpha_in[(size_t)j * np_ + i] * beta[j];
        res[i] = v - beta_in[i];
      }
      double mar = orc_mad(res.data(), np_);                          // :138
      bool failed = false;
      if (mar > DBL_EPSILON) {                                        // :139
        int e1 = orc_irls(d, np_, &mar, 5, 1, omega.data(), sol.data(), beta_in.data(), alpha_in.data(),
                          res.data(), dgelss);                        // :140-143
        if (e1 != 0) {
          if (k >= lier_k) { lier = 3; lier_k = k; }                  // :144-146 (CYCLE: rw untouched)
          failed = true;
        } else {
          for (int j = 0; j < d; ++j) a[k * d + j] = sol[j];          // :148
          mar = orc_mad(res.data(), np_);                             // :150
          if (mar > DBL_EPSILON) {                                    // :151
            e1 = orc_irls(d, np_, &mar, 5, 2, omega.data(), sol.data(), beta_in.data(), alpha_in.data(),
                          res.data(), dgelss);                        // :152-155
            if (e1 != 0) {
              if (k >= lier_k) { lier = 3; lier_k = k; }              // :156-158
              failed = true;
            } else {
              for (int j = 0; j < d; ++j) a[k * d + j] = sol[j];      // :160
            }
          }
        }
      }
      if (!failed) {
        for (int i = 0; i < np_; ++i) {                               // :164-174
          if (omega[i] < 0.8) {
            rw[k] = i == 0 ? dist[0] / 2.0 : (dist[i - 1] + dist[i]) / 2.0;
            break;
          }
        }
      }
      if (omega_out)
        for (int i = 0; i < np_; ++i) omega_out[k * np_ + i] = omega[i];
    }
#pragma omp critical
    {
      diam = std::max(diam, ldiam);
      if (lier_k > ier_k) {
        ier = lier;
        ier_k = lier_k;
      }
    }
  }
  const double half = std::sqrt(diam) / 2.0;                          // :177
  for (int64_t k = 0; k < n; ++k) rw[k] = std::min(half, rw[k]);
  return ier;
}

}  // extern "C"
