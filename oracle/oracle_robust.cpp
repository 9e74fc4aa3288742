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

// MAD as RIPPLE calls it (f95:602): MAD(RES(1:M+1), NP) -- the order statistics are taken at NP/2 (+1)
// of an array that holds only M+1 values (QUANTILE works on the array it is handed, MAD's NMID on NP).
static double mad_np(const double* res, int len, int np_) {
  const double factor = 1.0 / 0.6745;
  const int nmid = np_ / 2;
  std::vector<double> dev(len);
  if (np_ % 2 == 0) {
    const double med = (orc_quantile(res, len, nmid) + orc_quantile(res, len, nmid + 1)) / 2.0;
    for (int i = 0; i < len; ++i) dev[i] = std::fabs(res[i] - med);
    return factor * (orc_quantile(dev.data(), len, nmid) + orc_quantile(dev.data(), len, nmid + 1)) / 2.0;
  }
  const double med = orc_quantile(res, len, nmid + 1);
  for (int i = 0; i < len; ++i) dev[i] = std::fabs(res[i] - med);
  return factor * orc_quantile(dev.data(), len, nmid + 1);
}

// RIPPLE (f95:474-665).  x (n,d) row-major, a (n,d) row-major, rw (n).  Returns IER 0 / 1 / 2.
// The least-median-style start: for each of the NP chains of M+3 points grown through the neighbour
// lists, every way of leaving out two of its points (never the first) gives an (M+1) x M system; the
// one with the smallest residual component wins, then two IRLS stages on the neighbour set.
int orc_ripple(const double* x, int64_t n, int d, const double* f, double* a, double* rw, dgelss_fn dgelss) {
  if (d < 1 || n <= d + 3) return 1;                                             // :520-523
  const int np2 = (int)std::min<int64_t>(n - 1, (3 * d + 5) / 2);               // :525
  const int np0 = (int)std::min<int64_t>(n - 1, (3 * d + 1) / 2);               // :539
  const int mc = d + 3;                                                          // chain length
  std::vector<int64_t> S((size_t)n * np2);
  double diam = 0.0;
#pragma omp parallel
  {
    std::vector<Cand> c(np2);
    double ldiam = 0.0;
#pragma omp for schedule(dynamic, 16)
    for (int64_t k = 0; k < n; ++k) {                                            // :526-537
      double mx;
      knn_one(x, n, d, x + k * d, np2, k, c.data(), &mx);
      ldiam = std::max(ldiam, mx);
      for (int i = 0; i < np2; ++i) S[(size_t)k * np2 + i] = c[i].i;
    }
#pragma omp critical
    diam = std::max(diam, ldiam);
  }
  int ier = 0;
  int64_t ier_k = -1;
#pragma omp parallel
  {
    const int npmax = np0 + d + 1;
    std::vector<int64_t> D((size_t)np0 * mc), keep(d + 1), best(d + 1);
    std::vector<double> alpha_in((size_t)npmax * d), beta_in(npmax), alpha((size_t)(d + 1) * d), beta(d + 1), sigma(d),
        res(npmax), omega(npmax), sol(d), dist(d + 1), rows((size_t)mc * d), rb(mc);
    int lwork = 5 * d + 2 + 64 * (2 * d + 2);
    std::vector<double> work(lwork);
    int lier = 0;
    int64_t lier_k = -1;
#pragma omp for schedule(dynamic, 8)
    for (int64_t k = 0; k < n; ++k) {
      for (int i = 0; i < np0; ++i) D[(size_t)i * mc] = S[(size_t)k * np2 + i];  // :540
      for (int j = 1; j < mc; ++j)                                               // :541-548 (J = j+1)
        for (int i = 0; i < np0; ++i) {
          const int64_t prev = D[(size_t)i * mc + j - 1];
          for (int l = 0; l <= j; ++l) {                                         // L = 1..J: the last admissible L stays
            const int64_t cand = S[(size_t)prev * np2 + l];
            bool fresh = cand != k;
            for (int t = 0; t < j && fresh; ++t) fresh = D[(size_t)i * mc + t] != cand;
            if (fresh) D[(size_t)i * mc + j] = cand;
          }
        }
      double lse = std::numeric_limits<double>::max();                           // :549
      for (int j = 0; j < np0; ++j) {                                            // :550
        for (int i = 0; i < mc; ++i) {                                           // :551-554
          const int64_t p = D[(size_t)j * mc + i];
          rb[i] = f[p] - f[k];
          for (int t = 0; t < d; ++t) rows[(size_t)i * d + t] = x[p * d + t] - x[k * d + t];
        }
        for (int i = 2; i < mc; ++i)                                             // I = 3..M+3
          for (int l = 1; l < i; ++l) {                                          // L = 2..I-1
            int r = 0;
            for (int t = 0; t < mc; ++t) {                                       // :559-564: rows L and I left out
              if (t == l || t == i) continue;
              keep[r] = D[(size_t)j * mc + t];
              beta[r] = rb[t];
              for (int cdim = 0; cdim < d; ++cdim) alpha[(size_t)cdim * (d + 1) + r] = rows[(size_t)t * d + cdim];
              ++r;
            }
            int m_ = d + 1, n_ = d, nrhs = 1, lda = d + 1, ldb = d + 1, rank = 0, info = 0;
            double rc = (d + 1) * DBL_EPSILON;                                   // :555
            dgelss(&m_, &n_, &nrhs, alpha.data(), &lda, beta.data(), &ldb, sigma.data(), &rc, &rank, work.data(),
                   &lwork, &info);                                               // :565-566
            if (std::fabs(beta[d]) < lse) {                                      // :567-573
              lse = std::fabs(beta[d]);
              for (int t = 0; t < d; ++t) a[k * d + t] = beta[t];
              best = keep;
            }
          }
      }
      // :577-583 the winning M+1 points, their squared distances and residuals
      for (int i = 0; i <= d; ++i) {
        const int64_t p = best[i];
        beta_in[i] = f[p] - f[k];
        double s2 = 0.0;
        for (int t = 0; t < d; ++t) {
          const double v = x[p * d + t] - x[k * d + t];
          alpha_in[(size_t)t * npmax + i] = v;
          s2 += v * v;
        }
        dist[i] = s2;
      }
      for (int i = 0; i <= d; ++i) {
        double v = 0.0;
        for (int t = 0; t < d; ++t) v += alpha_in[(size_t)t * npmax + i] * a[k * d + t];
        res[i] = v - beta_in[i];
      }
      double mar = std::max(mad_np(res.data(), d + 1, np0), 1.0e-6);             // :584
      int np_ = np0;
      for (int i = 0; i < np0; ++i) {                                            // :585-588
        const int64_t p = S[(size_t)k * np2 + i];
        beta_in[i] = f[p] - f[k];
        for (int t = 0; t < d; ++t) alpha_in[(size_t)t * npmax + i] = x[p * d + t] - x[k * d + t];
      }
      double rwk = 0.0;                                                          // :589
      for (int t = 0; t < d; ++t) rwk += alpha_in[(size_t)t * npmax + np0 - 1] * alpha_in[(size_t)t * npmax + np0 - 1];
      {  // :590 SORT(DIST, IDIST, M+1, M+1): stable insertion, ties keep the earlier entry
        std::vector<int> ord(d + 1);
        std::iota(ord.begin(), ord.end(), 0);
        std::stable_sort(ord.begin(), ord.end(), [&](int p, int q) { return dist[p] < dist[q]; });
        std::vector<double> ds(d + 1);
        std::vector<int64_t> is(d + 1);
        for (int i = 0; i <= d; ++i) {
          ds[i] = dist[ord[i]];
          is[i] = best[ord[i]];
        }
        for (int i = 0; i <= d; ++i) {                                           // :591-598
          bool among = false;
          for (int t = 0; t < np0; ++t) among = among || S[(size_t)k * np2 + t] == is[i];
          if (ds[i] > rwk || (ds[i] == rwk && !among)) {
            beta_in[np_] = f[is[i]] - f[k];
            for (int t = 0; t < d; ++t) alpha_in[(size_t)t * npmax + np_] = x[is[i] * d + t] - x[k * d + t];
            ++np_;
          }
        }
      }
      rw[k] = std::sqrt(rwk);                                                    // :599
      // IRLS wants leading dimension = row count: compact the columns to np_ rows
      std::vector<double> ain((size_t)np_ * d);
      for (int t = 0; t < d; ++t)
        for (int i = 0; i < np_; ++i) ain[(size_t)t * np_ + i] = alpha_in[(size_t)t * npmax + i];
      for (int i = 0; i < np_; ++i) {                                            // :600
        double v = 0.0;
        for (int t = 0; t < d; ++t) v += ain[(size_t)t * np_ + i] * a[k * d + t];
        res[i] = v - beta_in[i];
      }
      int e1 = orc_irls(d, np_, &mar, 5, 1, omega.data(), sol.data(), beta_in.data(), ain.data(), res.data(), dgelss);  // :601-604
      if (e1 != 0) {
        if (k >= lier_k) { lier = 2; lier_k = k; }                               // :605-607
        continue;
      }
      for (int t = 0; t < d; ++t) a[k * d + t] = sol[t];                         // :609
      e1 = orc_irls(d, np_, &mar, 5, 2, omega.data(), sol.data(), beta_in.data(), ain.data(), res.data(), dgelss);      // :611-614
      if (e1 != 0) {
        if (k >= lier_k) { lier = 2; lier_k = k; }                               // :615-617
        continue;
      }
      for (int t = 0; t < d; ++t) a[k * d + t] = sol[t];                         // :619
      for (int i = 0; i < np_; ++i) {                                            // :621-634
        if (omega[i] < 0.8) {
          auto len = [&](int r) {
            double s2 = 0.0;
            for (int t = 0; t < d; ++t) s2 += ain[(size_t)t * np_ + r] * ain[(size_t)t * np_ + r];
            return std::sqrt(s2);
          };
          rw[k] = i == 0 ? len(0) / 2.0 : (len(i - 1) + len(i)) / 2.0;
          break;
        }
      }
    }
#pragma omp critical
    if (lier_k > ier_k) {
      ier = lier;
      ier_k = lier_k;
    }
  }
  const double half = std::sqrt(diam) / 2.0;                                     // :636
  for (int64_t k = 0; k < n; ++k) rw[k] = std::min(half, rw[k]);
  return ier;
}

}  // extern "C"
