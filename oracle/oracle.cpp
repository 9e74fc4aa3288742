// oracle.cpp -- CPU restatement of util.approximate's distance-based predictors.
//
// TEST INFRASTRUCTURE ONLY.  Nothing under util_b200/ may import, link or call
// this file; it is the checker used by tests/, __graft_entry__.smoke() and the
// cpu_baseline / --impl reference legs of bench.py.
//
// Parity pinning: the reference holds no golden vectors for this path
// (SURVEY.md section 4).  The kNN part is pinned against the real
// util.approximate.NearestNeighbor run in the build container
// (tests/golden/make_golden.py -> tests/golden/nn_*.npz).  The Fortran routines
// (WEIGHT, SHEPMOD*, LSHEP*, PREDICT) cannot be compiled here (no gfortran), so
// for them parity is UNPINNED beyond the invariants in tests/test_oracle.py.
//
// Build: g++ -O2 -fno-fast-math -ffp-contract=off -fopenmp -shared -fPIC
// (source-order fp64 sums, separate multiply/add roundings: the reference's
// gfortran -O3 build targets baseline x86-64 without FMA).
//
// All arrays are row-major (point index major) doubles unless stated; indices
// are 0-based int64 at this boundary (the Fortran code is 1-based int32).

#include "oracle_common.h"


extern "C" {

void orc_set_threads(int t) {
#ifdef _OPENMP
  if (t > 0) omp_set_num_threads(t);
#else
  (void)t;
#endif
}

int orc_max_threads() {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

// ---------------------------------------------------------------------------
// kNN: what NearestNeighbor._predict asks of BallTree.query
// (nearest_neighbor/__init__.py:75-76): the k nearest rows by Euclidean
// distance, ascending, ties broken by lowest index (BASELINE.json).
// idx (q,k) int64, d2 (q,k) squared distances (BallTree returns sqrt of these).
// ---------------------------------------------------------------------------
int orc_knn(const double* x, int64_t n, int d, const double* z, int64_t q, int k,
            int64_t* idx, double* d2) {
  if (k < 1 || k > n || d < 1) return 1;
#pragma omp parallel for schedule(dynamic, 16)
  for (int64_t p = 0; p < q; ++p) {
    std::vector<Cand> c(k);
    knn_one(x, n, d, z + p * d, k, -1, c.data());
    for (int j = 0; j < k; ++j) {
      idx[p * k + j] = c[j].i;
      if (d2) d2[p * k + j] = c[j].d2;
    }
  }
  return 0;
}

// ---------------------------------------------------------------------------
// WEIGHT (shepard/weight.f90:1-34).  pts (n,d) row-major, pt (d) -> wts (n).
// NORM2 is restated as sqrt of the sequential sum of squares (gfortran's
// scaled NORM2 can differ by <= 1 ulp).
// ---------------------------------------------------------------------------
void orc_shepard_weight(const double* pts, int64_t n, int d, const double* pt, double* wts) {
  std::vector<double> dist(n);
  for (int64_t i = 0; i < n; ++i) dist[i] = std::sqrt(sqdist(pts + i * d, pt, d));  // :13-15
  const double eps = std::sqrt(DBL_EPSILON);                                         // :17
  double mn = dist[0], mx = dist[0];
  for (int64_t i = 1; i < n; ++i) {
    mn = std::min(mn, dist[i]);
    mx = std::max(mx, dist[i]);
  }
  if (mx < eps) {  // :22-24
    for (int64_t i = 0; i < n; ++i) wts[i] = 1.0;
  } else if (mn < eps) {  // :25-28
    for (int64_t i = 0; i < n; ++i) wts[i] = dist[i] < eps ? 1.0 : 0.0;
  } else {  // :29-31
    for (int64_t i = 0; i < n; ++i) wts[i] = 1.0 / (dist[i] * dist[i]);
  }
  double s = 0.0;  // :34
  for (int64_t i = 0; i < n; ++i) s += wts[i];
  for (int64_t i = 0; i < n; ++i) wts[i] = wts[i] / s;
}

// ---------------------------------------------------------------------------
// SHEPMOD (modified_shepard.f95:10-60): RW(K) = distance from node K to its
// (M+1)-th nearest other node.  x (n,d) row-major.  ier: 0 ok, 1 too few.
// ---------------------------------------------------------------------------
int orc_shepmod_fit(const double* x, int64_t n, int d, double* rw) {
  if (d < 1 || n <= d + 1) return 1;  // :37-40
  const int np_ = d + 1;              // :42
#pragma omp parallel for schedule(dynamic, 16)
  for (int64_t k = 0; k < n; ++k) {
    std::vector<Cand> c(np_);
    knn_one(x, n, d, x + k * d, np_, k, c.data());  // :45-55 (SORT keeps the NP smallest)
    rw[k] = std::sqrt(c[np_ - 1].d2);               // :57
  }
  return 0;
}

// ---------------------------------------------------------------------------
// SHEPMODVAL (modified_shepard.f95:83-176), one query.  wts (n) dense.
// Returns IER (0 normal, 1 fallback to the M+1 nearest).
// ---------------------------------------------------------------------------
int orc_shepmod_val(const double* xp, int d, int64_t n, const double* x, const double* rw,
                    double* wts) {
  const int m1 = d + 1;
  std::vector<double> D(m1, 0.0);
  std::vector<int64_t> DI(m1, -1);
  int64_t cnt = 0;
  int j = 0;  // 0-based slot (Fortran J = 1)
  double temp = 0.0;
  for (int64_t i = 0; i < n; ++i) wts[i] = 0.0;  // :122
  for (int64_t k = 0; k < n; ++k) {
    double dist = std::sqrt(sqdist(xp, x + k * d, d));  // :132
    if (dist < rw[k]) {
      if (rw[k] - dist == rw[k]) {  // :136-140 exact hit (absorption test)
        for (int64_t i = 0; i < n; ++i) wts[i] = 0.0;
        wts[k] = 1.0;
        return 0;
      }
      double t = (rw[k] - dist) / (rw[k] * dist);  // :143
      wts[k] = t * t;
      ++cnt;
    } else {
      double w = 1.0 / dist;  // :150
      if (w > temp) {         // :152-158
        D[j] = w;
        DI[j] = k;
        j = (int)(std::min_element(D.begin(), D.end()) - D.begin());  // MINLOC: first minimum
        temp = D[j];
      }
    }
  }
  if (cnt == 0) {  // :163-170
    for (int i = 0; i < m1; ++i) D[i] = D[i] * D[i];
    int jj = j + 1;  // number of slots "assumed nonzero" (Fortran J)
    if (*std::min_element(D.begin(), D.end()) > 0.0) jj = m1;
    double s = 0.0;
    for (int i = 0; i < m1; ++i) s += D[i];
    for (int i = 0; i < jj; ++i)
      if (DI[i] >= 0) wts[DI[i]] = D[i] / s;
    return 1;
  }
  double s = 0.0;  // :172 SUM(WTS(1:N)) in index order
  for (int64_t i = 0; i < n; ++i) s += wts[i];
  for (int64_t i = 0; i < n; ++i) wts[i] = wts[i] / s;
  return 0;
}

// Batched wrapper = ShepMod._predict (modified_shepard/__init__.py:39-58):
// keeps wts > 0 in ascending index order.  ids/w are (q,cap); cnt (q) holds the
// true count (entries beyond cap are dropped, cnt still reports them).
int orc_shepmod_predict(const double* x, int64_t n, int d, const double* rw, const double* z,
                        int64_t q, int cap, int64_t* ids, double* w, int32_t* cnt, int32_t* ier) {
#pragma omp parallel
  {
    std::vector<double> wts(n);
#pragma omp for schedule(dynamic, 4)
    for (int64_t p = 0; p < q; ++p) {
      ier[p] = orc_shepmod_val(z + p * d, d, n, x, rw, wts.data());
      int c = 0;
      for (int64_t i = 0; i < n; ++i)
        if (wts[i] > 0) {
          if (c < cap) {
            ids[p * cap + c] = i;
            w[p * cap + c] = wts[i];
          }
          ++c;
        }
      cnt[p] = c;
    }
  }
  return 0;
}

}  // extern "C"
