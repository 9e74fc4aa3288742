// oracle_common.h -- shared helpers of the CPU oracle (TEST INFRASTRUCTURE ONLY,
// see oracle.cpp for the rules on who may use oracle/).
#pragma once
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <numeric>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

// Squared distance with the reference's operation order: direct differences,
// summed over ascending dimension (sklearn metrics/_dist_metrics.pxd rdist;
// modified_shepard.f95:51 DOT_PRODUCT(X(:,I)-X(:,K), X(:,I)-X(:,K))).
inline double sqdist(const double* a, const double* b, int d) {
  double s = 0.0;
  for (int j = 0; j < d; ++j) {
    double t = a[j] - b[j];
    s += t * t;
  }
  return s;
}

inline double dot(const double* a, const double* b, int d) {
  double s = 0.0;
  for (int j = 0; j < d; ++j) s += a[j] * b[j];
  return s;
}

struct Cand {
  double d2;
  int64_t i;
};
inline bool cand_less(const Cand& a, const Cand& b) {
  return a.d2 < b.d2 || (a.d2 == b.d2 && a.i < b.i);
}

// k smallest (d2, index) pairs of one query against all data, ascending.
// skip >= 0 excludes that data index (the "other nodes" scans of the fits).
inline void knn_one(const double* x, int64_t n, int d, const double* z, int k, int64_t skip,
             Cand* out, double* maxd2 = nullptr) {
  std::vector<Cand> heap;  // max-heap on (d2, i)
  heap.reserve(k + 1);
  double mx = 0.0;
  for (int64_t i = 0; i < n; ++i) {
    if (i == skip) continue;
    Cand c{sqdist(x + i * d, z, d), i};
    if (c.d2 > mx) mx = c.d2;
    if ((int)heap.size() < k) {
      heap.push_back(c);
      std::push_heap(heap.begin(), heap.end(), cand_less);
    } else if (cand_less(c, heap.front())) {
      std::pop_heap(heap.begin(), heap.end(), cand_less);
      heap.back() = c;
      std::push_heap(heap.begin(), heap.end(), cand_less);
    }
  }
  std::sort_heap(heap.begin(), heap.end(), cand_less);
  for (size_t j = 0; j < heap.size(); ++j) out[j] = heap[j];
  if (maxd2) *maxd2 = mx;
}

}  // namespace
