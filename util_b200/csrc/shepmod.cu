// shepmod.cu -- modified Shepard (ShepMod) fit and predict.
//
// Fit  = SHEPMOD (modified_shepard.f95:10-60): rw[K] is the distance from node K
//        to its (M+1)-th nearest OTHER node -> one self-excluding kNN with
//        k = M+1 (tensor or exact path) and a square root of the last column.
// Eval = SHEPMODVAL over the whole query batch (modified_shepard.f95:83-176) fused
//        with ShepMod._predict's `wts > 0` compaction
//        (modified_shepard/__init__.py:44-55) and WeightedApproximator.predict's
//        weighted sum (base.py:169).
//
// Eval kernel: one thread per query, nodes stream through shared memory in
// ascending index order, so the hit list, SUM(WTS) (f95:172) and the weighted
// sum are accumulated in exactly the reference's order; distances use direct
// differences in dimension order without FMA contraction.  Work per pair is
// 3M+2 fp64 operations: this stage is fp64-pipe bound, not HBM bound.
#include <math_constants.h>

#include "common.cuh"

namespace b200 {

int knn_dispatch(b200_model* m, const double* z_dev, int64_t q, int k, int64_t self_offset,
                 int32_t* idx_dev, double* d2_dev, cudaStream_t st);

namespace {

__global__ void absmax_kernel(const double* __restrict__ x, int64_t tot, unsigned long long* __restrict__ out) {
  double m = 0.0;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < tot; e += (int64_t)gridDim.x * blockDim.x)
    m = fmax(m, fabs(x[e]));  // NaN coordinates are ignored here; the exact path sees them
  for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0) atomicMax(out, (unsigned long long)__double_as_longlong(m));
}

__global__ void f32_image_kernel(const double* __restrict__ x, int64_t n, int d, int dsf, float* __restrict__ out) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n * dsf) return;
  const int64_t i = e / dsf;
  const int j = (int)(e - i * dsf);
  out[e] = j < d ? (float)x[i * d + j] : 0.0f;
}

__global__ void rw_from_knn_kernel(const double* __restrict__ d2, int64_t n, int k,
                                   double* __restrict__ rw) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) rw[i] = sqrt(d2[i * k + (k - 1)]);  // modified_shepard.f95:57
}

constexpr int SM_QB = 128;   // queries per block
constexpr int SM_FB = 72;    // fallback slots kept per thread (M+1 <= SM_FB)

template <int DMAX>  // query coordinates live in registers: DMAX >= d, loops fully unrolled
__global__ void __launch_bounds__(SM_QB)
shepmod_eval_kernel(const double* __restrict__ x, const float* __restrict__ x32, const double* __restrict__ rw,
                    int64_t n, int d,
                    const double* __restrict__ z, int64_t q, int cap, int tile_pts,
                    int32_t* __restrict__ ids, double* __restrict__ wts, int32_t* __restrict__ cnt,
                    int32_t* __restrict__ ier, const double* __restrict__ y, int r,
                    double* __restrict__ out, double x_absmax) {
  extern __shared__ double sm[];
  double* xs = sm;                        // [tile_pts][d]
  double* rs = xs + (size_t)tile_pts * d; // [tile_pts]
  // squared-radius band: outside [lo, hi) the test sqrt(a) < rk is decided without the square root
  double* r2lo = rs + tile_pts;
  double* r2hi = r2lo + tile_pts;
  // fp32 prefilter (common.cuh): float copies of the tile, rows padded to a multiple of 4
  constexpr int D4 = (DMAX + 3) / 4;
  const int dsf = ((d + 3) / 4) * 4;
  float* xf = reinterpret_cast<float*>(r2hi + tile_pts);  // [tile_pts][dsf]
  float* rkf = xf + (size_t)tile_pts * dsf;               // [tile_pts] prefilter threshold of the node's radius
  const int t = threadIdx.x;
  const int64_t p = (int64_t)blockIdx.x * SM_QB + t;
  const bool live = p < q;
  double zq[DMAX];
  float zf[D4 * 4];
  double cm = x_absmax;
#pragma unroll
  for (int j = 0; j < DMAX; ++j) {
    zq[j] = (live && j < d) ? z[p * d + j] : 0.0;
    cm = fmax(cm, fabs(zq[j]));
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
  const int m1 = d + 1;
  double D[SM_FB];
  int DI[SM_FB];
  for (int i = 0; i < m1; ++i) { D[i] = 0.0; DI[i] = -1; }  // :124-125
  int J = 0;
  double temp = 0.0;   // :126
  int c = 0;           // nodes whose radius contains the query (I in the source)
  double s = 0.0;      // running SUM(WTS) in index order
  int hit = -1;        // exact-hit node
  int32_t* myid = ids + p * cap;
  double* myw = wts + p * cap;

  for (int64_t base = 0; base < n; base += tile_pts) {
    const int tc = (int)imin64(tile_pts, n - base);
    __syncthreads();
    load_tile_f64_f32(x + base * d, x32 + base * dsf, tc, d, dsf, xs, xf, t, SM_QB);
    for (int e = t; e < tc; e += SM_QB) {
      const double rk = rw[base + e];
      rs[e] = rk;
      r2lo[e] = rk * rk * (1.0 - 1e-14);
      r2hi[e] = rk * rk * (1.0 + 1e-14);
      rkf[e] = pre_thr(pre_up(rk), g0, eps1);
    }
    __syncthreads();
    if (!live || hit >= 0) continue;
    // Nodes four at a time: four independent fp32 distance chains (the loop is latency bound), then
    // the few nodes the prefilter cannot rule out take the exact path in index order.  A node is
    // ruled out when it is certainly outside its radius and, before the first contributor, certainly
    // not among the M+1 nearest kept for the fallback; thresholds taken at the start of a group are
    // only ever looser than the current ones.
    for (int i0 = 0; i0 < tc && hit < 0; i0 += 4) {
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
      while (need) {
      const int i = i0 + __ffs(need) - 1;
      need &= need - 1;
      const double* xr = xs + (size_t)i * d;
      double a = 0.0;
#pragma unroll
      for (int j = 0; j < DMAX; ++j)
        if (j < d) {
          const double df = __dsub_rn(zq[j], xr[j]);
          a = __dadd_rn(a, __dmul_rn(df, df));
        }
      const double rk = rs[i];
      const bool inside = a < r2lo[i] ? true : (a >= r2hi[i] ? false : sqrt(a) < rk);  // DIST < RW(K), :134
      if (inside) {
        const double dist = sqrt(a);  // :132
        if (__dsub_rn(rk, dist) == rk) {  // :136-140
          hit = (int)(base + i);
          break;  // leaves the group; the outer loops test hit
        }
        const double u = __ddiv_rn(__dsub_rn(rk, dist), __dmul_rn(rk, dist));  // :143
        const double w = __dmul_rn(u, u);
        if (c < cap) {
          myid[c] = (int)(base + i);
          myw[c] = w;
        }
        s = __dadd_rn(s, w);
        ++c;
      } else if (c == 0) {
        const double w = __ddiv_rn(1.0, sqrt(a));  // :150
        if (w > temp) {                          // :152-158
          D[J] = w;
          DI[J] = (int)(base + i);
          int jm = 0;
          double dm = D[0];
          for (int u2 = 1; u2 < m1; ++u2)
            if (D[u2] < dm) { dm = D[u2]; jm = u2; }  // MINLOC: first minimum
          J = jm;
          temp = dm;
          // a node can still enter the fallback set only with 1/dist > temp, i.e. dist < 1/temp
          thr_fb = temp > 0.0 ? pre_thr(pre_up(1.0 / temp), g0, eps1) : CUDART_INF_F;
        }
      }
      }
    }
  }
  if (!live) return;
  int e = 0;
  if (hit >= 0) {  // one-hot (:137-139)
    myid[0] = hit;
    myw[0] = 1.0;
    c = 1;
  } else if (c == 0) {  // :163-170 original Shepard on the M+1 kept nodes
    e = 1;
    double sd = 0.0;
    for (int i = 0; i < m1; ++i) { D[i] = __dmul_rn(D[i], D[i]); }
    for (int i = 0; i < m1; ++i) sd = __dadd_rn(sd, D[i]);
    // ascending node index, as `base[wts > 0]` returns them
    for (int i = 1; i < m1; ++i) {
      const double dv = D[i];
      const int iv = DI[i];
      int j = i;
      while (j > 0 && DI[j - 1] > iv) { D[j] = D[j - 1]; DI[j] = DI[j - 1]; --j; }
      D[j] = dv;
      DI[j] = iv;
    }
    for (int i = 0; i < m1; ++i) {
      if (DI[i] < 0) continue;
      const double w = __ddiv_rn(D[i], sd);
      if (w > 0) {
        if (c < cap) { myid[c] = DI[i]; myw[c] = w; }
        ++c;
      }
    }
  } else {  // :172
    const int lim = c < cap ? c : cap;
    int kept = 0;
    for (int i = 0; i < lim; ++i) {
      const double w = __ddiv_rn(myw[i], s);
      if (w > 0) {  // a weight can underflow to zero; the wrapper drops those
        myid[kept] = myid[i];
        myw[kept] = w;
        ++kept;
      }
    }
    if (c <= cap) c = kept;
  }
  cnt[p] = c;
  ier[p] = e;
  if (out) {
    const int lim = c < cap ? c : cap;
    for (int col = 0; col < r; ++col) {
      double acc = 0.0;  // base.py:169
      for (int i = 0; i < lim; ++i) acc = __dadd_rn(acc, __dmul_rn(y[(int64_t)myid[i] * r + col], myw[i]));
      out[p * r + col] = acc;
    }
  }
}

// SHEPMODVAL on a candidate list (modified_shepard.f95:83-176): one warp per query.  cand (q, 64) holds
// fitted-row numbers of a SUPERSET of the nodes whose radius contains the query (the tensor-core radius
// filter, knn_tc.cu); the exact test DIST < RW(K), the exact-hit rule (lowest K wins, :136-140), the
// weights (:143) and their sums run on those rows in ascending node order, which is the order the
// reference's scan meets them in -- nodes outside the list contribute nothing to any of it.  A query with
// no node inside a radius (IER = 1: the M+1 nearest nodes take over, :147-170) or with an overflowed list
// is appended to `redo` for the dense scan.
__global__ void __launch_bounds__(256)
shepmod_cand_kernel(const double* __restrict__ x, const double* __restrict__ rw, int d, const double* __restrict__ z,
                    int64_t q, const int32_t* __restrict__ cand, const int32_t* __restrict__ cand_cnt, int cap,
                    int32_t* __restrict__ ids, double* __restrict__ wts, int32_t* __restrict__ cnt,
                    int32_t* __restrict__ ier, const double* __restrict__ y, int r, double* __restrict__ out,
                    int32_t* __restrict__ redo, int32_t* __restrict__ redo_cnt) {
  constexpr int W = 64;
  __shared__ int s_row[8][W];
  __shared__ double s_w[8][W];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t p = (int64_t)blockIdx.x * 8 + w;
  if (p >= q) return;
  const int nc = cand_cnt[p];
  if (nc > W || nc <= 0) {  // overflow, or certainly no radius contains the query
    if (lane == 0) redo[atomicAdd(redo_cnt, 1)] = (int32_t)p;
    return;
  }
  // rank the candidate rows ascending (each lane owns entries lane and lane + 32)
  int mine[2];
#pragma unroll
  for (int u = 0; u < 2; ++u) {
    const int j = lane + 32 * u;
    mine[u] = j < nc ? cand[p * W + j] : 0x7fffffff;
  }
  int rank[2] = {0, 0};
  for (int i = 0; i < nc; ++i) {
    const int v = __shfl_sync(0xffffffffu, i < 32 ? mine[0] : mine[1], i & 31);
#pragma unroll
    for (int u = 0; u < 2; ++u) rank[u] += v < mine[u] ? 1 : 0;
  }
#pragma unroll
  for (int u = 0; u < 2; ++u)
    if (lane + 32 * u < nc) s_row[w][rank[u]] = mine[u];
  __syncwarp();
  // exact distances and weights, two sorted entries per lane
  const double* zp = z + p * d;
  int hit = 0x7fffffff;
#pragma unroll
  for (int u = 0; u < 2; ++u) {
    const int j = lane + 32 * u;
    double wv = -1.0;  // not inside the node's radius
    if (j < nc) {
      const int node = s_row[w][j];
      const double a = sqdist_seq(zp, x + (int64_t)node * d, d);
      const double dist = sqrt(a);  // :132
      const double rk = rw[node];
      if (dist < rk) {  // :134
        if (__dsub_rn(rk, dist) == rk) hit = min(hit, node);  // :136
        const double t = __ddiv_rn(__dsub_rn(rk, dist), __dmul_rn(rk, dist));  // :143
        wv = __dmul_rn(t, t);
      }
    }
    if (j < W) s_w[w][j] = wv;
  }
  for (int o = 16; o > 0; o >>= 1) hit = min(hit, __shfl_xor_sync(0xffffffffu, hit, o));
  __syncwarp();
  // An exact hit returns one-hot -- but only if the scan reaches it before... it always does: the scan
  // stops at the FIRST node (lowest index) that is an exact hit, whatever came before (:137-139).
  if (lane == 0) {
    int c = 0;
    int e = 0;
    bool fallback = false;
    int32_t* myid = ids ? ids + p * cap : nullptr;
    double* myw = wts ? wts + p * cap : nullptr;
    double acc_w[8];
    int acc_id[8];
    (void)acc_w;
    (void)acc_id;
    if (hit != 0x7fffffff) {
      if (myid) myid[0] = hit;
      if (myw) myw[0] = 1.0;
      c = 1;
      if (out)
        for (int col = 0; col < r; ++col) out[p * r + col] = __dadd_rn(0.0, __dmul_rn(y[(int64_t)hit * r + col], 1.0));
    } else {
      double s = 0.0;  // SUM(WTS) in node order (:172)
      int hits = 0;
      for (int j = 0; j < nc; ++j)
        if (s_w[w][j] >= 0.0) {
          s = __dadd_rn(s, s_w[w][j]);
          ++hits;
        }
      if (hits == 0) {
        fallback = true;
      } else {
        for (int col = 0; out && col < r; ++col) out[p * r + col] = 0.0;
        for (int j = 0; j < nc; ++j) {
          if (s_w[w][j] < 0.0) continue;
          const double wn = __ddiv_rn(s_w[w][j], s);
          if (wn > 0) {  // a weight can underflow to zero; the wrapper drops those
            const int node = s_row[w][j];
            if (c < cap) {
              if (myid) myid[c] = node;
              if (myw) myw[c] = wn;
            }
            if (out && c < cap)
              for (int col = 0; col < r; ++col)
                out[p * r + col] = __dadd_rn(out[p * r + col], __dmul_rn(y[(int64_t)node * r + col], wn));  // base.py:169
            ++c;
          }
        }
      }
    }
    if (fallback) {
      redo[atomicAdd(redo_cnt, 1)] = (int32_t)p;
    } else {
      cnt[p] = c;
      ier[p] = e;
    }
  }
}

// The complete SHEPMODVAL for the queries listed in `sel`, one block per query: 256 nodes per step in
// ascending order, every thread one node; the sequential parts -- hit list and SUM(WTS) in node order, the
// replacement of the fallback slots by MINLOC (:147-158), the stop at the first exact hit -- are carried
// out by thread 0 on the few nodes the step flags.  Used for the queries the radius filter hands back (no
// node inside a radius, or more than 64 candidates), where a thread-per-query scan would leave the GPU
// empty.  Square roots and divisions only where a test on the squared distance cannot decide.
constexpr int SM_BT = 256;
__global__ void __launch_bounds__(SM_BT)
shepmod_block_kernel(const double* __restrict__ x, const double* __restrict__ rw, int64_t n, int d,
                     const double* __restrict__ z, const int32_t* __restrict__ sel, int cap,
                     int32_t* __restrict__ ids, double* __restrict__ wts, int32_t* __restrict__ cnt,
                     int32_t* __restrict__ ier, const double* __restrict__ y, int r, double* __restrict__ out) {
  __shared__ __align__(16) double s_z[SM_FB];
  __shared__ double s_D[SM_FB], s_a[SM_BT], s_rk[SM_BT];
  __shared__ int s_DI[SM_FB];
  __shared__ unsigned s_in[SM_BT / 32], s_cd[SM_BT / 32];
  __shared__ int s_c, s_hit, s_J;
  __shared__ double s_temp, s_sum;
  const int t = threadIdx.x, lane = t & 31, wid = t >> 5;
  const int64_t p = sel[blockIdx.x];
  const int m1 = d + 1;
  for (int j = t; j < d; j += SM_BT) s_z[j] = z[p * d + j];
  for (int i = t; i < m1; i += SM_BT) {
    s_D[i] = 0.0;   // :124-125
    s_DI[i] = -1;
  }
  if (t == 0) {
    s_c = 0;
    s_hit = -1;
    s_J = 0;
    s_temp = 0.0;  // :126
    s_sum = 0.0;
  }
  __syncthreads();
  int32_t* myid = ids + p * cap;
  double* myw = wts + p * cap;
  for (int64_t base = 0; base < n; base += SM_BT) {
    const int64_t k = base + t;
    const int c0 = s_c;
    const double temp0 = s_temp;
    if (s_hit >= 0) break;  // uniform: written before the last barrier
    double a = CUDART_INF, rk = 0.0;
    if (k < n) {
      a = sqdist_seq_wide<true>(x + k * d, s_z, d);  // the row's loads in flight together; (x - z)^2 == (z - x)^2 bit for bit
      rk = rw[k];
    }
    const double r2 = rk * rk;
    bool inside = false;
    if (k < n && a < r2 * (1.0 + 1e-14)) inside = a < r2 * (1.0 - 1e-14) ? true : sqrt(a) < rk;  // :134
    const double lim2 = temp0 > 0.0 ? (1.0 / temp0) * (1.0 / temp0) * (1.0 + 1e-12) : CUDART_INF;
    const bool maybe = c0 == 0 && k < n && !inside && a < lim2;  // may enter the fallback slots: 1/dist > temp
    const unsigned mi = __ballot_sync(0xffffffffu, inside), mc = __ballot_sync(0xffffffffu, maybe);
    if (lane == 0) {
      s_in[wid] = mi;
      s_cd[wid] = mc;
    }
    s_a[t] = a;
    s_rk[t] = rk;
    if (!__syncthreads_or(inside || maybe)) continue;
    if (t == 0) {
      int c = s_c, J = s_J, hit = -1;
      double temp = s_temp, sum = s_sum;
      for (int wq = 0; wq < SM_BT / 32 && hit < 0; ++wq) {
        unsigned both = s_in[wq] | s_cd[wq];
        while (both) {
          const int b = __ffs(both) - 1;
          both &= both - 1;
          const int i = wq * 32 + b;
          const double dist = sqrt(s_a[i]);  // :132
          const double rki = s_rk[i];
          if ((s_in[wq] >> b) & 1u) {
            if (__dsub_rn(rki, dist) == rki) {  // :136-140
              hit = (int)(base + i);
              break;
            }
            const double u = __ddiv_rn(__dsub_rn(rki, dist), __dmul_rn(rki, dist));  // :143
            const double wv = __dmul_rn(u, u);
            if (c < cap) {
              myid[c] = (int)(base + i);
              myw[c] = wv;
            }
            sum = __dadd_rn(sum, wv);
            ++c;
          } else if (c == 0) {
            const double wv = __ddiv_rn(1.0, dist);  // :150
            if (wv > temp) {                         // :152-158
              s_D[J] = wv;
              s_DI[J] = (int)(base + i);
              int jm = 0;
              double dm = s_D[0];
              for (int u2 = 1; u2 < m1; ++u2)
                if (s_D[u2] < dm) { dm = s_D[u2]; jm = u2; }  // MINLOC: first minimum
              J = jm;
              temp = dm;
            }
          }
        }
      }
      s_c = c;
      s_J = J;
      s_temp = temp;
      s_sum = sum;
      s_hit = hit;
    }
    __syncthreads();
  }
  __syncthreads();
  if (t != 0) return;
  int c = s_c;
  const int hit = s_hit;
  const double s = s_sum;
  int e = 0;
  if (hit >= 0) {  // one-hot (:137-139)
    myid[0] = hit;
    myw[0] = 1.0;
    c = 1;
  } else if (c == 0) {  // :163-170 original Shepard on the M+1 kept nodes
    e = 1;
    double sd = 0.0;
    for (int i = 0; i < m1; ++i) s_D[i] = __dmul_rn(s_D[i], s_D[i]);
    for (int i = 0; i < m1; ++i) sd = __dadd_rn(sd, s_D[i]);
    for (int i = 1; i < m1; ++i) {  // ascending node index, as `base[wts > 0]` returns them
      const double dv = s_D[i];
      const int iv = s_DI[i];
      int j = i;
      while (j > 0 && s_DI[j - 1] > iv) { s_D[j] = s_D[j - 1]; s_DI[j] = s_DI[j - 1]; --j; }
      s_D[j] = dv;
      s_DI[j] = iv;
    }
    for (int i = 0; i < m1; ++i) {
      if (s_DI[i] < 0) continue;
      const double wn = __ddiv_rn(s_D[i], sd);
      if (wn > 0) {
        if (c < cap) { myid[c] = s_DI[i]; myw[c] = wn; }
        ++c;
      }
    }
  } else {  // :172
    const int lim = c < cap ? c : cap;
    int kept = 0;
    for (int i = 0; i < lim; ++i) {
      const double wn = __ddiv_rn(myw[i], s);
      if (wn > 0) {
        myid[kept] = myid[i];
        myw[kept] = wn;
        ++kept;
      }
    }
    if (c <= cap) c = kept;
  }
  cnt[p] = c;
  ier[p] = e;
  if (out) {
    const int lim = c < cap ? c : cap;
    for (int col = 0; col < r; ++col) {
      double acc = 0.0;  // base.py:169
      for (int i = 0; i < lim; ++i) acc = __dadd_rn(acc, __dmul_rn(y[(int64_t)myid[i] * r + col], myw[i]));
      out[p * r + col] = acc;
    }
  }
}

// rows of a compact (ns, w) array scattered to the rows listed in `sel` of a (q, w) array
template <class T>
__global__ void scatter_rows_kernel(const T* __restrict__ src, const int32_t* __restrict__ sel, int ns, int w,
                                    T* __restrict__ dst) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= (int64_t)ns * w) return;
  const int i = (int)(e / w), j = (int)(e - (int64_t)i * w);
  dst[(int64_t)sel[i] * w + j] = src[e];
}

}  // namespace

int data_absmax(b200_model* m, cudaStream_t st, double* out) {
  if (m->x_absmax < 0.0) {
    B200_TRY(m->flags.reserve(256));
    unsigned long long* bits = reinterpret_cast<unsigned long long*>(m->flags.as<char>() + 128);
    B200_CUDA(cudaMemsetAsync(bits, 0, 8, st));
    absmax_kernel<<<1024, 256, 0, st>>>(m->x, m->n * m->d, bits);
    B200_CUDA(cudaGetLastError());
    unsigned long long h = 0;
    B200_CUDA(cudaMemcpyAsync(&h, bits, 8, cudaMemcpyDeviceToHost, st));
    B200_CUDA(cudaStreamSynchronize(st));
    double v;
    memcpy(&v, &h, 8);
    m->x_absmax = (v == v && v < 1e300) ? v : 1e300;  // inf -> the prefilter switches itself off
    m->st.launches += 1;
  }
  *out = m->x_absmax;
  return 0;
}

int data_f32(b200_model* m, cudaStream_t st, const float** out) {
  const int dsf = ((m->d + 3) / 4) * 4;
  const size_t bytes = (size_t)m->n * dsf * 4;
  if (!m->x32.p) {
    B200_TRY(m->x32.reserve(bytes));
    const int64_t tot = m->n * dsf;
    f32_image_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(m->x, m->n, m->d, dsf, m->x32.as<float>());
    B200_CUDA(cudaGetLastError());
    m->st.launches += 1;
  }
  *out = m->x32.as<float>();
  return 0;
}

int shepmod_fit(b200_model* m, cudaStream_t st, int* ier) {
  const int64_t n = m->n;
  const int d = m->d;
  *ier = 0;
  if (d < 1 || n <= d + 1) {  // :37-40
    *ier = 1;
    return 0;
  }
  const int k = d + 1;  // :42
  B200_TRY(m->idx32.reserve((size_t)n * k * 4));
  B200_TRY(m->d2.reserve((size_t)n * k * 8));
  B200_TRY(m->rw.reserve((size_t)n * 8));
  B200_TRY(knn_dispatch(m, m->x, n, k, 0, m->idx32.as<int32_t>(), m->d2.as<double>(), st));
  rw_from_knn_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(m->d2.as<double>(), n, k, m->rw.as<double>());
  B200_CUDA(cudaGetLastError());
  m->st.launches += 1;
  m->shepmod_ready = true;
  return 0;
}

static int shepmod_dense(b200_model* m, const double* z_dev, int64_t q, int cap, int32_t* ids_dev,
                         double* wts_dev, int32_t* cnt_dev, int32_t* ier_dev, const double* y_dev, int r,
                         double* out_dev, cudaStream_t st);

int shepmod_predict(b200_model* m, const double* z_dev, int64_t q, int cap, int32_t* ids_dev,
                    double* wts_dev, int32_t* cnt_dev, int32_t* ier_dev, const double* y_dev, int r,
                    double* out_dev, cudaStream_t st) {
  if (!m->shepmod_ready) {
    set_error("shepmod_predict: model not fitted (call b200_shepmod_fit)");
    return 1;
  }
  if (q == 0) return 0;
  const int d = m->d;
  if (d + 1 > SM_FB) {
    set_error("shepmod_predict: dimension %d > %d unsupported", d, SM_FB - 1);
    return 1;
  }
  if (cap < d + 1) {
    set_error("shepmod_predict: cap=%d must be at least M+1=%d", cap, d + 1);
    return 1;
  }
  // Large problems: the radius test DIST < RW(K) is first run as a tensor-core filter (the same fused
  // contraction as the kNN path with a per-node threshold), then SHEPMODVAL's arithmetic runs on the few
  // candidate nodes of each query; the dense scan below takes what the filter hands back.
  const bool filter = m->knn_path != B200_PATH_EXACT && shepmod_filter_supported(m) && (double)m->n * (double)q >= 2.5e8 &&
                      m->n >= 4096 && d >= 5;
  if (!filter) return shepmod_dense(m, z_dev, q, cap, ids_dev, wts_dev, cnt_dev, ier_dev, y_dev, r, out_dev, st);
  B200_TRY(m->part_idx.reserve((size_t)q * 64 * 4));
  B200_TRY(m->misc.reserve((size_t)q * 8 + 64));
  int32_t* cand_rows = m->part_idx.as<int32_t>();
  int32_t* cand_cnt = m->misc.as<int32_t>();
  int32_t* redo = cand_cnt + q;
  int32_t* redo_cnt = redo + q;
  B200_TRY(shepmod_filter(m, z_dev, q, cand_rows, cand_cnt, st));
  B200_CUDA(cudaMemsetAsync(redo_cnt, 0, 4, st));
  shepmod_cand_kernel<<<(unsigned)((q + 7) / 8), 256, 0, st>>>(m->x, m->rw.as<double>(), d, z_dev, q, cand_rows, cand_cnt, cap,
                                                             ids_dev, wts_dev, cnt_dev, ier_dev, y_dev, r,
                                                             y_dev ? out_dev : nullptr, redo, redo_cnt);
  B200_CUDA(cudaGetLastError());
  m->st.launches += 1;
  int nr = 0;
  B200_CUDA(cudaMemcpyAsync(&nr, redo_cnt, 4, cudaMemcpyDeviceToHost, st));
  B200_CUDA(cudaStreamSynchronize(st));
  m->st.uncertified += nr;
  if (nr > 0) {
    // queries the filter hands back (no radius holds them, or more than 64 candidates): the full scan
    shepmod_block_kernel<<<(unsigned)nr, SM_BT, 0, st>>>(m->x, m->rw.as<double>(), m->n, d, z_dev, redo, cap, ids_dev, wts_dev,
                                                         cnt_dev, ier_dev, y_dev, r, y_dev ? out_dev : nullptr);
    B200_CUDA(cudaGetLastError());
    m->st.launches += 1;
  }
  return 0;
}

static int shepmod_dense(b200_model* m, const double* z_dev, int64_t q, int cap, int32_t* ids_dev,
                         double* wts_dev, int32_t* cnt_dev, int32_t* ier_dev, const double* y_dev, int r,
                         double* out_dev, cudaStream_t st) {
  const int d = m->d;
  if (d + 1 > SM_FB) {
    set_error("shepmod_predict: dimension %d > %d unsupported", d, SM_FB - 1);
    return 1;
  }
  if (cap < d + 1) {
    set_error("shepmod_predict: cap=%d must be at least M+1=%d", cap, d + 1);
    return 1;
  }
  const int dsf = ((d + 3) / 4) * 4;
  const size_t per_pt = (size_t)(d + 3) * 8 + (size_t)(dsf + 1) * 4;
  int tile_pts = (int)((40 * 1024) / per_pt);  // small tiles: five blocks per SM hide the scan's latencies
  if (tile_pts > 512) tile_pts = 512;
  tile_pts &= ~3;  // whole groups of four; keeps the float rows 16-byte aligned behind the double arrays
  if (tile_pts < 4) tile_pts = 4;
  const size_t smem = (size_t)tile_pts * per_pt;
  const unsigned gb = (unsigned)((q + SM_QB - 1) / SM_QB);
  double absmax = 0.0;
  B200_TRY(data_absmax(m, st, &absmax));
  const float* x32 = nullptr;
  B200_TRY(data_f32(m, st, &x32));
#define SM_LAUNCH(DM)                                                                              \
  do {                                                                                             \
    auto kern = shepmod_eval_kernel<DM>;                                                           \
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    kern<<<gb, SM_QB, smem, st>>>(m->x, x32, m->rw.as<double>(), m->n, d, z_dev, q, cap, tile_pts, ids_dev, \
                                  wts_dev, cnt_dev, ier_dev, y_dev, r, y_dev ? out_dev : nullptr, absmax); \
  } while (0)
  if (d <= 2) SM_LAUNCH(2);
  else if (d <= 4) SM_LAUNCH(4);
  else if (d <= 8) SM_LAUNCH(8);
  else if (d <= 12) SM_LAUNCH(12);
  else if (d <= 16) SM_LAUNCH(16);
  else if (d <= 24) SM_LAUNCH(24);
  else if (d <= 32) SM_LAUNCH(32);
  else if (d <= 48) SM_LAUNCH(48);
  else SM_LAUNCH(72);
#undef SM_LAUNCH
  B200_CUDA(cudaGetLastError());
  m->st.launches += 1;
  return 0;
}

}  // namespace b200
