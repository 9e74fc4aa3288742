// knn_tc.cu -- tensor-core kNN: fused distance + candidate selection, fp64 re-rank.
//
// Replaces BallTree.query over the whole batch (nearest_neighbor/__init__.py:75-76)
// for large problems.  Three stages:
//
//  1. knn_tc_kernel (tcgen05 / TMEM / bulk-TMA, persistent, warp specialised):
//     dot(q, j) = -2 x_q . y_j as ONE contraction of mean-centred, power-of-two scaled fp16
//     operands with fp32 accumulation in TMEM, K = pad16(d): no padding of the contraction for
//     the squared norms.  The fitted points are streamed in order of increasing centred norm,
//     so the norms inside one 32-column chunk span a tiny interval and the filter works on
//         s'(q, j) = dot(q, j) + nlo(chunk of j),   nlo = lower bound of ||y_j||^2 over the chunk,
//     a lower bound of the true score ||y_j||^2 - 2 x_q . y_j that is tight to ~1e-4 of the
//     gaps between neighbours.  Three kinds of warps share the work:
//       * epilogue warps (one thread = one TMEM lane = one query): read each 128-column
//         accumulator tile, reduce every 32-column chunk to its minimum and compare it with the
//         row's threshold.  A row whose chunk may hold a survivor pushes the chunk's 32 raw
//         values into a shared-memory queue -- a handful of stores -- and moves on, so the
//         accumulator goes back to the MMA issuers after a fixed, short time;
//       * selection warps (one per epilogue warp) drain the queues: the 32 values of an entry
//         sit one per lane, a ballot gives the survivors, and the row's candidate list -- kept
//         sorted in shared memory, `keep` entries, threshold = its last entry -- takes them by
//         warp-wide insertion (or a rank-merge when many arrive at once);
//       * one bulk-copy (TMA engine) producer warp and two single-thread MMA issuers.
//     The distance matrix is never written to HBM.
//  2. rerank_kernel: exact fp64 squared distances of the <= keep candidates of each query with
//     the reference's arithmetic (direct differences, dimension order, no FMA), sorted by
//     (d2, index) -> bit-exact neighbours.  Each query is CERTIFIED: every non-candidate has
//     s' >= tau, so with a rigorous bound E on |approximate - exact| score, exact
//     d2_k < (tau - E + ||x_q||^2)/s^2 proves no non-candidate can enter the top k.
//  3. Queries that fail the certificate are resolved exactly (knn_exact.cu: a rescan listing
//     every point within the candidates' k-th distance when they are few, the split
//     brute-force scan otherwise) -- so the result is always the exact answer.
//
// Operand layout in HBM: per 128-row tile a ready-to-use UMMA image [K/8][128][8 halfs]
// (K-major, no swizzle: 8x16-byte core matrices contiguous), so one cp.async.bulk (TMA engine)
// brings a whole tile and the smem descriptor is fixed: LBO = 2048 B (next K chunk),
// SBO = 128 B (next 8 rows).
#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"
#include "ptx.cuh"

namespace b200 {

namespace tc {

constexpr int TILE = 128;          // query rows per accumulator half = data columns per tile
constexpr int CHUNK = 32;          // columns per filter chunk = values per queue entry
constexpr int QWORDS = 36;         // 32-bit words per queue entry: 32 values + {row, column base, nlo bits, -}
constexpr double U16 = 4.8828125e-4;               // 2^-11: fp16 unit roundoff
constexpr double TINY16 = 2.98023223876953125e-8;  // 2^-25: half the smallest fp16 subnormal
constexpr double GAMMA_ACC = 1.52587890625e-5;     // 2^-16: bound on tensor-core fp32 accumulation error, relative to sum |terms| (tests/test_gpu_tensor.py measures it)
constexpr float NLO_PAD = 3.0e38f;                 // norm bound of a chunk that holds only padding columns

__host__ __device__ inline int padK(int d) { return ((d + 15) / 16) * 16; }

// ------------------------------------------------------------------ statistics
__global__ void colsum_kernel(const double* __restrict__ x, int64_t n, int d, double* __restrict__ sum) {
  extern __shared__ double sh[];  // [d]
  for (int j = threadIdx.x; j < d; j += blockDim.x) sh[j] = 0.0;
  __syncthreads();
  const int64_t tot = n * d;
  const int64_t per = (tot + gridDim.x - 1) / gridDim.x;
  const int64_t lo = (int64_t)blockIdx.x * per, hi = imin64(tot, lo + per);
  for (int64_t e = lo + threadIdx.x; e < hi; e += blockDim.x) atomicAdd(&sh[e % d], x[e]);
  __syncthreads();
  for (int j = threadIdx.x; j < d; j += blockDim.x) atomicAdd(&sum[j], sh[j]);
}

__global__ void maxabs_kernel(const double* __restrict__ x, int64_t n, int d, const double* __restrict__ mu,
                              unsigned long long* __restrict__ out /* [0] centred, [1] raw */) {
  const int64_t tot = n * d;
  double mc = 0.0, mr = 0.0;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < tot; e += (int64_t)gridDim.x * blockDim.x) {
    const double v = x[e];
    mc = fmax(mc, fabs(v - mu[e % d]));
    mr = fmax(mr, fabs(v));
  }
  for (int o = 16; o > 0; o >>= 1) {
    mc = fmax(mc, __shfl_xor_sync(0xffffffffu, mc, o));
    mr = fmax(mr, __shfl_xor_sync(0xffffffffu, mr, o));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMax(out, (unsigned long long)__double_as_longlong(mc));
    atomicMax(out + 1, (unsigned long long)__double_as_longlong(mr));
  }
}

// ------------------------------------------------------------- data order
// The fitted points are streamed past every query in order of increasing centred norm.  The score
// ||y||^2 - 2 x.y of a low-norm point is systematically smaller, so the running thresholds tighten
// early and far fewer later points survive the filter; and the norms inside one 32-column chunk
// are nearly equal, which is what lets the filter use one norm bound per chunk.  Any order gives
// the same exact result.  Key: fp32 bits of the squared norm (non-negative, so unsigned order =
// numeric order).
// rw != NULL (the ShepMod radius filter): the order is by c = s^2 rw^2 - ||y||^2 instead, the quantity
// whose spread inside a chunk decides how tight the per-chunk bound is there; c may be negative, so the
// float bits are mapped to an order-preserving unsigned key.
__global__ void norm_key_kernel(const double* __restrict__ x, int64_t n, int d, const double* __restrict__ mu,
                                double scale, const double* __restrict__ rw, uint32_t* __restrict__ key,
                                int32_t* __restrict__ val) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double n2 = 0.0;
  for (int j = 0; j < d; ++j) {
    const double yc = (x[i * d + j] - mu[j]) * scale;
    n2 = fma(yc, yc, n2);
  }
  if (rw) {
    const double r = rw[i] * scale;
    const uint32_t b = __float_as_uint((float)(r * r - n2));
    key[i] = (b & 0x80000000u) ? ~b : (b | 0x80000000u);
  } else {
    key[i] = __float_as_uint((float)n2);
  }
  val[i] = (int32_t)i;
}

// fitted rows in image order (fp64), for the re-rank's gathers
__global__ void gather_rows_kernel(const double* __restrict__ x, int64_t n, int d, const int32_t* __restrict__ perm,
                                   double* __restrict__ xs) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n * d) return;
  const int64_t i = e / d;
  xs[e] = x[(int64_t)perm[i] * d + (e - i * d)];
}

__global__ void iota_kernel(int32_t* __restrict__ v, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) v[i] = (int32_t)i;
}

// ------------------------------------------------------------- operand images
// Data side: one thread per (padded) image position, one warp per 32-column chunk.  Writes the
// position's row of the tile image, the chunk's norm bound nlo (a float that is <= the exact
// squared norm of every real point of the chunk), and folds max(||yc||, ||fp16(yc)||) into ymax.
// Padding positions hold zero rows; the selection warps drop them by their position.
__global__ void __launch_bounds__(128)
data_img_kernel(const double* __restrict__ x, int64_t n, int d, int K, const double* __restrict__ mu, double scale,
                const int32_t* __restrict__ perm, const double* __restrict__ rw, __half* __restrict__ img,
                float* __restrict__ nlo, int64_t npad, unsigned long long* __restrict__ ymax_bits) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;  // image position (npad is a multiple of 128)
  double nm = 0.0;
  float lo = NLO_PAD;
  {
    const int64_t src = i < n ? (int64_t)perm[i] : 0;  // fitted row that sits at this position
    const int64_t t = i / TILE;
    const int r = (int)(i - t * TILE);
    uint4* dst = reinterpret_cast<uint4*>(img + (size_t)t * TILE * K);  // chunk c at dst[c*TILE + r]
    const int KC = K / 8;
    double n2 = 0.0, n2h = 0.0;
    for (int c = 0; c < KC; ++c) {
      __half h[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int j = c * 8 + e;
        __half hh = __float2half_rn(0.0f);
        if (i < n && j < d) {
          const double yc = (x[src * d + j] - mu[j]) * scale;
          hh = __double2half(yc);
          n2 = fma(yc, yc, n2);
          const double yh = (double)__half2float(hh);
          n2h = fma(yh, yh, n2h);
        }
        h[e] = hh;
      }
      dst[(size_t)c * TILE + r] = *reinterpret_cast<uint4*>(h);
    }
    nm = sqrt(fmax(n2, n2h));
    // rounded towards zero with room for the fp64 rounding of n2 itself
    if (i < n) lo = __double2float_rd(n2 * (1.0 - 1e-12));
    if (i < n && rw) {  // radius filter: the chunk bound is -max c, c = s^2 rw^2 - ||y||^2, rounded away from the data
      const double r = rw[src] * scale;
      const double r2 = r * r;
      lo = __double2float_rd(-(r2 - n2) - 1e-12 * (r2 + n2));
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    nm = fmax(nm, __shfl_xor_sync(0xffffffffu, nm, o));
    lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMax(ymax_bits, (unsigned long long)__double_as_longlong(nm));
    nlo[i / CHUNK] = lo;
  }
}

struct QMeta {
  double qn2;   // ||(z - mu) * scale||^2
  double err;   // E_q: bound on |tensor dot - exact dot| (scaled units), +inf when fp16 overflows
  double cmax;  // max |z_j|
};

// Query side: x~ = fp16(-2 * (z - mu) * scale).
__global__ void query_img_kernel(const double* __restrict__ z, int64_t q, int d, int K, const double* __restrict__ mu,
                                 double scale, double ymax, __half* __restrict__ img, int64_t qpad,
                                 QMeta* __restrict__ meta) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= qpad) return;
  const int64_t t = i / TILE;
  const int r = (int)(i - t * TILE);
  uint4* dst = reinterpret_cast<uint4*>(img + (size_t)t * TILE * K);
  const int KC = K / 8;
  double qn2 = 0.0, xn2 = 0.0, x1 = 0.0, cmax = 0.0;
  bool bad = false;
  for (int c = 0; c < KC; ++c) {
    __half h[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int j = c * 8 + e;
      __half hh = __float2half_rn(0.0f);
      if (i < q && j < d) {
        const double zj = z[i * d + j];
        const double zc = (zj - mu[j]) * scale;
        const double xv = -2.0 * zc;
        qn2 = fma(zc, zc, qn2);
        xn2 = fma(xv, xv, xn2);
        x1 += fabs(xv);
        cmax = fmax(cmax, fabs(zj));
        if (!(fabs(xv) < 60000.0)) bad = true;  // also catches NaN
        hh = __double2half(xv);
      }
      h[e] = hh;
    }
    dst[(size_t)c * TILE + r] = *reinterpret_cast<uint4*>(h);
  }
  if (i < q) {
    const double xn = sqrt(xn2) * 1.001;
    // rounding of both operands (Cauchy-Schwarz), the fp16 subnormal floor, and the tensor-core
    // accumulation -- DESIGN.md "certificate".  The rounding of dot + nlo is added by the re-rank.
    double E = 2.0 * U16 * xn * ymax + TINY16 * (x1 + sqrt((double)d) * ymax) + GAMMA_ACC * xn * ymax;
    if (bad) E = CUDART_INF;
    meta[i].qn2 = qn2;
    meta[i].err = E;
    meta[i].cmax = cmax;
  }
}

// ------------------------------------------------------------------ main kernel
struct KParams {
  const __half* a_img;   // query tile images  [qblocks*HALVES][K/8][128][8]
  const __half* b_img;   // data tile images   [ntiles][K/8][128][8]
  const float* nlo;      // [ntiles*4] per-chunk lower bound of the squared norms
  int K;                 // contraction length, multiple of 16
  int ntiles;            // 128-column data tiles of this launch
  int ncols;             // real image positions of this launch (the rest of the last tile is padding)
  int qblocks;           // query blocks of HALVES*128 rows
  int slots;             // depth of the data-tile ring (power of two)
  int qent;              // entries per selection queue (power of two, >= 16)
  int parts;             // bulk copies per data tile (1, 2 or 4)
  int keep;              // candidates kept per query (<= 32*U)
  int mode;              // 0 normal; ceilings: 1 = epilogue reads and reduces but queues nothing, 2 = epilogue only releases
  int32_t* cand_idx;     // [qpad][32*U]
  int32_t* cand_cnt;     // [qpad]
  float* cand_tau;       // [qpad]
  float* cand_sc;        // glists: [qpad][32*U] scores of the lists kept in global memory (beside cand_idx)
  int glists;            // append-and-trim lists live in cand_idx / cand_sc instead of shared memory
  const float* row_thr;  // THR kernels: [qpad] fixed threshold of every query row
  float* dbg;            // optional [qpad][ntiles*128] raw tensor dots (tests only)
  unsigned long long* stats;  // optional [16] counters of the instrumented build (tuning only): events, events with a
                              // survivor, survivors, rank-merges, selection busy clk, queue-full clk, ready-wait clk,
                              // accumulator-wait clk (clk summed over the warps of all CTAs)
};

// Spin limits: a protocol bug must end in a trap, not in a hung GPU (2^27 probes: seconds).
#define B200_SPIN_GUARD(spins)            \
  do {                                    \
    if (++(spins) > (1u << 26)) __trap(); \
  } while (0)
// Tuning switches kept for A/B builds (profiles/microbench/ab.py).  While every row of a block is still filling
// its list the selection warps are what the pipeline waits for, and they share their SM sub-partitions with
// waiting warps; every probe a waiting warp issues is an issue slot they lose:
//   B200_PAIR         two queue entries of different rows per selection step, one per half-warp
//   B200_PRESLEEP     ns of plain sleep before an epilogue warp suspends on its accumulator barrier (every mbarrier
//                     arrival in the CTA wakes the suspended warps to probe again)
//   B200_QWAIT        ns between an epilogue warp's looks at the tail of its full queue (100 -> 1000: -2 %)
//   B200_ISSUER_HINT  the MMA issuers wait for a free accumulator with a suspend hint instead of probing (-1 %)
//   B200_PWAIT        ns between the producer's looks at a full tile ring (64 -> 300: neutral at d = 64, -0.6 % at d = 16)
//   B200_ALL_HINT     the same for the producer's and the issuers' tile-ring waits (slower at d = 64, not used)
#ifndef B200_PAIR
#define B200_PAIR 1
#endif
#ifndef B200_PWAIT
#define B200_PWAIT 300
#endif
#ifndef B200_ALL_HINT
#define B200_ALL_HINT false
#endif
#ifndef B200_ISSUER_HINT
#define B200_ISSUER_HINT true
#endif
#ifndef B200_QWAIT
#define B200_QWAIT 1000
#endif
#ifndef B200_PRESLEEP
#define B200_PRESLEEP 64
#endif
// Wait with a suspend-time hint: the warp sleeps in hardware until the phase completes (or the hint
// elapses) instead of re-issuing the probe; sixteen epilogue warps probing in a tight loop took a fifth
// of all issue slots.
__device__ __forceinline__ void mbar_wait_suspend(uint32_t bar, uint32_t parity) {
  if (B200_PRESLEEP > 0) {
    if (ptx::mbar_test_wait(bar, parity)) return;
    __nanosleep(B200_PRESLEEP);
  }
  uint32_t spins = 0;
  while (!ptx::mbar_try_wait_hint(bar, parity, 4000)) B200_SPIN_GUARD(spins);
}

__device__ __forceinline__ float min3(float a, float b, float c) { return fminf(fminf(a, b), c); }
__device__ __forceinline__ float chunk_min(const float* v) {
  float g[4];
#pragma unroll
  for (int gi = 0; gi < 4; ++gi) {
    const float* w = v + gi * 8;
    g[gi] = fminf(min3(w[0], w[1], w[2]), min3(min3(w[3], w[4], w[5]), w[6], w[7]));
  }
  return fminf(fminf(g[0], g[1]), fminf(g[2], g[3]));
}

// The sorted candidate list of one query row, U entries per lane: entry e = u*32 + lane, ascending
// score; entries never filled hold +inf / -1.  `tau` = entry keep-1.
template <int U>
struct RowList {
  float sc[U];
  int id[U];
  __device__ __forceinline__ void load(const float* ls, const int* li, int lane) {
#pragma unroll
    for (int u = 0; u < U; ++u) {
      sc[u] = ls[u * 32 + lane];
      id[u] = li[u * 32 + lane];
    }
  }
  __device__ __forceinline__ void store(float* ls, int* li, int lane) const {
#pragma unroll
    for (int u = 0; u < U; ++u) {
      ls[u * 32 + lane] = sc[u];
      li[u * 32 + lane] = id[u];
    }
  }
  __device__ __forceinline__ float entry(int e) const {  // e warp-uniform
    float v = sc[0];
#pragma unroll
    for (int u = 1; u < U; ++u) v = (e >> 5) == u ? sc[u] : v;
    return __shfl_sync(0xffffffffu, v, e & 31);
  }
  // insert (sv, col) behind the entries <= sv; the last entry drops out
  __device__ __forceinline__ void insert(float sv, int col, int lane) {
    int pos = 0;
#pragma unroll
    for (int u = 0; u < U; ++u) pos += __popc(__ballot_sync(0xffffffffu, sc[u] <= sv));
#pragma unroll
    for (int u = U - 1; u >= 0; --u) {
      float ps = __shfl_up_sync(0xffffffffu, sc[u], 1);
      int pi = __shfl_up_sync(0xffffffffu, id[u], 1);
      if (u > 0) {
        const float cs = __shfl_sync(0xffffffffu, sc[u - 1], 31);
        const int ci = __shfl_sync(0xffffffffu, id[u - 1], 31);
        if (lane == 0) {
          ps = cs;
          pi = ci;
        }
      }
      const int e = u * 32 + lane;
      if (e > pos) {
        sc[u] = ps;
        id[u] = pi;
      } else if (e == pos) {
        sc[u] = sv;
        id[u] = col;
      }
    }
  }
};

// One queue entry -> the row's list.  All lanes call; `s` is this lane's value of the entry
// (already dot + nlo), `ok` whether it may enter (below the threshold, a real column), m = ballot(ok) != 0.
template <int U>
__device__ __forceinline__ float take_entry(float* ls, int* li, float s, bool ok, unsigned m, int base, float tau,
                                            int keep, int lane) {
  RowList<U> L;
  L.load(ls, li, lane);
  if (__popc(m) < 6) {
    // few survivors (the usual case is one): insert them one at a time
    while (m) {
      const int bl = __ffs(m) - 1;
      m &= m - 1;
      const float sv = __shfl_sync(0xffffffffu, s, bl);
      if (!(sv < tau)) continue;  // the threshold moved since the mask was taken
      L.insert(sv, base + bl, lane);
      tau = L.entry(keep - 1);
    }
    L.store(ls, li, lane);
    return tau;
  }
  // many survivors (the first tiles of a block): rank every old and new element in the merged
  // order -- old before new on equal scores, new ones by lane -- and scatter the first 32*U
  const float sn = ok ? s : CUDART_INF_F;
  int rn = 0;   // new elements before mine
  int ro[U];    // new elements before my old entry u
#pragma unroll
  for (int u = 0; u < U; ++u) ro[u] = 0;
#pragma unroll 8
  for (int j = 0; j < 32; ++j) {
    const float bj = __shfl_sync(0xffffffffu, sn, j);
    rn += (bj < sn || (bj == sn && j < lane)) ? 1 : 0;
#pragma unroll
    for (int u = 0; u < U; ++u) ro[u] += bj < L.sc[u] ? 1 : 0;
  }
  int cle = 0;  // old entries <= my new value
#pragma unroll
  for (int u = 0; u < U; ++u) {
#pragma unroll 8
    for (int j = 0; j < 32; ++j) {
      const float oj = __shfl_sync(0xffffffffu, L.sc[u], j);
      cle += oj <= sn ? 1 : 0;
    }
  }
  __syncwarp();
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const int r = u * 32 + lane + ro[u];
    if (r < 32 * U) {
      ls[r] = L.sc[u];
      li[r] = L.id[u];
    }
  }
  {
    const int r = rn + cle;
    if (r < 32 * U) {
      ls[r] = sn;
      li[r] = ok ? base + lane : -1;
    }
  }
  __syncwarp();
  const int e = keep - 1;
  return ls[e];
}

// ---- append-and-trim lists (U >= 2, keep <= 32U - 12).  A sorted insert into a 64- or 128-entry list costs
// a warp ~150 clk per survivor; here survivors are appended unsorted and, when the list is full, the row is
// cut to the entries below its keep-th smallest score T (one bisection on the score bits), T becoming the
// row's threshold.  Everything dropped or refused has score >= the threshold of that moment >= the final
// one, which is all the certificate needs; the threshold is just up to 32U - keep appends staler.
__device__ __forceinline__ uint32_t score_key(float v) {  // order-preserving map float -> uint32
  const uint32_t b = __float_as_uint(v);
  return b ^ ((uint32_t)((int32_t)b >> 31) | 0x80000000u);
}
__device__ __forceinline__ float key_score(uint32_t k) {
  return __uint_as_float((k & 0x80000000u) ? (k ^ 0x80000000u) : ~k);
}

// Cut the row's first `have` entries to those below the keep-th smallest (have >= keep); returns that score.
template <int U>
__device__ __forceinline__ float trim_row(float* ls, int* li, int keep, int& have, int lane) {
  uint32_t key[U];
  int id[U];
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const int e = u * 32 + lane;
    key[u] = e < have ? score_key(ls[e]) : 0xffffffffu;
    id[u] = li[e];
  }
  uint32_t T = 0;  // max T with #(key < T) < keep  ==  the keep-th smallest key
  for (int bit = 31; bit >= 0; --bit) {
    const uint32_t cand = T | (1u << bit);
    int c = 0;
#pragma unroll
    for (int u = 0; u < U; ++u) c += __popc(__ballot_sync(0xffffffffu, key[u] < cand));
    if (c < keep) T = cand;
  }
  __syncwarp();
  int pos = 0;
  const unsigned lt = (1u << lane) - 1u;
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const bool kf = key[u] < T;
    const unsigned km = __ballot_sync(0xffffffffu, kf);
    if (kf) {
      const int r = pos + __popc(km & lt);
      ls[r] = key_score(key[u]);
      li[r] = id[u];
    }
    pos += __popc(km);
  }
  __syncwarp();
  have = pos;
  return key_score(T);
}

// One queue entry -> the row's unsorted list; `have` = entries in it (updated).  Returns the threshold.
template <int U>
__device__ __forceinline__ float take_append(float* ls, int* li, float s, unsigned m, int base, float tau, int keep,
                                             int& have, int lane) {
  constexpr int W = 32 * U;
  const unsigned lt = (1u << lane) - 1u;
  for (;;) {
    const int room = W - have;
    const int r = __popc(m & lt);
    const bool mine = (m >> lane) & 1u;
    if (mine && r < room) {
      ls[have + r] = s;
      li[have + r] = base + lane;
    }
    const int add = __popc(m);
    __syncwarp();
    if (add <= room) {
      have += add;
      break;
    }
    have = W;
    tau = trim_row<U>(ls, li, keep, have, lane);
    m = __ballot_sync(0xffffffffu, mine && r >= room && s < tau);
    if (m == 0) break;
  }
  return tau;
}

// Warp roles (E = 4*HALVES):
//   [0, 2E)     epilogue, two groups of E warps: group g reads columns [64g, 64g+64) of every tile, so an
//               accumulator is handed back after two TMEM reads per warp instead of four -- the round
//               trip MMA -> epilogue -> MMA on a stage is what the tensor pipe waits for;
//               warp w = g*E + half*4 + quad, thread = query row (half*128 + quad*32 + lane)
//   [2E, 3E)    selection: warp 2E + c owns rows c*32 .. c*32+31 and drains the queues of epilogue
//               warps c (columns 0-63) and E + c (columns 64-127)
//   3E          producer (bulk copies);  3E+1, 3E+2  MMA issuers (issuer j = accumulator stage j)
// THR = true turns the selection into a fixed-threshold filter (ShepMod's radius test): the row's
// threshold is a constant handed in through p.row_thr, the per-chunk bound p.nlo is -max_j c_j of the
// chunk (c_j = s^2 RW_j^2 - ||y_j||^2), and the selection warp appends every column with
// dot + nlo < threshold to the row's list (ids only, no order); a row with more than 32*U such columns is
// marked overflowed (count 32*U + 1) and left to the dense scan.
template <int U, int HALVES, bool DBG, bool THR = false>
__global__ void __launch_bounds__((HALVES * 12 + 3) * 32, 1) knn_tc_kernel(const KParams p) {
  constexpr int ROWS = HALVES * TILE;
  constexpr int E = HALVES * 4;
  constexpr int W = 32 * U;                  // list entries per row
  constexpr uint32_t ACC_COLS = HALVES * TILE;  // TMEM columns per accumulator stage
  extern __shared__ __align__(1024) uint8_t smem[];
  const int K = p.K;
  const uint32_t tile_bytes = (uint32_t)TILE * K * 2;  // one 128-row tile image
  const int QE = p.qent;
  uint8_t* sA = smem;                                  // HALVES query tiles
  uint8_t* sB = sA + (size_t)HALVES * tile_bytes;      // ring of p.slots data tiles
  float* lsc = reinterpret_cast<float*>(sB + (size_t)p.slots * tile_bytes);  // [ROWS][W] list scores
  // (p.glists: the lists live in global memory -- append-and-trim lists are write-mostly, and 128-entry lists of two
  //  halves do not fit beside the tiles -- and these two planes are empty)
  const size_t list_elems = p.glists ? 0 : (size_t)ROWS * W;
  int* lid = reinterpret_cast<int*>(lsc + list_elems);                       // [ROWS][W] list positions
  float* que = reinterpret_cast<float*>(lid + list_elems);                   // [2E][QE][QWORDS]
  float* tau_sh = que + (size_t)2 * E * QE * QWORDS;                         // [ROWS] thresholds (written by selection warps)
  // queue control words, indexed [row block][column group]: heads, tails, done marks; then ready[row block]
  uint32_t* ctl = reinterpret_cast<uint32_t*>((reinterpret_cast<uintptr_t>(tau_sh + ROWS) + 7) & ~(uintptr_t)7);
  uint32_t* q_head = ctl;
  uint32_t* q_tail = ctl + 2 * E;
  uint32_t* q_done = ctl + 4 * E;
  uint32_t* ready = ctl + 6 * E;
  uint64_t* bars = reinterpret_cast<uint64_t*>((reinterpret_cast<uintptr_t>(ready + E) + 15) & ~(uintptr_t)15);
  // bars: full[8], empty[8], accf[stage][half], acce[stage][half], afull, aempty
  uint64_t* bar_full = bars;
  uint64_t* bar_empty = bars + 8;
  uint64_t* bar_accf = bars + 16;
  uint64_t* bar_acce = bars + 20;
  uint64_t* bar_afull = bars + 24;
  uint64_t* bar_aempty = bars + 25;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 26);

  // shfl-broadcast so the compiler knows the role index is warp-uniform: the role bodies then run on
  // the uniform datapath (UR operands for UTCHMMA / UBLKCP without R2UR round trips)
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.slots; ++s) {
      ptx::mbar_init(ptx::smem_u32(&bar_full[s]), 1);
      ptx::mbar_init(ptx::smem_u32(&bar_empty[s]), 1);
    }
    for (int a = 0; a < 4; ++a) {
      ptx::mbar_init(ptx::smem_u32(&bar_accf[a]), 1);
      ptx::mbar_init(ptx::smem_u32(&bar_acce[a]), 8);  // the eight epilogue warps (4 quadrants x 2 column groups) of one half
    }
    ptx::mbar_init(ptx::smem_u32(bar_afull), 1);
    ptx::mbar_init(ptx::smem_u32(bar_aempty), 2);  // one commit per issuer warp
    ptx::fence_barrier_init();
  }
  if (threadIdx.x < 7 * E) ctl[threadIdx.x] = 0;
  if (warp == 3 * E + 1) {  // first MMA warp owns the TMEM allocation
    ptx::tmem_alloc(ptx::smem_u32(tmem_slot), 512);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (tmem_base != 0) {  // one CTA per SM allocating all 512 columns always gets column 0
    if (threadIdx.x == 0) printf("b200: unexpected TMEM base %u\n", tmem_base);
    __trap();
  }

  const int nqb = (p.qblocks - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;  // my query blocks

  if (warp == 3 * E) {
    // ===================== producer: bulk copies on the TMA engine =====================
    if (lane == 0) {
      uint32_t s = 0, ph = 0;
      const uint32_t full0 = ptx::smem_u32(&bar_full[0]), empty0 = ptx::smem_u32(&bar_empty[0]);
      const uint32_t sB0 = ptx::smem_u32(sB);
      for (int b = 0; b < nqb; ++b) {
        const int qb = (int)blockIdx.x + b * (int)gridDim.x;
        ptx::mbar_wait(B200_ALL_HINT, ptx::smem_u32(bar_aempty), (b & 1) ^ 1, B200_ALL_HINT ? 0 : B200_PWAIT);
        ptx::mbar_arrive_expect_tx(ptx::smem_u32(bar_afull), HALVES * tile_bytes);
        ptx::bulk_g2s(ptx::smem_u32(sA), reinterpret_cast<const uint8_t*>(p.a_img) + (size_t)qb * HALVES * tile_bytes,
                      HALVES * tile_bytes, ptx::smem_u32(bar_afull));
        const uint8_t* src = reinterpret_cast<const uint8_t*>(p.b_img);
        for (int t = 0; t < p.ntiles; ++t, src += tile_bytes) {
          ptx::mbar_wait(B200_ALL_HINT, empty0 + s * 8, ph ^ 1, B200_ALL_HINT ? 0 : B200_PWAIT);
          ptx::mbar_arrive_expect_tx(full0 + s * 8, tile_bytes);
          // one tile as `parts` bulk copies: more requests in flight per slot (tile_bytes is a multiple of 4 KB)
          const uint32_t part = tile_bytes / (uint32_t)p.parts;
          for (int c = 0; c < p.parts; ++c)
            ptx::bulk_g2s(sB0 + s * tile_bytes + c * part, src + c * part, part, full0 + s * 8);
          if (++s == (uint32_t)p.slots) {
            s = 0;
            ph ^= 1;
          }
        }
      }
    }
  } else if (warp > 3 * E) {
    // ===================== MMA issuers: two warps, alternate tiles =====================
    // Issuer j owns accumulator stage j.  One thread per issuer warp runs a lean loop: barrier probes,
    // fence, K/16 MMAs per half (M128 N128 K16, 64 clk each), commits.  Descriptors are built once and
    // K steps only bump the address field.
    const int j = warp - (3 * E + 1);
    const uint32_t idesc = ptx::umma_idesc_f16(TILE, TILE);
    const uint32_t lbo = TILE * 16, sbo = 128, kstep = (TILE * 32) >> 4;
    const int ksteps = K / 16;
    uint64_t a_desc[HALVES];
#pragma unroll
    for (int h = 0; h < HALVES; ++h) a_desc[h] = ptx::umma_desc(ptx::smem_u32(sA + (size_t)h * tile_bytes), lbo, sbo);
    const uint64_t b_desc0 = ptx::umma_desc(ptx::smem_u32(sB), lbo, sbo);
    const uint32_t slot_step = tile_bytes >> 4;
    const uint32_t full0 = ptx::smem_u32(&bar_full[0]), empty0 = ptx::smem_u32(&bar_empty[0]);
    const uint32_t accf = ptx::smem_u32(&bar_accf[j * 2]), acce = ptx::smem_u32(&bar_acce[j * 2]);
    const uint32_t slot_mask = p.slots - 1, slot_shift = 31 - __clz(p.slots);
    if (lane == 0) {
      uint32_t it0 = 0;   // global index of the current block's first tile
      uint32_t mine = 0;  // tiles this issuer has done (phase of its accumulator stage)
      const uint32_t afull = ptx::smem_u32(bar_afull), aempty = ptx::smem_u32(bar_aempty);
      const uint32_t ntiles = (uint32_t)p.ntiles;
      for (int b = 0; b < nqb; ++b, it0 += ntiles) {
        ptx::mbar_wait(B200_ALL_HINT, afull, b & 1);
        for (uint32_t t = (j + it0) & 1; t < ntiles; t += 2, ++mine) {
          const uint32_t it = it0 + t, s = it & slot_mask, ph = (it >> slot_shift) & 1;
          const uint64_t bd = b_desc0 + (uint64_t)(s * slot_step);
#pragma unroll
          for (int h = 0; h < HALVES; ++h) {
            // the accumulator half is free once its four epilogue warps have read the previous tile
            ptx::mbar_wait(B200_ISSUER_HINT, acce + h * 8, (mine & 1) ^ 1);
            if (h == 0) ptx::mbar_wait(B200_ALL_HINT, full0 + s * 8, ph);
            ptx::tc_fence_after();
            // (the accumulator base is the literal 0: this CTA owns all 512 columns, checked at start)
            const uint32_t d_tmem = j * ACC_COLS + h * TILE;
#define B200_ISSUE(KS)                                                                                  \
  _Pragma("unroll") for (int ks = 0; ks < KS; ++ks)                                                      \
    ptx::umma_f16(d_tmem, a_desc[h] + (uint64_t)(ks * kstep), bd + (uint64_t)(ks * kstep), idesc, ks > 0 ? 1u : 0u);
            switch (ksteps) {
              case 1: B200_ISSUE(1) break;
              case 2: B200_ISSUE(2) break;
              case 3: B200_ISSUE(3) break;
              case 4: B200_ISSUE(4) break;
              default:
                for (int ks = 0; ks < ksteps; ++ks)
                  ptx::umma_f16(d_tmem, a_desc[h] + (uint64_t)(ks * kstep), bd + (uint64_t)(ks * kstep), idesc,
                                ks > 0 ? 1u : 0u);
            }
#undef B200_ISSUE
            ptx::umma_commit(accf + h * 8);                           // this half's 128 columns are ready
            if (h == HALVES - 1) ptx::umma_commit(empty0 + s * 8);    // tile slot free once these MMAs retire
          }
        }
        ptx::umma_commit(aempty);  // my MMAs on these A tiles are tracked
      }
    }
  } else if (warp < 2 * E) {
    // ===================== epilogue: threshold filter, survivors' chunks to the queue ==============
    // The kernel is bound by instruction issue (~1000 warp instructions per tile and SM sub-partition
    // against 512 clk of tensor time), so this loop is written for instruction count.
    const int g = warp / E;            // column group: this warp reads columns [64g, 64g+64) of every tile
    const int rb = warp - g * E;       // row block: half*4 + quad
    const int half = rb >> 2, quad = rb & 3;
    const int row = rb * 32 + lane;    // my query row inside the block
    const uint32_t t_acc0 = ((uint32_t)(quad * 32) << 16) + half * TILE + g * 64;
    float* myq = que + (size_t)warp * QE * QWORDS;
    const uint32_t c_head = ptx::smem_u32(&q_head[rb * 2 + g]), c_tail = ptx::smem_u32(&q_tail[rb * 2 + g]);
    const uint32_t c_done = ptx::smem_u32(&q_done[rb * 2 + g]), c_ready = ptx::smem_u32(&ready[rb]);
    const uint32_t my_tau = ptx::smem_u32(&tau_sh[row]);
    const uint32_t bar_f0 = ptx::smem_u32(&bar_accf[half]), bar_e0 = ptx::smem_u32(&bar_acce[half]);
    const uint32_t lt_mask = (1u << lane) - 1u;
    const int colg = g * 64;
    // keep these loop constants in registers: left alone, the compiler re-derives the shared-memory
    // addresses from the kernel parameters in every tile step (a dozen instructions each)
#define B200_PIN32(x) asm volatile("" : "+r"(x))
    uint32_t a_tacc = t_acc0, a_head = c_head, a_tail = c_tail, a_tau = my_tau, a_bf = bar_f0, a_be = bar_e0;
    B200_PIN32(a_tacc);
    B200_PIN32(a_head);
    B200_PIN32(a_tail);
    B200_PIN32(a_tau);
    B200_PIN32(a_bf);
    B200_PIN32(a_be);
    asm volatile("" : "+l"(myq));
#undef B200_PIN32
    uint32_t head = 0, tail_c = 0;  // entries pushed so far / last tail seen
    uint32_t it = 0;                // global tile index: stage it & 1, phase (it >> 1) & 1
    unsigned long long st_npush = 0;  // instrumented build only
    for (int b = 0; b < nqb; ++b) {
      const int qb = (int)blockIdx.x + b * (int)gridDim.x;
      {  // the selection warp of my rows has flushed the previous block and reset lists and thresholds
        uint32_t spins = 0;
        while (ptx::ld_acquire_shared(c_ready) != (uint32_t)(b + 1)) {
          __nanosleep(64);
          B200_SPIN_GUARD(spins);
        }
      }
      const float2* nlp = reinterpret_cast<const float2*>(p.nlo) + g;  // my two chunks' norm bounds of tile 0
      float v[32];
      bool pending = false;  // entries stored but not yet published (warp-uniform)
      // Queue push in two steps.  STORE: the flagged lanes of `fm_` (at most QE of them) take
      // consecutive slots and write their 32 values; PUBLISH: the new head becomes visible to the
      // selection warp.  The first chunk is only stored before the accumulator goes back to the MMA
      // issuers; it is published after the release.
#define B200_PUBLISH()                                                                               \
  do {                                                                                               \
    if (pending) {                                                                                   \
      __syncwarp();                                                                                  \
      if (lane == 0) ptx::st_release_shared(a_head, head);                                           \
      pending = false;                                                                               \
    }                                                                                                \
  } while (0)
#define B200_STORE(fm_, col_, nlc)                                                                   \
  do {                                                                                               \
    const uint32_t nadd_ = __popc(fm_);                                                              \
    if (head + nadd_ - tail_c > (uint32_t)QE) {                                                      \
      B200_PUBLISH(); /* the selection warp must see what is stored before this warp waits for it */ \
      uint32_t spins_ = 0;                                                                           \
      do {                                                                                           \
        tail_c = ptx::ld_acquire_shared(a_tail);                                                     \
        if (head + nadd_ - tail_c <= (uint32_t)QE) break;                                            \
        __nanosleep(B200_QWAIT);                                                                     \
        B200_SPIN_GUARD(spins_);                                                                     \
      } while (true);                                                                                \
    }                                                                                                \
    if ((fm_ >> lane) & 1u) {                                                                        \
      const uint32_t slot_ = (head + __popc(fm_ & lt_mask)) & (uint32_t)(QE - 1);                    \
      float4* dst_ = reinterpret_cast<float4*>(myq + (size_t)slot_ * QWORDS);                        \
      _Pragma("unroll") for (int i = 0; i < 8; ++i)                                                  \
        dst_[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);                   \
      reinterpret_cast<int4*>(dst_)[8] = make_int4(lane, col_, __float_as_int(nlc), 0);              \
    }                                                                                                \
    head += nadd_;                                                                                   \
    pending = true;                                                                                  \
    if (DBG) ++st_npush;                                                                             \
  } while (0)
      // one 32-column chunk in v: dump (tests), reduce to its minimum, compare with the row's threshold
      // on the raw dot -- dot + nlo < tau, widened by the rounding of both sides (an inexact test here
      // only adds queue entries; the selection warp decides) -- and store the flagged rows' values
#define B200_CHUNK(col_, nlc)                                                                        \
  do {                                                                                               \
    if (DBG) {                                                                                       \
      if (p.dbg) {                                                                                   \
        float* o = p.dbg + ((size_t)qb * ROWS + row) * ((size_t)p.ntiles * TILE) + (col_);           \
        _Pragma("unroll") for (int i = 0; i < 32; ++i) o[i] = v[i];                                  \
      }                                                                                              \
    }                                                                                                \
    const bool fl_ = chunk_min(v) < (tau - (nlc)) + fmaf(fabsf(nlc), 4.76837158e-7f, at);            \
    const unsigned fa_ = (!DBG || (p.mode & 3) == 0) ? __ballot_sync(0xffffffffu, fl_) : 0u;         \
    if (fa_) {                                                                                       \
      if (QE >= 32) {                                                                                \
        B200_STORE(fa_, col_, nlc);                                                                  \
      } else {                                                                                       \
        const unsigned lo_ = fa_ & 0xffffu, hi_ = fa_ & 0xffff0000u;                                 \
        if (lo_) B200_STORE(lo_, col_, nlc);                                                         \
        if (hi_) B200_STORE(hi_, col_, nlc);                                                         \
      }                                                                                              \
    }                                                                                                \
  } while (0)
      // one tile on accumulator stage A: wait, two TMEM reads, release, filter
#define B200_TILE(A, PH)                                                                             \
  do {                                                                                               \
    const float2 nl = __ldg(nlp);                                                                    \
    nlp += 2;                                                                                        \
    mbar_wait_suspend(a_bf + (A) * 16, (PH));                                                  \
    ptx::tc_fence_after();                                                                           \
    if (!DBG || (p.mode & 3) != 2) {                                                                 \
      const float tau = ptx::ld_volatile_shared_f32(a_tau);                                         \
      const float at = fabsf(tau) * 4.76837158e-7f;                                                  \
      ptx::tmem_ld32(a_tacc + (A) * ACC_COLS, v);                                                      \
      ptx::tmem_ld_wait();                                                                           \
      B200_CHUNK(col, nl.x);                                                                         \
      ptx::tmem_ld32(a_tacc + (A) * ACC_COLS + 32, v);                                               \
      ptx::tmem_ld_wait();                                                                           \
      /* my columns of the tile have been read: hand the accumulator back before the last chunk */   \
      ptx::tc_fence_before();                                                                        \
      __syncwarp();                                                                                  \
      if (lane == 0) ptx::mbar_arrive(a_be + (A) * 16);                                        \
      B200_CHUNK(col + 32, nl.y);                                                                    \
      B200_PUBLISH();                                                                                \
    } else {                                                                                         \
      ptx::tc_fence_before();                                                                        \
      __syncwarp();                                                                                  \
      if (lane == 0) ptx::mbar_arrive(a_be + (A) * 16);                                        \
    }                                                                                                \
    col += TILE;                                                                                     \
  } while (0)
      int col = colg;  // first of my columns in the current tile
      // (one instance of the tile step with run-time stage selects: instantiating it per stage made the
      // loop four times as long and cost more in instruction fetch than the selects do)
      const long long blk0 = (DBG && p.stats) ? clock64() : 0;
      for (int t = 0; t < p.ntiles; ++t, ++it) {
        const uint32_t a = it & 1;
        B200_TILE(a, (it >> 1) & 1);
        if (DBG && p.stats && warp == 0 && lane == 0) {  // block timeline: clocks to reach tiles 64, 512, 2048, the end
          if (t == 63) atomicAdd(p.stats + 12, (unsigned long long)(clock64() - blk0));
          if (t == 511) atomicAdd(p.stats + 13, (unsigned long long)(clock64() - blk0));
          if (t == 2047) atomicAdd(p.stats + 14, (unsigned long long)(clock64() - blk0));
          if (t == p.ntiles - 1) atomicAdd(p.stats + 15, (unsigned long long)(clock64() - blk0));
        }
      }
#undef B200_TILE
#undef B200_CHUNK
#undef B200_STORE
#undef B200_PUBLISH
      __syncwarp();
      if (lane == 0) ptx::st_release_shared(c_done, (uint32_t)(b + 1));  // everything of this block is queued
    }
    if (DBG && p.stats && lane == 0) atomicAdd(p.stats + 11, st_npush);
  } else {
    // ===================== selection: queue entries -> sorted candidate lists ======================
    const int c = warp - 2 * E;  // row block: rows c*32 .. c*32+31, fed by epilogue warps c and E + c
    float* ls0 = lsc + (size_t)c * 32 * W;
    int* li0 = lid + (size_t)c * 32 * W;
    const bool glists = U >= 2 && !THR && p.glists != 0;
    float* tau0 = tau_sh + c * 32;
    const uint32_t c_ready = ptx::smem_u32(&ready[c]);
    const int keep = p.keep;
    uint32_t tail[2] = {0u, 0u};
    const bool append = !THR && U >= 2 && keep <= W - 12;  // append-and-trim lists instead of sorted ones
    int cntreg = 0;         // append lists: entries in row c*32 + lane
    long long st_busy = 0;  // instrumented build only
    unsigned long long st_ev = 0, st_hit = 0, st_surv = 0, st_bulk = 0, st_trim = 0, st_pair = 0;
    for (int b = 0; b < nqb; ++b) {
      const int qb = (int)blockIdx.x + b * (int)gridDim.x;
      if (glists) {  // the rows' lists are their slices of the output arrays (only entries below the count are read)
        ls0 = p.cand_sc + ((int64_t)qb * ROWS + c * 32) * W;
        li0 = p.cand_idx + ((int64_t)qb * ROWS + c * 32) * W;
      } else {
        for (int e = lane; e < 32 * W; e += 32) {
          ls0[e] = CUDART_INF_F;
          li0[e] = -1;
        }
      }
      int* cnt0 = reinterpret_cast<int*>(ls0);  // THR: entries per row (the score plane is unused there)
      if (THR) {
        cnt0[lane] = 0;
        tau0[lane] = p.row_thr[((int64_t)qb * ROWS + c * 32) + lane];
      } else {
        tau0[lane] = CUDART_INF_F;
      }
      cntreg = 0;
      __syncwarp();
      if (lane == 0) ptx::st_release_shared(c_ready, (uint32_t)(b + 1));
      uint32_t spins = 0, nap = 32;
      uint32_t a_heads = ptx::smem_u32(&q_head[c * 2]), a_tails = ptx::smem_u32(&q_tail[c * 2]);
      uint32_t a_dones = ptx::smem_u32(&q_done[c * 2]);
      asm volatile("" : "+r"(a_heads));
      asm volatile("" : "+r"(a_tails));
      asm volatile("" : "+r"(a_dones));
      for (;;) {
        // idle probe: both heads in one 8-byte load (the acquire load follows only when there is work)
        const uint2 hh = ptx::ld_volatile_shared_v2(a_heads);
        if (hh.x == tail[0] && hh.y == tail[1]) {
          // both queues empty: finished once both epilogue warps have queued everything (done is
          // written after the last head update, so re-reading the heads after it is conclusive)
          const uint2 dd = ptx::ld_volatile_shared_v2(a_dones);
          if (dd.x == (uint32_t)(b + 1) && dd.y == (uint32_t)(b + 1)) {
            if (ptx::ld_acquire_shared(a_heads) == tail[0] && ptx::ld_acquire_shared(a_heads + 4) == tail[1]) break;
            continue;
          }
          __nanosleep(nap);  // back off: an entry then waits at most ~0.3 us, and the queues hold 16
          if (nap < 2048) nap *= 2;
          B200_SPIN_GUARD(spins);
          continue;
        }
        spins = 0;
        nap = 32;
#pragma unroll
        for (int g = 0; g < 2; ++g) {
          const float* myq = que + (size_t)(g * E + c) * QE * QWORDS;
          const uint32_t h = ptx::ld_acquire_shared(a_heads + 4 * g);
          uint32_t tl = tail[g];
          const long long c0 = (DBG && p.stats && tl != h) ? clock64() : 0;
          while (tl != h) {
            if (B200_PAIR && U == 1 && !THR && h - tl >= 2) {
              // Two entries of different rows at once, one per half-warp: a row's 32-entry list is spread over 16
              // lanes (two entries each), shuffles run 16 wide.  One entry is a chain of dependent shuffles and
              // shared-memory round trips (~500 clk for ~60 instructions); while every row of a block is still
              // filling its list the selection warps are what the whole pipeline waits for, and the second
              // half-warp's chain rides in the same instructions.
              const float* ea = myq + (size_t)(tl & (uint32_t)(QE - 1)) * QWORDS;
              const float* eb = myq + (size_t)((tl + 1) & (uint32_t)(QE - 1)) * QWORDS;
              const int4 ma = *reinterpret_cast<const int4*>(ea + 32);
              const int4 mb = *reinterpret_cast<const int4*>(eb + 32);
              if (ma.x != mb.x) {
                const int hw = lane >> 4, sub = lane & 15;
                const float* em = hw ? eb : ea;
                const int4 mt = hw ? mb : ma;
                const float nlc = __int_as_float(mt.z);
                const float v0 = em[sub] + nlc, v1 = em[16 + sub] + nlc;
                float tq = tau0[mt.x];
                const unsigned b0 = __ballot_sync(0xffffffffu, v0 < tq && mt.y + sub < p.ncols);
                const unsigned b1 = __ballot_sync(0xffffffffu, v1 < tq && mt.y + 16 + sub < p.ncols);
                const unsigned mA = (b0 & 0xffffu) | (b1 << 16), mB = (b0 >> 16) | (b1 & 0xffff0000u);
                if (__popc(mA) < 6 && __popc(mB) < 6) {
                  tl += 2;
                  if (lane == 0) ptx::st_volatile_shared_u32(a_tails + 4 * g, tl);
                  if (DBG) {
                    st_pair += 2;
                    st_ev += 2;
                    st_hit += (mA != 0) + (mB != 0);
                    st_surv += __popc(mA) + __popc(mB);
                  }
                  if (mA | mB) {
                    unsigned mm = hw ? mB : mA;
                    const bool mine = mm != 0;
                    float* lsr = ls0 + (size_t)mt.x * W;
                    int* lir = li0 + (size_t)mt.x * W;
                    float sc0 = lsr[sub], sc1 = lsr[16 + sub];
                    int id0 = lir[sub], id1 = lir[16 + sub];
                    const float t_in = tq;
                    const int ke = keep - 1, sh = hw * 16;
                    unsigned any = 1u;
                    while (any) {
                      const bool act = mm != 0;
                      const int bl = act ? __ffs(mm) - 1 : 0;
                      mm &= mm - 1;  // (0 stays 0)
                      float sv = __shfl_sync(0xffffffffu, (bl & 16) ? v1 : v0, bl & 15, 16);
                      sv = (act && sv < tq) ? sv : CUDART_INF_F;  // +inf lands behind the last entry: nothing moves
                      const unsigned c0 = __ballot_sync(0xffffffffu, sc0 <= sv), c1 = __ballot_sync(0xffffffffu, sc1 <= sv);
                      const int pos = __popc((c0 >> sh) & 0xffffu) + __popc((c1 >> sh) & 0xffffu);
                      float p1s = __shfl_up_sync(0xffffffffu, sc1, 1, 16);
                      int p1i = __shfl_up_sync(0xffffffffu, id1, 1, 16);
                      const float cs = __shfl_sync(0xffffffffu, sc0, 15, 16);
                      const int ci = __shfl_sync(0xffffffffu, id0, 15, 16);
                      const float p0s = __shfl_up_sync(0xffffffffu, sc0, 1, 16);
                      const int p0i = __shfl_up_sync(0xffffffffu, id0, 1, 16);
                      if (sub == 0) {
                        p1s = cs;
                        p1i = ci;
                      }
                      const int col = mt.y + bl;
                      if (16 + sub > pos) {
                        sc1 = p1s;
                        id1 = p1i;
                      } else if (16 + sub == pos) {
                        sc1 = sv;
                        id1 = col;
                      }
                      if (sub > pos) {
                        sc0 = p0s;
                        id0 = p0i;
                      } else if (sub == pos) {
                        sc0 = sv;
                        id0 = col;
                      }
                      tq = __shfl_sync(0xffffffffu, (ke & 16) ? sc1 : sc0, ke & 15, 16);
                      any = __ballot_sync(0xffffffffu, mm != 0);
                    }
                    if (mine) {
                      lsr[sub] = sc0;
                      lsr[16 + sub] = sc1;
                      lir[sub] = id0;
                      lir[16 + sub] = id1;
                      if (sub == 0 && tq != t_in) ptx::st_volatile_shared_f32(ptx::smem_u32(&tau0[mt.x]), tq);
                    }
                  }
                  __syncwarp();
                  continue;
                }
              }
            }
            if (B200_PAIR && U >= 2 && !THR && append && h - tl >= 2) {
              // append lists: two entries of different rows per step as well, one per half-warp (two values per
              // lane); a pair that would overfill one of the two lists is left to the single path and its trim
              const float* ea = myq + (size_t)(tl & (uint32_t)(QE - 1)) * QWORDS;
              const float* eb = myq + (size_t)((tl + 1) & (uint32_t)(QE - 1)) * QWORDS;
              const int4 ma = *reinterpret_cast<const int4*>(ea + 32);
              const int4 mb = *reinterpret_cast<const int4*>(eb + 32);
              if (ma.x != mb.x) {
                const int hw = lane >> 4, sub = lane & 15;
                const float* em = hw ? eb : ea;
                const int4 mt = hw ? mb : ma;
                const float nlc = __int_as_float(mt.z);
                const float v0 = em[sub] + nlc, v1 = em[16 + sub] + nlc;
                const float tq = tau0[mt.x];
                const unsigned b0 = __ballot_sync(0xffffffffu, v0 < tq && mt.y + sub < p.ncols);
                const unsigned b1 = __ballot_sync(0xffffffffu, v1 < tq && mt.y + 16 + sub < p.ncols);
                const unsigned mA = (b0 & 0xffffu) | (b1 << 16), mB = (b0 >> 16) | (b1 & 0xffff0000u);
                const int haveA = __shfl_sync(0xffffffffu, cntreg, ma.x), haveB = __shfl_sync(0xffffffffu, cntreg, mb.x);
                if (haveA + __popc(mA) <= W && haveB + __popc(mB) <= W) {
                  tl += 2;
                  if (lane == 0) ptx::st_volatile_shared_u32(a_tails + 4 * g, tl);
                  if (DBG) {
                    st_pair += 2;
                    st_ev += 2;
                    st_hit += (mA != 0) + (mB != 0);
                    st_surv += __popc(mA) + __popc(mB);
                  }
                  const unsigned mm = hw ? mB : mA;
                  const int have = hw ? haveB : haveA;
                  float* lsr = ls0 + (size_t)mt.x * W;
                  int* lir = li0 + (size_t)mt.x * W;
                  if ((mm >> sub) & 1u) {
                    const int r = have + __popc(mm & ((1u << sub) - 1u));
                    lsr[r] = v0;
                    lir[r] = mt.y + sub;
                  }
                  if ((mm >> (16 + sub)) & 1u) {
                    const int r = have + __popc(mm & ((1u << (16 + sub)) - 1u));
                    lsr[r] = v1;
                    lir[r] = mt.y + 16 + sub;
                  }
                  if (lane == ma.x) cntreg = haveA + __popc(mA);
                  if (lane == mb.x) cntreg = haveB + __popc(mB);
                  __syncwarp();
                  continue;
                }
              }
            }
            const float* e = myq + (size_t)(tl & (uint32_t)(QE - 1)) * QWORDS;
            const int4 meta = *reinterpret_cast<const int4*>(e + 32);
            const float s = e[lane] + __int_as_float(meta.z);  // dot + nlo of the chunk
            ++tl;
            const int rl = meta.x, base = meta.y;
            const float tau = tau0[rl];
            const bool ok = s < tau && base + lane < p.ncols;
            const unsigned m = __ballot_sync(0xffffffffu, ok);
            // the entry's words are in registers (the ballot consumed them): the slot may be refilled
            if (lane == 0) ptx::st_volatile_shared_u32(a_tails + 4 * g, tl);
            if (DBG) {
              ++st_ev;
              st_hit += m != 0;
              st_surv += __popc(m);
              st_bulk += __popc(m) >= 6;
            }
            if (m == 0) continue;
            if (THR) {
              const int have = cnt0[rl];
              const int add = __popc(m);
              __syncwarp();
              if (have + add <= W) {
                if (ok) li0[(size_t)rl * W + have + __popc(m & ((1u << lane) - 1u))] = base + lane;
                if (lane == 0) cnt0[rl] = have + add;
              } else if (lane == 0) {
                cnt0[rl] = W + 1;  // overflow: stays (W + 1 + anything > W)
              }
              __syncwarp();
              continue;
            }
            float nt;
            if (U >= 2 && append) {
              int have = __shfl_sync(0xffffffffu, cntreg, rl);
              const int before = have + __popc(m);
              nt = take_append<U>(ls0 + (size_t)rl * W, li0 + (size_t)rl * W, s, m, base, tau, keep, have, lane);
              if (DBG) st_trim += have < before;
              if (lane == rl) cntreg = have;
            } else {
              nt = take_entry<U>(ls0 + (size_t)rl * W, li0 + (size_t)rl * W, s, ok, m, base, tau, keep, lane);
            }
            if (lane == 0 && nt != tau) ptx::st_volatile_shared_f32(ptx::smem_u32(&tau0[rl]), nt);
            __syncwarp();
          }
          if (DBG && c0) st_busy += clock64() - c0;
          tail[g] = tl;
        }
      }
      // ---- flush this block's candidate lists (coalesced, one row at a time)
      const int64_t q0 = (int64_t)qb * ROWS + c * 32;
      int my_cnt = 0;
      if (U >= 2 && append) {
        // rows holding more than keep entries are cut once more: the re-rank then gathers keep - 1 rows
        for (int r = 0; r < 32; ++r) {
          int have = __shfl_sync(0xffffffffu, cntreg, r);
          if (have > keep) {
            const float nt = trim_row<U>(ls0 + (size_t)r * W, li0 + (size_t)r * W, keep, have, lane);
            if (lane == r) {
              cntreg = have;
              tau0[r] = nt;
            }
          }
          for (int e = have + lane; e < W; e += 32) li0[(size_t)r * W + e] = -1;
        }
        __syncwarp();
      }
      if (!glists) {
        for (int r = 0; r < 32; ++r) {
          int nv = 0;
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const int e = u * 32 + lane;
            const int id = li0[(size_t)r * W + e];
            p.cand_idx[(q0 + r) * W + e] = id;
            nv += __popc(__ballot_sync(0xffffffffu, id >= 0 && e < keep));
          }
          if (lane == r) my_cnt = nv;
        }
      }
      if (THR) my_cnt = min(cnt0[lane], W + 1);
      if (U >= 2 && append) my_cnt = cntreg;
      p.cand_cnt[q0 + lane] = my_cnt;
      p.cand_tau[q0 + lane] = tau0[lane];
      __syncwarp();
    }
    if (DBG && p.stats && lane == 0) {
      atomicAdd(p.stats + 0, st_ev);
      atomicAdd(p.stats + 1, st_hit);
      atomicAdd(p.stats + 2, st_surv);
      atomicAdd(p.stats + 3, st_bulk);
      atomicAdd(p.stats + 4, (unsigned long long)st_busy);
      atomicAdd(p.stats + 5, st_trim);
      atomicAdd(p.stats + 6, st_pair);
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 3 * E + 1) ptx::tmem_dealloc(tmem_base, 512);
}

// ------------------------------------------------------------------ fp64 re-rank
// One warp per query: exact squared distances of its candidates (reference
// arithmetic), rank by (d2, index), emit the k best, test the certificate.
// Candidate lists of one query when the data stream was split over several launches (the last, partly
// filled round of query blocks): list r holds positions relative to its sub-range, `off[r]` is that
// sub-range's first position.  nlists = 1: the usual single list.
struct RerankLists {
  int nlists;
  int64_t list_elems;   // elements between lists in cand_idx
  int64_t meta_elems;   // elements between lists in cand_cnt / cand_tau
  int off[4];
};

template <int CAP>
__global__ void __launch_bounds__(256)
rerank_kernel(const double* __restrict__ x, int64_t n, int d, const double* __restrict__ z, int64_t q,
              int k, int64_t self_offset, const int32_t* __restrict__ cand_idx, int cstride,
              const int32_t* __restrict__ cand_cnt, const float* __restrict__ cand_tau,
              const QMeta* __restrict__ meta, double scale, double data_cmax, const int32_t* __restrict__ perm,
              const double* __restrict__ xs, int32_t* __restrict__ out_idx, double* __restrict__ out_d2, int32_t* __restrict__ fail_list,
              int32_t* __restrict__ fail_cnt, const RerankLists L) {
  constexpr int U = CAP / 32;
  __shared__ double s_d2[8][CAP];
  __shared__ int s_id[8][CAP];
  __shared__ double s_dk[8];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t p = (int64_t)blockIdx.x * 8 + w;
  if (p >= q) return;
  const int cnt = min(cand_cnt[p], min(cstride, CAP));
  const int self = self_offset >= 0 ? (int)(self_offset + p) : -1;
  // the query row goes to shared memory once (streaming loads: it is used by this warp only and should
  // not push gathered rows out of L2); the lanes then read it as broadcasts
  extern __shared__ double zsm[];
  double* zp = zsm + (size_t)w * ((d + 1) & ~1);
  for (int j = lane; j < d; j += 32) zp[j] = __ldcs(z + p * d + j);
  __syncwarp();
  double dd[U];
  int id[U];
  int nvalid = 0;
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const int j = lane + 32 * u;
    int pos;
    if (L.nlists > 1) {  // pass u takes list u (each at most 32 entries after the kernel's final trim)
      const int cu = u < L.nlists ? min(cand_cnt[u * L.meta_elems + p], 32) : 0;
      pos = lane < cu ? cand_idx[u * L.list_elems + p * cstride + lane] + L.off[u] : -1;
    } else {
      pos = j < cnt ? cand_idx[p * cstride + j] : -1;
    }
    const int c = (pos >= 0 && pos < n) ? perm[pos] : -1;  // image position -> fitted row (padding positions drop out)
    const bool ok = c >= 0 && c != self;
    // xs: the same rows in image (norm) order.  Neighbours are mostly low-norm points, so two thirds of
    // the gathered rows come from the first quarter of xs and stay in L2; in fitted order they are scattered.
    const double* rowp = xs ? xs + (int64_t)pos * d : x + (int64_t)c * d;
    dd[u] = ok ? sqdist_seq_wide<true>(rowp, zp, d) : CUDART_INF;
    id[u] = ok ? c : 0x7fffffff;
    s_d2[w][j] = dd[u];
    s_id[w][j] = id[u];
    nvalid += __popc(__ballot_sync(0xffffffffu, ok));
  }
  __syncwarp();
  int rank[U];
#pragma unroll
  for (int u = 0; u < U; ++u) rank[u] = 0;
  for (int i = 0; i < CAP; ++i) {
    const double di = s_d2[w][i];
    const int ii = s_id[w][i];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int j = lane + 32 * u;
      const bool less = di < dd[u] || (di == dd[u] && (ii < id[u] || (ii == id[u] && i < j)));
      rank[u] += less ? 1 : 0;
    }
  }
#pragma unroll
  for (int u = 0; u < U; ++u) {
    if (rank[u] < k) {
      out_idx[p * k + rank[u]] = id[u] == 0x7fffffff ? -1 : id[u];
      if (out_d2) out_d2[p * k + rank[u]] = dd[u];
    }
    if (rank[u] == k - 1) s_dk[w] = dd[u];
  }
  __syncwarp();
  if (lane == 0) {
    bool good = nvalid >= k;
    float tau = cand_tau[p];
    for (int r = 1; r < L.nlists; ++r) tau = fminf(tau, cand_tau[r * L.meta_elems + p]);  // bound for every non-candidate
    const QMeta mq = meta[p];
    good = good && mq.err < CUDART_INF;  // a query the fp16 image cannot hold is never certified
    if (good && tau < CUDART_INF_F) {
      // scaled units: every non-candidate j has  dot_j + nlo_j >= tau (1 - 2^-24) (fp32 rounding of the sum),
      // hence  d2_j * s^2 >= tau - 2^-23 |tau| - E + ||x_q||^2
      const double lb = (double)tau - 1.1920928955078125e-7 * fabs((double)tau) - mq.err + mq.qn2;
      const double cm = fmax(data_cmax, mq.cmax) * scale;
      const double slack = 1e-7 * fabs(lb) + 1e-9 * (double)d * cm * cm;  // fp64 rounding of the centred form
      good = s_dk[w] * scale * scale < lb - slack;
    }
    if (!good) fail_list[atomicAdd(fail_cnt, 1)] = (int32_t)p;
  }
}

}  // namespace tc

// ------------------------------------------------------------------ host side
namespace {

// rw == NULL: the kNN images (m->tc).  rw != NULL: the images of the ShepMod radius filter (m->tc_sm),
// ordered and bounded by c = s^2 rw^2 - ||y||^2.
int tc_prepare(b200_model* m, cudaStream_t st, const double* rw = nullptr) {
  TcData& t = rw ? m->tc_sm : m->tc;
  if (t.ready) return 0;
  const int64_t n = m->n;
  const int d = m->d;
  t.K = tc::padK(d);
  t.tiles = (n + tc::TILE - 1) / tc::TILE;
  const int64_t npad = t.tiles * tc::TILE;
  B200_TRY(t.mu.reserve((size_t)d * 8 + 64));
  B200_TRY(t.img.reserve((size_t)npad * t.K * 2));
  B200_TRY(t.nlo.reserve((size_t)t.tiles * 4 * sizeof(float) + 64));
  B200_TRY(m->flags.reserve(256));
  B200_CUDA(cudaMemsetAsync(t.mu.p, 0, (size_t)d * 8, st));
  B200_CUDA(cudaMemsetAsync(m->flags.p, 0, 256, st));
  tc::colsum_kernel<<<(unsigned)std::min<int64_t>(1024, (n * d + 4095) / 4096), 256, (size_t)d * 8, st>>>(
      m->x, n, d, t.mu.as<double>());
  std::vector<double> mu(d);
  B200_CUDA(cudaMemcpyAsync(mu.data(), t.mu.p, (size_t)d * 8, cudaMemcpyDeviceToHost, st));
  B200_CUDA(cudaStreamSynchronize(st));
  for (int j = 0; j < d; ++j) mu[j] /= (double)n;
  B200_CUDA(cudaMemcpyAsync(t.mu.p, mu.data(), (size_t)d * 8, cudaMemcpyHostToDevice, st));
  unsigned long long* bits = reinterpret_cast<unsigned long long*>(m->flags.as<char>() + 192);
  tc::maxabs_kernel<<<1024, 256, 0, st>>>(m->x, n, d, t.mu.as<double>(), bits);
  unsigned long long hb[2];
  B200_CUDA(cudaMemcpyAsync(hb, bits, 16, cudaMemcpyDeviceToHost, st));
  B200_CUDA(cudaStreamSynchronize(st));
  double mc, mr;
  memcpy(&mc, &hb[0], 8);
  memcpy(&mr, &hb[1], 8);
  if (!(mc < 1e300) || !(mr < 1e300)) {
    set_error("tensor path: non-finite coordinates");
    return 1;
  }
  int ex = 0;
  t.scale = 1.0;
  if (mc > 0) {
    frexp(mc, &ex);              // mc = f * 2^ex, f in [0.5, 1)
    t.scale = ldexp(1.0, -ex);   // |centred value| * scale < 1
  }
  t.cmax = mr;
  // image position -> fitted row, by increasing centred norm
  B200_TRY(t.perm.reserve((size_t)n * 4 + 64));
  const unsigned nb = (unsigned)((n + 255) / 256);
  {
    DevBuf keys, vals, tmp;
    B200_TRY(keys.reserve((size_t)n * 8 + 64));
    B200_TRY(vals.reserve((size_t)n * 4 + 64));
    uint32_t* k_in = keys.as<uint32_t>();
    uint32_t* k_out = k_in + n;
    tc::norm_key_kernel<<<nb, 256, 0, st>>>(m->x, n, d, t.mu.as<double>(), t.scale, rw, k_in, vals.as<int32_t>());
    size_t tmp_bytes = 0;
    B200_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, k_in, k_out, vals.as<int32_t>(), t.perm.as<int32_t>(),
                                              (int)n, 0, 32, st));
    B200_TRY(tmp.reserve(tmp_bytes + 64));
    B200_CUDA(cub::DeviceRadixSort::SortPairs(tmp.p, tmp_bytes, k_in, k_out, vals.as<int32_t>(), t.perm.as<int32_t>(),
                                              (int)n, 0, 32, st));
    B200_CUDA(cudaStreamSynchronize(st));
    keys.release();
    vals.release();
    tmp.release();
    m->st.launches += 2;
  }
  B200_CUDA(cudaGetLastError());
  // a copy of the rows in image order for the re-rank (skipped for very large sets: it doubles the data)
  t.xs.release();
  if (!rw && (size_t)n * d * 8 <= ((size_t)16 << 30)) {
    B200_TRY(t.xs.reserve((size_t)n * d * 8));
    const int64_t tot = n * d;
    tc::gather_rows_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(m->x, n, d, t.perm.as<int32_t>(), t.xs.as<double>());
    B200_CUDA(cudaGetLastError());
    m->st.launches += 1;
  }
  B200_CUDA(cudaMemsetAsync(bits, 0, 8, st));
  tc::data_img_kernel<<<(unsigned)(npad / 128), 128, 0, st>>>(m->x, n, d, t.K, t.mu.as<double>(), t.scale, t.perm.as<int32_t>(),
                                                              rw, t.img.as<__half>(), t.nlo.as<float>(), npad, bits);
  B200_CUDA(cudaGetLastError());
  B200_CUDA(cudaMemcpyAsync(hb, bits, 8, cudaMemcpyDeviceToHost, st));
  B200_CUDA(cudaStreamSynchronize(st));
  memcpy(&t.ymax, &hb[0], 8);
  m->st.launches += 3;
  t.ready = true;
  return 0;
}

struct TcConfig {
  int u, halves, slots, keep, qent;
  int glists;  // the rows' lists live in global memory (128-entry append-and-trim lists)
  size_t smem;
};

// Kernel shape for k (+1 when the query's own row is dropped later): the sorted list of a row holds
// 32*u entries (u = 1, 2, 4), of which the first keep = k + margin count; two 128-row halves per CTA
// when shared memory allows, else one; the data-tile ring and the selection queues share what is left.
bool tc_config(int d, int k, int margin, bool self, int force_halves, TcConfig* c, int force_slots = 0,
               bool shared_lists_only = false) {
  const int K = tc::padK(d);
  const int base = k + (self ? 1 : 0);
  int keep = base + (margin < 0 ? std::max(4, std::min(k / 2, 16)) : margin);
  if (base > 128) return false;
  if (keep > 128) keep = 128;
  // keep <= 32: sorted 32-entry lists in shared memory.  keep <= 52: 64-entry append-and-trim lists, in shared
  // memory when that leaves room for two halves per CTA and a ring of four tiles, else in global memory.
  // keep <= 116: 128-entry append-and-trim lists in global memory.  Above: sorted 128-entry lists in shared memory
  // (one half per CTA: half the selection warps per SM, ~1.5x slower).
  const size_t tile = (size_t)tc::TILE * K * 2;
  const size_t limit = 227 * 1024 - 1024;
  // Append-and-trim lists (keep <= 32u - 12) are write-mostly -- read only when a row is trimmed -- so they may live
  // in global memory, in the output arrays themselves, which leaves shared memory for two halves per CTA and a
  // deeper tile ring.
  auto fit = [&](int u, int glists, TcConfig* out) -> bool {
    if (glists && (shared_lists_only || keep > 32 * u - 12)) return false;
    if (keep > 32 * u) return false;
    for (int halves = 2; halves >= 1; --halves) {
      if (force_halves && halves != force_halves) continue;
      const int rows = halves * tc::TILE, epi = halves * 4;
      const size_t fixed = halves * tile + (glists ? 0 : (size_t)rows * 32 * u * 8) + (size_t)rows * 4 + (size_t)epi * 36 + 512;
      const int choices[6][2] = {{8, 32}, {4, 32}, {8, 16}, {4, 16}, {2, 32}, {2, 16}};  // {ring slots, queue entries}
      for (const auto& ch : choices) {
        if (force_slots && ch[0] != force_slots) continue;
        const size_t tot = fixed + (size_t)ch[0] * tile + (size_t)2 * epi * ch[1] * tc::QWORDS * 4;
        if (tot > limit) continue;
        out->u = u;
        out->halves = halves;
        out->slots = ch[0];
        out->qent = ch[1];
        out->keep = keep;
        out->glists = glists;
        out->smem = tot;
        return true;
      }
    }
    return false;
  };
  // candidates in order of preference; the first that runs two halves per CTA on a ring of >= 4 tiles wins,
  // else the first that fits at all
  int cand[3][2];  // {u, lists in global memory}
  int ncand = 0;
  auto add = [&](int u, int g) {
    cand[ncand][0] = u;
    cand[ncand][1] = g;
    ++ncand;
  };
  if (keep <= 32) {
    add(1, 0);
  } else if (keep <= 52) {
    add(2, 0);
    add(2, 1);
  } else if (keep <= 116) {
    add(4, 1);
    if (keep <= 64) add(2, 0);
    add(4, 0);
  } else {
    add(4, 0);
  }
  TcConfig first{}, t{};
  bool have_first = false;
  for (int ci = 0; ci < ncand; ++ci) {
    const int* cd = cand[ci];
    if (!fit(cd[0], cd[1], &t)) continue;
    if (t.halves == 2 && t.slots >= 4) {
      *c = t;
      return true;
    }
    if (!have_first) {
      first = t;
      have_first = true;
    }
  }
  if (have_first) *c = first;
  return have_first;
}

template <int HALVES>
int launch_thr(b200_model* m, const tc::KParams& kp, const TcConfig& cfg, cudaStream_t st) {
  const int threads = (HALVES * 12 + 3) * 32;
  const int grid = std::min(kp.qblocks, m->sm_count);
  auto kern = tc::knn_tc_kernel<2, HALVES, false, true>;
  B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg.smem));
  kern<<<grid, threads, cfg.smem, st>>>(kp);
  B200_CUDA(cudaGetLastError());
  return 0;
}

template <int U, int HALVES>
int launch_tc(b200_model* m, const tc::KParams& kp, const TcConfig& cfg, bool dbg, cudaStream_t st) {
  const int threads = (HALVES * 12 + 3) * 32;
  const int grid = std::min(kp.qblocks, m->sm_count);
  if (dbg) {
    auto kern = tc::knn_tc_kernel<U, HALVES, true>;
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg.smem));
    kern<<<grid, threads, cfg.smem, st>>>(kp);
  } else {
    auto kern = tc::knn_tc_kernel<U, HALVES, false>;
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg.smem));
    kern<<<grid, threads, cfg.smem, st>>>(kp);
  }
  B200_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace

namespace tc {
// Fixed threshold of each query row for the radius filter: a node K can only contain the query in its
// radius if  dot_K + (-c_K) < E - ||x||^2  in scaled units, with E the bound on the tensor dot's error;
// widened by the fp64 rounding of the centred form, rounded up to float.  Queries the fp16 image cannot
// hold get +inf (everything is flagged, the row overflows and goes to the dense scan); padding rows -inf.
__global__ void row_thr_kernel(const QMeta* __restrict__ meta, int64_t q, int64_t qpad, int d, double scale,
                               double data_cmax, float* __restrict__ thr) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= qpad) return;
  if (i >= q) {
    thr[i] = -CUDART_INF_F;
    return;
  }
  const QMeta mq = meta[i];
  if (!(mq.err < CUDART_INF)) {
    thr[i] = CUDART_INF_F;
    return;
  }
  const double cm = fmax(data_cmax, mq.cmax) * scale;
  const double slack = 1e-7 * (fabs(mq.qn2) + mq.err) + 1e-9 * (double)d * cm * cm;
  thr[i] = __double2float_ru(mq.err - mq.qn2 + slack);
}
__global__ void cand_rows_kernel(const int32_t* __restrict__ cand_pos, const int32_t* __restrict__ cand_cnt,
                                 const int32_t* __restrict__ perm, int64_t q, int W, int64_t n, int32_t* __restrict__ rows,
                                 int32_t* __restrict__ cnt_out) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= q * W) return;
  const int64_t p = e / W;
  const int j = (int)(e - p * W);
  const int c = cand_cnt[p];
  if (j == 0) cnt_out[p] = c;
  const int pos = (j < c && c <= W) ? cand_pos[e] : -1;
  rows[e] = (pos >= 0 && pos < n) ? perm[pos] : -1;
}
}  // namespace tc

// ShepMod's radius test as a tensor-core filter: for every query the positions (fitted rows, through
// perm) of all nodes that may hold it inside their radius -- a superset of {K : dist < RW(K)} -- at most
// 64 per query; cand_cnt = 65 marks a query with more (or one the filter cannot handle).
// cand_rows (q, 64) int32 fitted-row numbers, cand_cnt (q).
int shepmod_filter(b200_model* m, const double* z_dev, int64_t q, int32_t* cand_rows, int32_t* cand_cnt_out,
                   cudaStream_t st) {
  B200_TRY(tc_prepare(m, st, m->rw.as<double>()));
  const TcData& t = m->tc_sm;
  TcConfig cfg;
  // shape as for a 64-entry list (u = 2); keep is not used by the threshold kernel
  if (!tc_config(m->d, 40, 0, false, 0, &cfg, 0, /*shared_lists_only=*/true)) {  // the radius filter keeps its hit lists in shared memory
    set_error("radius filter unsupported for d=%d", m->d);
    return 1;
  }
  const int rows = cfg.halves * tc::TILE, W = 64;
  const int qblocks = (int)((q + rows - 1) / rows);
  const int64_t qpad = (int64_t)qblocks * rows;
  B200_TRY(m->zimg.reserve((size_t)qpad * t.K * 2));
  B200_TRY(m->cand.reserve((size_t)qpad * W * 4));
  B200_TRY(m->cand_meta.reserve((size_t)qpad * (12 + sizeof(tc::QMeta)) + 64));
  int32_t* cand_cnt = m->cand_meta.as<int32_t>();
  float* cand_tau = reinterpret_cast<float*>(cand_cnt + qpad);
  float* row_thr = cand_tau + qpad;
  tc::QMeta* meta = reinterpret_cast<tc::QMeta*>(row_thr + qpad);
  cudaEventRecord(m->ev3, st);  // (ev0 / ev1 belong to the caller's stage timer)
  tc::query_img_kernel<<<(unsigned)((qpad + 127) / 128), 128, 0, st>>>(z_dev, q, m->d, t.K, t.mu.as<double>(), t.scale, t.ymax,
                                                                        m->zimg.as<__half>(), qpad, meta);
  tc::row_thr_kernel<<<(unsigned)((qpad + 255) / 256), 256, 0, st>>>(meta, q, qpad, m->d, t.scale, t.cmax, row_thr);
  B200_CUDA(cudaGetLastError());
  tc::KParams kp{};
  kp.a_img = m->zimg.as<__half>();
  kp.b_img = t.img.as<__half>();
  kp.nlo = t.nlo.as<float>();
  kp.K = t.K;
  kp.ntiles = (int)t.tiles;
  kp.ncols = (int)m->n;
  kp.qblocks = qblocks;
  kp.slots = cfg.slots;
  kp.qent = cfg.qent;
  kp.parts = 1;
  kp.keep = W;
  kp.mode = 0;
  kp.cand_idx = m->cand.as<int32_t>();
  kp.cand_cnt = cand_cnt;
  kp.cand_tau = cand_tau;
  kp.row_thr = row_thr;
  kp.dbg = nullptr;
  kp.stats = nullptr;
  if (cfg.halves == 2) B200_TRY(launch_thr<2>(m, kp, cfg, st));
  else B200_TRY(launch_thr<1>(m, kp, cfg, st));
  // positions -> fitted rows, compacted to the caller's (q, 64) layout
  tc::cand_rows_kernel<<<(unsigned)((q * W + 255) / 256), 256, 0, st>>>(kp.cand_idx, cand_cnt, t.perm.as<int32_t>(), q, W, m->n,
                                                                        cand_rows, cand_cnt_out);
  B200_CUDA(cudaGetLastError());
  cudaEventRecord(m->ev2, st);
  B200_CUDA(cudaEventSynchronize(m->ev2));
  float ms1 = 0;
  cudaEventElapsedTime(&ms1, m->ev3, m->ev2);
  m->st.ms_stage1 += ms1;  // also inside the caller's ms_weights
  m->st.launches += 4;
  m->st.tc_queries += q;
  return 0;
}

bool tc_supported(const b200_model* m, int k) {
  TcConfig c;
  return tc_config(m->d, k, 0, true, 0, &c);
}

// the radius filter's shape: 64-entry hit lists in shared memory (the same call as in shepmod_filter)
bool shepmod_filter_supported(const b200_model* m) {
  TcConfig c;
  return tc_config(m->d, 40, 0, false, 0, &c, 0, /*shared_lists_only=*/true);
}

static bool t_tiles_ok(const b200_model* m, int split) { return m->tc.tiles >= (int64_t)split * 64; }

int knn_tc(b200_model* m, const double* z_dev, int64_t q, int k, int64_t self_offset,
           int32_t* idx_dev, double* d2_dev, cudaStream_t st) {
  if (q == 0) return 0;
  if (k < 1 || k > m->n - (self_offset >= 0 ? 1 : 0)) {
    set_error("knn: k=%d out of range for n=%lld", k, (long long)m->n);
    return 1;
  }
  B200_TRY(tc_prepare(m, st));
  TcConfig cfg;
  if (!tc_config(m->d, k, m->margin, self_offset >= 0, m->tc_halves, &cfg, m->tc_slots)) {
    set_error("tensor kNN path unsupported for d=%d k=%d", m->d, k);
    return 1;
  }
  // Wave quantisation: the persistent grid takes sm_count 256-query blocks per round.  When the last
  // round would keep at most half the SMs busy, those queries go to a second call whose data stream is
  // split over S concurrent launches on the otherwise idle SMs, S candidate lists per query merged by
  // the re-rank.  8 GPUs x 125 k queries: 3.3 rounds instead of 4.
  if (cfg.halves == 2 && cfg.u == 1 && m->tc_halves == 0 && m->tc_split == 0 && m->tc_mode == 0 && !m->dbg_scores &&
      !m->no_tail_split) {
    const int64_t blocks = (q + 2 * tc::TILE - 1) / (2 * tc::TILE);
    const int64_t rounds = blocks / m->sm_count, tail = blocks - rounds * m->sm_count;
    if (rounds >= 1 && tail > 0 && 2 * tail <= m->sm_count) {
      const int split = (int)std::min<int64_t>(4, m->sm_count / tail);
      if (split >= 2 && t_tiles_ok(m, split)) {
        const int64_t q_main = rounds * m->sm_count * 2 * tc::TILE;
        B200_TRY(knn_tc(m, z_dev, q_main, k, self_offset, idx_dev, d2_dev, st));
        m->tc_split = split;
        m->defer_base += q_main;
        const int rc = knn_tc(m, z_dev + (size_t)q_main * m->d, q - q_main, k,
                              self_offset >= 0 ? self_offset + q_main : -1, idx_dev + (size_t)q_main * k,
                              d2_dev + (size_t)q_main * k, st);
        m->defer_base -= q_main;
        m->tc_split = 0;
        return rc;
      }
    }
  }
  const TcData& t = m->tc;
  const int rows = cfg.halves * tc::TILE;
  const int W = 32 * cfg.u;
  const int qblocks = (int)((q + rows - 1) / rows);
  const int64_t qpad = (int64_t)qblocks * rows;
  const int S = m->tc_split > 1 ? m->tc_split : 1;  // concurrent launches over sub-ranges of the data stream
  B200_TRY(m->zimg.reserve((size_t)qpad * t.K * 2));
  B200_TRY(m->cand.reserve((size_t)S * qpad * W * 4 * (cfg.glists ? 2 : 1)));  // glists: scores behind the positions
  B200_TRY(m->cand_meta.reserve((size_t)qpad * (8 * S + sizeof(tc::QMeta)) + 64));
  B200_TRY(m->flags.reserve(256));
  B200_TRY(m->misc.reserve((size_t)q * 4 + 64));
  int32_t* cand_cnt = m->cand_meta.as<int32_t>();
  float* cand_tau = reinterpret_cast<float*>(cand_cnt + (size_t)S * qpad);
  tc::QMeta* meta = reinterpret_cast<tc::QMeta*>(cand_tau + (size_t)S * qpad);
  int32_t* fail_cnt = m->flags.as<int32_t>() + 8;
  int32_t* fail_list = m->misc.as<int32_t>();
  float* dbg = m->dbg_scores;

  cudaEventRecord(m->ev3, st);
  tc::query_img_kernel<<<(unsigned)((qpad + 127) / 128), 128, 0, st>>>(z_dev, q, m->d, t.K, t.mu.as<double>(), t.scale, t.ymax,
                                                                        m->zimg.as<__half>(), qpad, meta);
  B200_CUDA(cudaGetLastError());
  cudaEventRecord(m->ev0, st);
  tc::KParams kp{};
  kp.a_img = m->zimg.as<__half>();
  kp.b_img = t.img.as<__half>();
  kp.nlo = t.nlo.as<float>();
  kp.K = t.K;
  kp.ntiles = (int)t.tiles;
  kp.ncols = (int)m->n;
  kp.qblocks = qblocks;
  kp.slots = cfg.slots;
  kp.qent = cfg.qent;
  kp.parts = m->tc_parts;
  kp.keep = cfg.keep;
  kp.mode = m->tc_mode;
  kp.cand_idx = m->cand.as<int32_t>();
  kp.cand_cnt = cand_cnt;
  kp.cand_tau = cand_tau;
  kp.glists = cfg.glists;
  kp.cand_sc = cfg.glists ? reinterpret_cast<float*>(kp.cand_idx + (size_t)S * qpad * W) : nullptr;
  kp.dbg = dbg;
  kp.stats = nullptr;
  if (m->tc_stats) {  // instrumented build: event and stall counters (tuning only)
    kp.stats = reinterpret_cast<unsigned long long*>(m->flags.as<char>() + 64);
    B200_CUDA(cudaMemsetAsync(kp.stats, 0, 128, st));
  }
  const bool dg = dbg != nullptr || kp.stats != nullptr || m->tc_mode != 0;  // ceilings run in the instrumented build
  auto launch_one = [&](const tc::KParams& kq, cudaStream_t s_) -> int {
    if (cfg.u == 1 && cfg.halves == 2) return launch_tc<1, 2>(m, kq, cfg, dg, s_);
    if (cfg.u == 1 && cfg.halves == 1) return launch_tc<1, 1>(m, kq, cfg, dg, s_);
    if (cfg.u == 2 && cfg.halves == 2) return launch_tc<2, 2>(m, kq, cfg, dg, s_);
    if (cfg.u == 2 && cfg.halves == 1) return launch_tc<2, 1>(m, kq, cfg, dg, s_);
    if (cfg.u == 4 && cfg.halves == 2) return launch_tc<4, 2>(m, kq, cfg, dg, s_);
    if (cfg.u == 4 && cfg.halves == 1) return launch_tc<4, 1>(m, kq, cfg, dg, s_);
    set_error("tensor kNN: no kernel for u=%d halves=%d", cfg.u, cfg.halves);
    return 1;
  };
  int rc = 1;
  tc::RerankLists L;
  L.nlists = S;
  L.list_elems = (int64_t)qpad * W;
  L.meta_elems = qpad;
  for (int r = 0; r < 4; ++r) L.off[r] = 0;
  if (S == 1) {
    rc = launch_one(kp, st);
  } else {
    for (int r = 0; r < S; ++r) {
      if (!m->aux_stream[r]) {
        B200_CUDA(cudaStreamCreateWithFlags(&m->aux_stream[r], cudaStreamNonBlocking));
        B200_CUDA(cudaEventCreateWithFlags(&m->aux_event[r], cudaEventDisableTiming));
      }
      const int64_t lo = t.tiles * r / S, hi = t.tiles * (r + 1) / S;
      tc::KParams kq = kp;
      kq.b_img = kp.b_img + (size_t)lo * tc::TILE * t.K;
      kq.nlo = kp.nlo + (size_t)lo * 4;
      kq.ntiles = (int)(hi - lo);
      kq.ncols = (int)std::min<int64_t>(m->n - lo * tc::TILE, (hi - lo) * tc::TILE);
      kq.cand_idx = kp.cand_idx + (size_t)r * L.list_elems;
      kq.cand_cnt = kp.cand_cnt + (size_t)r * qpad;
      kq.cand_tau = kp.cand_tau + (size_t)r * qpad;
      L.off[r] = (int)(lo * tc::TILE);
      B200_CUDA(cudaStreamWaitEvent(m->aux_stream[r], m->ev0, 0));  // the query images are ready
      rc = launch_one(kq, m->aux_stream[r]);
      if (rc) break;
      B200_CUDA(cudaEventRecord(m->aux_event[r], m->aux_stream[r]));
      B200_CUDA(cudaStreamWaitEvent(st, m->aux_event[r], 0));
    }
  }
  B200_TRY(rc);
  cudaEventRecord(m->ev1, st);
  m->st.last_cap = W;
  m->st.last_keep = cfg.keep;
  m->st.last_slots = cfg.slots;
  m->st.last_halves = cfg.halves;
  if (m->tc_mode != 0) {  // ceilings: outputs are meaningless, only stage 1 is timed
    B200_CUDA(cudaEventSynchronize(m->ev1));
    float msd = 0;
    cudaEventElapsedTime(&msd, m->ev0, m->ev1);
    m->st.ms_stage1 += msd;
    m->st.launches += 2;
    return 0;
  }
  B200_CUDA(cudaMemsetAsync(fail_cnt, 0, 4, st));
  const unsigned rb = (unsigned)((q + 7) / 8);
#define RERANK(CAPV)                                                                                     \
  tc::rerank_kernel<CAPV><<<rb, 256, (size_t)8 * ((m->d + 1) & ~1) * 8, st>>>(                           \
      m->x, m->n, m->d, z_dev, q, k, self_offset, kp.cand_idx, W, cand_cnt, cand_tau, meta, t.scale, t.cmax, \
      t.perm.as<int32_t>(), t.xs.as<double>(), idx_dev, d2_dev, fail_list, fail_cnt, L)
  if (S > 1) RERANK(128);  // pass r of the 128-slot instance takes list r (u = 1: at most 32 entries each)
  else if (cfg.u == 1) RERANK(32);
  else if (cfg.u == 2) RERANK(64);
  else RERANK(128);
#undef RERANK
  B200_CUDA(cudaGetLastError());
  cudaEventRecord(m->ev2, st);
  int h_fail = 0;
  B200_CUDA(cudaMemcpyAsync(&h_fail, fail_cnt, 4, cudaMemcpyDeviceToHost, st));
  B200_CUDA(cudaStreamSynchronize(st));
  if (kp.stats) {
    unsigned long long hs[16];
    B200_CUDA(cudaMemcpy(hs, kp.stats, 128, cudaMemcpyDeviceToHost));
    for (int i = 0; i < 16; ++i) m->tc_stat_vals[i] += (double)hs[i];
  }
  float ms1 = 0, ms2 = 0, ms0 = 0;
  cudaEventElapsedTime(&ms1, m->ev0, m->ev1);
  cudaEventElapsedTime(&ms2, m->ev1, m->ev2);
  cudaEventElapsedTime(&ms0, m->ev3, m->ev0);
  m->st.ms_prep += ms0;
  m->st.ms_stage1 += ms1;
  m->st.ms_rerank += ms2;
  m->st.launches += 3 + (S - 1);  // query image, S stage-1 launches, re-rank
  m->st.tc_queries += q;
  m->st.uncertified += h_fail;
  if (h_fail > 0 && m->defer_fallback && h_fail <= 64 && m->deferred.size() + h_fail <= 512) {
    // the caller resolves these after its last chunk (one pass over the fitted points instead of one per chunk)
    std::vector<int32_t> loc(h_fail);
    B200_CUDA(cudaMemcpy(loc.data(), fail_list, (size_t)h_fail * 4, cudaMemcpyDeviceToHost));
    for (int32_t v : loc) m->deferred.push_back((int32_t)(v + m->defer_base));
    return 0;
  }
  if (h_fail > 0) {
    cudaEventRecord(m->ev0, st);
    if (m->force_split_fallback) B200_TRY(knn_exact(m, z_dev, q, k, self_offset, idx_dev, d2_dev, fail_list, h_fail, st));
    else B200_TRY(knn_rescan(m, z_dev, q, k, self_offset, idx_dev, d2_dev, fail_list, h_fail, st));
    cudaEventRecord(m->ev1, st);
    B200_CUDA(cudaEventSynchronize(m->ev1));
    cudaEventElapsedTime(&ms1, m->ev0, m->ev1);
    m->st.ms_exact += ms1;
  }
  return 0;
}

}  // namespace b200
