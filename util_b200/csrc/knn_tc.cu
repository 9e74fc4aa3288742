// knn_tc.cu -- tensor-core kNN: fused distance + candidate selection, fp64 re-rank.
//
// Replaces BallTree.query over the whole batch (nearest_neighbor/__init__.py:75-76)
// for large problems.  Three stages:
//
//  1. knn_tc_kernel (tcgen05 / TMEM / bulk-TMA, persistent, warp specialised):
//     score(q, j) = ||y_j||^2 - 2 x_q . y_j as ONE contraction
//         [ -2 x_q | 1 1 1 | 0.. ] . [ y_j | n_hi n_mid n_lo | 0.. ],   K = pad16(d + 3)
//     on mean-centred, power-of-two scaled fp16 operands with fp32 accumulation
//     in TMEM (the squared norm rides in three extra K slots as a 33-bit fp16
//     triple, so the epilogue needs no add).  Query rows are the M side: each
//     epilogue thread owns one TMEM lane = one query and streams every data
//     tile past a running threshold tau (min3 trees over 8-column groups);
//     survivors go to a per-row shared-memory buffer that a warp-cooperative
//     rank-select trims to the C best whenever it fills.  The distance matrix
//     is never written to HBM.
//  2. rerank_kernel: exact fp64 squared distances of the <= CAP candidates of
//     each query with the reference's arithmetic (direct differences, dimension
//     order, no FMA), sorted by (d2, index) -> bit-exact neighbours.  Each query
//     is CERTIFIED: every non-candidate has approximate score >= tau, so with a
//     rigorous bound E on |approximate - exact| score, exact d2_k < (tau - E +
//     ||x_q||^2)/s^2 proves no non-candidate can enter the top k.
//  3. Queries that fail the certificate are resolved exactly (knn_exact.cu: a rescan
//     listing every point within the candidates' k-th distance when they are few, the
//     split brute-force scan otherwise) -- so the result is always the exact answer.
//
// The images hold the fitted points in order of increasing centred norm (position ->
// row map `perm`, applied by the re-rank): thresholds tighten early, fewer survivors.
//
// Operand layout in HBM: per 128-row tile a ready-to-use UMMA image
// [K/8][128][8 halfs] (K-major, no swizzle: 8x16-byte core matrices contiguous),
// so one cp.async.bulk (TMA engine) brings a whole tile and the smem descriptor
// is fixed: LBO = 2048 B (next K chunk), SBO = 128 B (next 8 rows).
#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"
#include "ptx.cuh"

namespace b200 {

namespace tc {

constexpr int TILE = 128;          // query rows per accumulator half = data columns per epilogue tile
constexpr int NORM_SLOTS = 3;
constexpr double U16 = 4.8828125e-4;             // 2^-11: fp16 unit roundoff
constexpr double TINY16 = 2.98023223876953125e-8;  // 2^-25: half the smallest fp16 subnormal
constexpr double GAMMA_ACC = 1.52587890625e-5;   // 2^-16: bound on tensor-core fp32 accumulation error, relative to sum |terms|
constexpr float PAD_NORM = 60000.0f;             // squared norm given to padding rows: never selected

__host__ __device__ inline int padK(int d) { return ((d + NORM_SLOTS + 15) / 16) * 16; }

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
// early and far fewer later points survive the filter (about half as many after the first 1 % of the
// stream on uniform data) -- any order gives the same exact result.  Key: fp32 bits of the squared
// norm (non-negative, so unsigned order = numeric order).
__global__ void norm_key_kernel(const double* __restrict__ x, int64_t n, int d, const double* __restrict__ mu,
                                double scale, uint32_t* __restrict__ key, int32_t* __restrict__ val) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double n2 = 0.0;
  for (int j = 0; j < d; ++j) {
    const double yc = (x[i * d + j] - mu[j]) * scale;
    n2 = fma(yc, yc, n2);
  }
  key[i] = __float_as_uint((float)n2);
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
// Data side: one thread per (padded) point; writes its row of the tile image and
// folds max(||yc||, ||fp16(yc)||) into ymax.
__global__ void data_img_kernel(const double* __restrict__ x, int64_t n, int d, int K, int SUBROWS, int rg,
                                const double* __restrict__ mu, double scale, const int32_t* __restrict__ perm,
                                __half* __restrict__ img, int64_t npad, unsigned long long* __restrict__ ymax_bits) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;  // image position
  double nm = 0.0;
  if (i < npad) {
    const int64_t src = i < n ? (int64_t)perm[i] : 0;  // fitted row that sits at this position
    const int64_t t = i / SUBROWS;
    const int r = (int)(i - t * SUBROWS);
    uint4* dst = reinterpret_cast<uint4*>(img + (size_t)t * SUBROWS * K);  // chunk c at dst[c*SUBROWS + r]
    const int KC = K / 8;
    double n2 = 0.0, n2h = 0.0;
    for (int c = 0; c < KC; ++c) {
      __half h[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int j = c * 8 + e;
        float hv = 0.0f;
        if (i < n && j < d) {
          const double yc = (x[src * d + j] - mu[j]) * scale;
          const __half hh = __double2half(yc);
          n2 = fma(yc, yc, n2);
          const double yh = (double)__half2float(hh);
          n2h = fma(yh, yh, n2h);
          h[e] = hh;
          continue;
        }
        if (j >= d && j < d + NORM_SLOTS) {
          if (i >= n) hv = (j == d) ? PAD_NORM : 0.0f;
          else {
            // 33-bit fp16 triple of the exact (fp64) squared norm
            const __half a = __double2half(n2);
            const double r1 = n2 - (double)__half2float(a);
            const __half b = __double2half(r1);
            const double r2 = r1 - (double)__half2float(b);
            const __half c3 = __double2half(r2);
            hv = __half2float(j == d ? a : (j == d + 1 ? b : c3));
          }
        }
        h[e] = __float2half_rn(hv);
      }
      dst[rg ? ((size_t)(r >> 3) * KC + c) * 8 + (r & 7) : (size_t)c * SUBROWS + r] = *reinterpret_cast<uint4*>(h);
    }
    nm = sqrt(fmax(n2, n2h));
  }
  for (int o = 16; o > 0; o >>= 1) nm = fmax(nm, __shfl_xor_sync(0xffffffffu, nm, o));
  if ((threadIdx.x & 31) == 0) atomicMax(ymax_bits, (unsigned long long)__double_as_longlong(nm));
}

struct QMeta {
  double qn2;   // ||(z - mu) * scale||^2
  double err;   // E_q: bound on |approximate - exact| score (scaled units), +inf when fp16 overflows
  double cmax;  // max |z_j|
};

// Query side: x~ = fp16(-2 * (z - mu) * scale), then the three 1.0 slots.
__global__ void query_img_kernel(const double* __restrict__ z, int64_t q, int d, int K, int rg,
                                 const double* __restrict__ mu, double scale, double ymax,
                                 __half* __restrict__ img, int64_t qpad, QMeta* __restrict__ meta) {
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
      float hv = 0.0f;
      if (i < q && j < d) {
        const double zj = z[i * d + j];
        const double zc = (zj - mu[j]) * scale;
        const double xv = -2.0 * zc;
        qn2 = fma(zc, zc, qn2);
        xn2 = fma(xv, xv, xn2);
        x1 += fabs(xv);
        cmax = fmax(cmax, fabs(zj));
        if (!(fabs(xv) < 60000.0)) bad = true;  // also catches NaN
        h[e] = __double2half(xv);
        continue;
      }
      if (i < q && j >= d && j < d + NORM_SLOTS) hv = 1.0f;
      h[e] = __float2half_rn(hv);
    }
    dst[rg ? ((size_t)(r >> 3) * KC + c) * 8 + (r & 7) : (size_t)c * TILE + r] = *reinterpret_cast<uint4*>(h);
  }
  if (i < q) {
    const double xn = sqrt(xn2) * 1.001;
    // rounding of both operands (Cauchy-Schwarz), fp16 subnormal floor, the norm triple,
    // and the tensor-core accumulation -- see DESIGN.md "certificate"
    double E = 2.0 * U16 * xn * ymax + TINY16 * (x1 + sqrt((double)d) * ymax) + 2.0 * TINY16 +
               GAMMA_ACC * (xn * ymax + ymax * ymax);
    if (bad) E = CUDART_INF;
    meta[i].qn2 = qn2;
    meta[i].err = E;
    meta[i].cmax = cmax;
  }
}


// ------------------------------------------------------------------ main kernel
struct KParams {
  const __half* a_img;   // query tile images  [qtiles][K/8][128][8]
  const __half* b_img;   // data sub-tile images [2*ntiles][K/8][64][8]
  int K;                 // contraction length, multiple of 16
  int ntiles;            // 128-column data tiles (two 64-row sub-tiles each)
  int qblocks;           // query blocks of HALVES*128 rows
  int sub;               // data rows per B sub-tile: 64 or 128 (one TMA copy, one UMMA N)
  int flags;             // bit 0: mbarrier waits pass a suspend-time hint
  int slots;             // depth of the B ring, in sub-tiles
  int stages;            // TS kernel: 64-column accumulator stages per half
  int keep;              // C: candidates a compaction keeps (<= CAP - 8)
  int mode;              // 0 normal; debug ceilings: 1 = epilogue reads TMEM only, 2 = epilogue only releases
  int32_t* cand_idx;     // [qpad][CAP]
  int32_t* cand_cnt;     // [qpad]
  float* cand_tau;       // [qpad]
  float* dbg;            // optional [qpad][ntiles*128] raw scores (tests only)
  int trace_start;       // first tile recorded in `trace`
  long long* trace;      // optional [3 roles][TRACE_TILES][4] clock64 stamps of CTA 0 (tuning only)
};
constexpr int TRACE_TILES = 96;
#define B200_TRACE(role, tile, ev)                                                         \
  do {                                                                                     \
    const int tt_ = (int)(tile) - p.trace_start;                                           \
    if (p.trace && blockIdx.x == 0 && tt_ >= 0 && tt_ < TRACE_TILES)                       \
      p.trace[((role) * TRACE_TILES + tt_) * 4 + (ev)] = clock64();                        \
  } while (0)


__device__ __forceinline__ float min3(float a, float b, float c) { return fminf(fminf(a, b), c); }

// Trim buffer column `col` (one query row) to its `keep` best entries; returns the new
// threshold (the keep-th smallest score).  Warp-cooperative, all lanes must call.
// bsc/bid are [CAP][RS] (slot-major, RS = rows + 1 keeps slot walks conflict-free).
// Rank by counting: lane l ranks entries l and l+32 against broadcast reads of all n entries.
// Equal scores would share a rank; they are detected by the rank checksum and re-ranked with
// the slot number as tie-break.
template <int CAP>
__device__ __forceinline__ float compact_row(float* bsc, int* bid, int RS, int col, int n_valid, int keep,
                                             int lane) {
  constexpr int U = (CAP + 31) / 32;
  float sc[U];
  int id[U], rank[U];
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const int j = lane + 32 * u;
    sc[u] = j < n_valid ? bsc[j * RS + col] : CUDART_INF_F;
    id[u] = j < n_valid ? bid[j * RS + col] : 0;
    rank[u] = 0;
  }
  const float* ps = bsc + col;
  int i = 0;
  for (; i + 8 <= n_valid; i += 8) {
    float s8[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) s8[e] = ps[(i + e) * RS];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      int r0 = 0, r1 = 0;
#pragma unroll
      for (int e = 0; e < 8; e += 2) {
        r0 += s8[e] < sc[u] ? 1 : 0;
        r1 += s8[e + 1] < sc[u] ? 1 : 0;
      }
      rank[u] += r0 + r1;
    }
  }
  for (; i < n_valid; ++i) {
    const float s0 = ps[i * RS];
#pragma unroll
    for (int u = 0; u < U; ++u) rank[u] += s0 < sc[u] ? 1 : 0;
  }
  int sum = 0;
#pragma unroll
  for (int u = 0; u < U; ++u) sum += (lane + 32 * u < n_valid) ? rank[u] : 0;
  sum = __reduce_add_sync(0xffffffffu, sum);
  if (sum != n_valid * (n_valid - 1) / 2) {  // equal scores: redo with the slot number as tie-break
#pragma unroll
    for (int u = 0; u < U; ++u) rank[u] = 0;
    for (i = 0; i < n_valid; ++i) {
      const float s0 = ps[i * RS];
#pragma unroll
      for (int u = 0; u < U; ++u) rank[u] += (s0 < sc[u] || (s0 == sc[u] && i < lane + 32 * u)) ? 1 : 0;
    }
  }
  __syncwarp();
  float tau = 0.0f;
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const bool valid = lane + 32 * u < n_valid;
    if (valid && rank[u] < keep) {
      bsc[rank[u] * RS + col] = sc[u];
      bid[rank[u] * RS + col] = id[u];
    }
    const unsigned bm = __ballot_sync(0xffffffffu, valid && rank[u] == keep - 1);
    if (bm) tau = __shfl_sync(0xffffffffu, sc[u], __ffs(bm) - 1);
  }
  __syncwarp();
  return tau;
}

// Survivors of one 32-column chunk held in registers v[32] -> candidate buffer of my row.
// mask: values below my threshold (32 independent compares, four short OR chains).  A row whose
// buffer cannot take all its survivors is trimmed FIRST (warp-cooperative compact_row: keep best
// entries, threshold lowered to the keep-th), the mask is taken again under the new threshold, and
// only then are the survivors stored -- nothing is ever dropped, so the buffer holds every value
// below the final threshold.  In the first tiles one chunk can hold more survivors than the
// CAP - keep (>= 16) free slots a trim leaves: the chunk is then taken in two 16-column halves.
// The usual case late in the stream is one survivor, the chunk minimum `mcc` already held.
// Uses: tau, cnt, keep, lane, my_sc, my_id, bsc, bid, RS, row0, CAP (of the enclosing kernel).
#define B200_MASK32(v, out)                                                                  \
  do {                                                                                       \
    unsigned m4_[4] = {0u, 0u, 0u, 0u};                                                      \
    _Pragma("unroll") for (int i_ = 0; i_ < 32; ++i_) m4_[i_ & 3] |= v[i_] < tau ? (1u << i_) : 0u; \
    out = (m4_[0] | m4_[1]) | (m4_[2] | m4_[3]);                                             \
  } while (0)
#define B200_APPEND_CHUNK(v, base, mcc)                                                      \
  do {                                                                                       \
    unsigned todo_ = 0xffffffffu;                                                            \
    do {                                                                                     \
      unsigned cur_ = todo_, mask_;                                                          \
      B200_MASK32(v, mask_);                                                                 \
      mask_ &= todo_;                                                                        \
      unsigned need_ = __ballot_sync(0xffffffffu, cnt + __popc(mask_) > CAP);                \
      if (need_) {                                                                           \
        while (need_) {                                                                      \
          const int r_ = __ffs(need_) - 1;                                                   \
          need_ &= need_ - 1;                                                                \
          const int nv_ = __shfl_sync(0xffffffffu, cnt, r_);                                 \
          const float nt_ = compact_row<CAP>(bsc, bid, RS, row0 + r_, nv_, keep, lane);      \
          if (lane == r_) {                                                                  \
            tau = fminf(tau, nt_);                                                           \
            cnt = keep;                                                                      \
          }                                                                                  \
        }                                                                                    \
        B200_MASK32(v, mask_);                                                               \
        mask_ &= todo_;                                                                      \
        if (todo_ == 0xffffffffu && __any_sync(0xffffffffu, cnt + __popc(mask_) > CAP)) {    \
          cur_ = 0x0000ffffu;                                                                \
          mask_ &= cur_;                                                                     \
        }                                                                                    \
      }                                                                                      \
      const int nh_ = __popc(mask_);                                                         \
      if (nh_ == 1 && cur_ == 0xffffffffu) {                                                 \
        if (cnt < CAP) {                                                                     \
          my_sc[cnt * RS] = mcc;                                                             \
          my_id[cnt * RS] = (base) + __ffs(mask_) - 1;                                       \
          ++cnt;                                                                             \
        } else {                                                                             \
          tau = mcc; /* cannot happen (space was made above); kept so the filter stays valid */ \
        }                                                                                    \
      } else if (nh_ > 0) {                                                                  \
        _Pragma("unroll") for (int i_ = 0; i_ < 32; ++i_) {                                  \
          const bool lt_ = (mask_ >> i_) & 1u;                                               \
          const bool put_ = lt_ && cnt < CAP;                                                \
          if (put_) {                                                                        \
            my_sc[cnt * RS] = v[i_];                                                         \
            my_id[cnt * RS] = (base) + i_;                                                   \
          }                                                                                  \
          tau = (lt_ && !put_) ? v[i_] : tau;                                                \
          cnt += put_ ? 1 : 0;                                                               \
        }                                                                                    \
      }                                                                                      \
      todo_ &= ~cur_;                                                                        \
    } while (todo_);                                                                         \
  } while (0)

template <int CAP, int HALVES, bool DBG>
__global__ void __launch_bounds__((HALVES * 4 + 3) * 32, 1) knn_tc_kernel(const KParams p) {
  constexpr int ROWS = HALVES * TILE;
  constexpr int RS = ROWS + 1;
  constexpr int EPI_WARPS = HALVES * 4;
  constexpr uint32_t ACC_COLS = HALVES * TILE;  // TMEM columns per accumulator stage
  extern __shared__ __align__(1024) uint8_t smem[];
  const int K = p.K;
  const uint32_t tile_bytes = (uint32_t)TILE * K * 2;  // one 128-row query tile
  const int SUB = p.sub;
  const uint32_t sub_bytes = (uint32_t)SUB * K * 2;    // one data sub-tile
  uint8_t* sA = smem;                                  // HALVES query tiles
  uint8_t* sB = sA + (size_t)HALVES * tile_bytes;      // ring of p.slots sub-tiles
  float* bsc = reinterpret_cast<float*>(sB + (size_t)p.slots * sub_bytes);  // [CAP][RS] scores
  int* bid = reinterpret_cast<int*>(bsc + (size_t)CAP * RS);                // [CAP][RS] data indices
  uint64_t* bars = reinterpret_cast<uint64_t*>((reinterpret_cast<uintptr_t>(bid + (size_t)CAP * RS) + 15) & ~(uintptr_t)15);
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
      ptx::mbar_init(ptx::smem_u32(&bar_acce[a]), 4);  // the four warps of one half
    }
    ptx::mbar_init(ptx::smem_u32(bar_afull), 1);
    ptx::mbar_init(ptx::smem_u32(bar_aempty), 2);  // one commit per issuer warp
    ptx::fence_barrier_init();
  }
  if (warp == EPI_WARPS + 1) {  // MMA warp owns the TMEM allocation
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
  const int per_tile = TILE / SUB;
  const int nsub = per_tile * p.ntiles;
  const bool hint = p.flags & 1;
  const uint32_t bsleep = (p.flags & 16) ? 0 : 64;  // ns between barrier probes of the producer / MMA warps

  if (warp == EPI_WARPS) {
    // ===================== producer: bulk copies on the TMA engine =====================
    if (lane == 0) {
      uint32_t jt = 0, s = 0, ph = 0;
      const uint32_t full0 = ptx::smem_u32(&bar_full[0]), empty0 = ptx::smem_u32(&bar_empty[0]);
      const uint32_t sB0 = ptx::smem_u32(sB);
      for (int b = 0; b < nqb; ++b) {
        const int qb = (int)blockIdx.x + b * (int)gridDim.x;
        ptx::mbar_wait(hint, ptx::smem_u32(bar_aempty), (b & 1) ^ 1, bsleep);
        ptx::mbar_arrive_expect_tx(ptx::smem_u32(bar_afull), HALVES * tile_bytes);
        ptx::bulk_g2s(ptx::smem_u32(sA), reinterpret_cast<const uint8_t*>(p.a_img) + (size_t)qb * HALVES * tile_bytes,
                      HALVES * tile_bytes, ptx::smem_u32(bar_afull));
        const uint8_t* src = reinterpret_cast<const uint8_t*>(p.b_img);
        for (int st = 0; st < nsub; ++st, ++jt, src += sub_bytes) {
          ptx::mbar_wait(hint, empty0 + s * 8, ph ^ 1, bsleep);
          B200_TRACE(0, (int)jt, 0);
          if ((p.mode & 4) && jt >= (uint32_t)p.slots) {  // debug: no copy, the slot keeps its old tile
            ptx::mbar_arrive(full0 + s * 8);
          } else {
            ptx::mbar_arrive_expect_tx(full0 + s * 8, sub_bytes);
            ptx::bulk_g2s(sB0 + s * sub_bytes, src, sub_bytes, full0 + s * 8);
          }
          if (++s == (uint32_t)p.slots) {
            s = 0;
            ph ^= 1;
          }
        }
      }
    }
  } else if (warp > EPI_WARPS) {
    // ===================== MMA issuers: two warps, alternate tiles =====================
    // An issuing thread spends, per 128-column tile, ~730 clk issuing 10 MMAs into the short tensor
    // queue plus ~700 clk in commits and barrier waits during which it queues nothing (measured with
    // the clock trace, profiles/r01).  Two issuers on alternate tiles (issuer j owns accumulator
    // stage j) interleave those phases so the tensor pipe always has work.  Each warp runs in uniform
    // control flow; one elected lane issues the MMAs and the commits that track them.
    {
      const int j = warp - (EPI_WARPS + 1);  // issuer index = accumulator stage
      const uint32_t idesc = ptx::umma_idesc_f16(TILE, SUB);
      // operand image layout: chunk-major [K/8][rows][16 B] or row-group-major [rows/8][K/8][8][16 B]
      const bool rg = p.flags & 2;
      const uint32_t lbo_a = rg ? 128 : TILE * 16, sbo_a = rg ? (K / 8) * 128 : 128, kstep_a = rg ? 256 : TILE * 32;
      const uint32_t lbo_b = rg ? 128 : SUB * 16, sbo_b = sbo_a, kstep_b = rg ? 256 : SUB * 32;
      const bool skip_mma = p.mode & 8;
      const int ksteps = K / 16;
      uint64_t a_desc[HALVES];
#pragma unroll
      for (int h = 0; h < HALVES; ++h) a_desc[h] = ptx::umma_desc(ptx::smem_u32(sA + (size_t)h * tile_bytes), lbo_a, sbo_a);
      const uint64_t b_desc0 = ptx::umma_desc(ptx::smem_u32(sB), lbo_b, sbo_b);
      const uint32_t a_step = kstep_a >> 4, b_step = kstep_b >> 4, slot_step = sub_bytes >> 4;
      const uint32_t full0 = ptx::smem_u32(&bar_full[0]), empty0 = ptx::smem_u32(&bar_empty[0]);
      const uint32_t accf = ptx::smem_u32(&bar_accf[j * 2]), acce = ptx::smem_u32(&bar_acce[j * 2]);
      const uint32_t slot_mask = p.slots - 1, slot_shift = 31 - __clz(p.slots);  // slots is a power of two
      // One thread per issuer warp, lean barrier probes, no per-tile parameter reloads: the per-tile
      // serial path of the issuer is what bounds the pipeline once the epilogue keeps up.
      if (lane == 0) {
        uint32_t it0 = 0;   // global index of the current block's first tile
        uint32_t mine = 0;  // tiles this issuer has done (phase of its accumulator stage)
        const uint32_t afull = ptx::smem_u32(bar_afull), aempty = ptx::smem_u32(bar_aempty);
        const uint32_t ntiles = (uint32_t)p.ntiles;
        for (int b = 0; b < nqb; ++b, it0 += ntiles) {
          while (!ptx::mbar_try_wait(afull, b & 1)) {}
          for (uint32_t t = (j + it0) & 1; t < ntiles; t += 2, ++mine) {
            const uint32_t it = it0 + t;
            for (int hs = 0; hs < per_tile; ++hs) {
              const uint32_t jt = it * per_tile + hs, s = jt & slot_mask, ph = (jt >> slot_shift) & 1;
              const uint64_t bd = b_desc0 + (uint64_t)(s * slot_step);
#pragma unroll
              for (int h = 0; h < HALVES; ++h) {
                // the accumulator half is free once its four epilogue warps have read the previous tile
                if (hs == 0) {
                  while (!ptx::mbar_try_wait(acce + h * 8, (mine & 1) ^ 1)) {}
                }
                if (h == 0) {
                  while (!ptx::mbar_try_wait(full0 + s * 8, ph)) {}
                }
                ptx::tc_fence_after();
                // (the accumulator base is the literal 0: this CTA owns all 512 columns, checked at start)
                const uint32_t d_tmem = j * ACC_COLS + h * TILE + hs * SUB;
                if (DBG && h == 0) B200_TRACE(1, (int)it, 0);
                if (!(DBG && skip_mma)) {
#define B200_ISSUE(KS)                                                                                   \
  _Pragma("unroll") for (int ks = 0; ks < KS; ++ks)                                                       \
    ptx::umma_f16(d_tmem, a_desc[h] + (uint64_t)(ks * a_step), bd + (uint64_t)(ks * b_step), idesc,        \
                  ks > 0 ? 1u : 0u);
                  switch (ksteps) {
                    case 2: B200_ISSUE(2) break;
                    case 3: B200_ISSUE(3) break;
                    case 4: B200_ISSUE(4) break;
                    case 5: B200_ISSUE(5) break;
                    default:
                      for (int ks = 0; ks < ksteps; ++ks)
                        ptx::umma_f16(d_tmem, a_desc[h] + (uint64_t)(ks * a_step), bd + (uint64_t)(ks * b_step),
                                      idesc, ks > 0 ? 1u : 0u);
                  }
#undef B200_ISSUE
                }
                if (hs == per_tile - 1) ptx::umma_commit(accf + h * 8);  // this half's 128 columns are ready
                if (h == HALVES - 1) {
                  if (DBG) B200_TRACE(1, (int)it, 1);
                  ptx::umma_commit(empty0 + s * 8);  // sub-tile slot free once these MMAs retire
                  if (DBG) B200_TRACE(1, (int)it, 3);
                }
              }
            }
          }
          ptx::umma_commit(aempty);  // my MMAs on these A tiles are tracked
        }
      }
    }
  } else {
    // ===================== epilogue: threshold filter + candidate buffers =====================
    // Fast path per 128-column tile: four tcgen05.ld of 32 columns (software pipelined), an FMNMX3
    // tree per chunk, ONE warp vote per tile.  Only when some lane beats its threshold are the
    // offending chunks read again from TMEM and scanned.  The four warps of a half synchronise with
    // the MMA issuers on their own per-half barriers, so a warp that is compacting holds up three
    // others, not seven.
    const int half = warp >> 2, quad = warp & 3;
    const int row0 = half * TILE + quad * 32;
    const int row = row0 + lane;  // my query row inside the block
    const uint32_t t_lane = ((uint32_t)(quad * 32) << 16) + half * TILE;
    const int keep = p.keep;
    float* my_sc = bsc + row;
    int* my_id = bid + row;
    uint32_t it = 0;
    for (int b = 0; b < nqb; ++b) {
      const int qb = (int)blockIdx.x + b * (int)gridDim.x;
      float tau = CUDART_INF_F;
      int cnt = 0;
      for (int t = 0; t < p.ntiles; ++t, ++it) {
        const uint32_t a = it & 1, aph = (it >> 1) & 1;
        ptx::mbar_wait(hint, ptx::smem_u32(&bar_accf[a * 2 + half]), aph);
        if (warp == 0 && lane == 0) B200_TRACE(2, (int)it, 0);
        ptx::tc_fence_after();
        const uint32_t t_acc = t_lane + a * ACC_COLS;
        if ((p.mode & 3) != 2) {
          float va[32], vb[32], mc[4];
#define B200_CHUNKMIN(v, cc)                                                                       \
  do {                                                                                             \
    if (DBG) {                                                                                     \
      if (p.dbg) {                                                                                 \
        float* o = p.dbg + ((size_t)qb * ROWS + row) * ((size_t)p.ntiles * TILE) + (size_t)t * TILE + (cc) * 32; \
        _Pragma("unroll") for (int i = 0; i < 32; ++i) o[i] = v[i];                                \
      }                                                                                            \
    }                                                                                              \
    float g_[4];                                                                                   \
    _Pragma("unroll") for (int gi = 0; gi < 4; ++gi) {                                             \
      const float* w = v + gi * 8;                                                                 \
      g_[gi] = fminf(min3(w[0], w[1], w[2]), min3(min3(w[3], w[4], w[5]), w[6], w[7]));            \
    }                                                                                              \
    mc[cc] = fminf(fminf(g_[0], g_[1]), fminf(g_[2], g_[3]));                                      \
  } while (0)
          ptx::tmem_ld32(t_acc, va);
          ptx::tmem_ld_wait();
          ptx::tmem_ld32(t_acc + 32, vb);
          B200_CHUNKMIN(va, 0);
          ptx::tmem_ld_wait();
          ptx::tmem_ld32(t_acc + 64, va);
          B200_CHUNKMIN(vb, 1);
          ptx::tmem_ld_wait();
          ptx::tmem_ld32(t_acc + 96, vb);
          B200_CHUNKMIN(va, 2);
          ptx::tmem_ld_wait();
          B200_CHUNKMIN(vb, 3);
#undef B200_CHUNKMIN
          const float mt = fminf(fminf(mc[0], mc[1]), fminf(mc[2], mc[3]));
          // which 32-column chunks hold a survivor for some row of this warp: one warp-wide OR
          unsigned hitc = (mc[0] < tau ? 1u : 0u) | (mc[1] < tau ? 2u : 0u) | (mc[2] < tau ? 4u : 0u) | (mc[3] < tau ? 8u : 0u);
          hitc = (p.mode & 3) == 0 ? __reduce_or_sync(0xffffffffu, hitc) : 0u;
          if (p.mode & 128) hitc |= 1u;  // experiment: keep the slow-path code hot (chunk 0 rescanned every tile)
          // ---- slow path (uniform loop over the flagged chunks): read the chunk again, append every
          // value below the row's threshold with branch-free predicated stores
          while (hitc) {
            const int c = __ffs(hitc) - 1;
            hitc &= hitc - 1;
            ptx::tmem_ld32(t_acc + c * 32, va);
            ptx::tmem_ld_wait();
            const int base = t * TILE + c * 32;
            const float mcc = c == 0 ? mc[0] : (c == 1 ? mc[1] : (c == 2 ? mc[2] : mc[3]));
            B200_APPEND_CHUNK(va, base, mcc);
          }
        }
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&bar_acce[a * 2 + half]));
        if (warp == 0 && lane == 0) B200_TRACE(2, (int)it, 1);
      }
      // ---- final trim: a row left with more than 32 entries is cut to its keep best, so the re-rank
      // takes every query's candidates in one pass of its warp (one candidate row per lane)
      if (p.keep <= 32 && !(p.flags & 1024)) {
        const int keep = p.keep;
        unsigned need = __ballot_sync(0xffffffffu, cnt > 32);
        while (need) {
          const int r = __ffs(need) - 1;
          need &= need - 1;
          const int nv = __shfl_sync(0xffffffffu, cnt, r);
          const float nt = compact_row<CAP>(bsc, bid, RS, row0 + r, nv, keep, lane);
          if (lane == r) {
            tau = fminf(tau, nt);
            cnt = keep;
          }
        }
      }
      // ---- flush this block's candidate lists (coalesced, one row at a time)
      const int64_t q0 = (int64_t)qb * ROWS + row0;
      p.cand_cnt[q0 + lane] = cnt;
      p.cand_tau[q0 + lane] = tau;
      for (int r = 0; r < 32; ++r) {
        const int nv = __shfl_sync(0xffffffffu, cnt, r);
        for (int j = lane; j < nv; j += 32) p.cand_idx[(q0 + r) * CAP + j] = bid[j * RS + row0 + r];
      }
      __syncwarp();
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == EPI_WARPS + 1) ptx::tmem_dealloc(tmem_base, 512);
}

#include "knn_tc_ts.cuh"

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
    if (good && tau < CUDART_INF_F) {
      // scaled units: every non-candidate j has  d2_j * s^2 >= tau - E + ||x_q||^2
      const QMeta mq = meta[p];
      const double lb = (double)tau - mq.err + mq.qn2;
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

// Two stage-1 kernels.  knn_tc_kernel (SS form, default) takes 128- or 64-row sub-tile images in
// either layout.  knn_ts_kernel (TS form, tc_flags bit 6; 64-row chunk-major images) has the faster
// pipeline -- its distance + min-tree path alone runs at 0.97 of the measured cuBLAS peak against
// 0.86 -- but its epilogue warps are saturated, so survivor handling is not hidden and the whole
// kernel is ~8 % slower at C5 (profiles/r01/README.md); kept selectable for the next round.
inline bool use_ts(const b200_model* m) { return (m->tc_flags & 64) != 0; }

int tc_prepare(b200_model* m, cudaStream_t st) {
  TcData& t = m->tc;
  const int want_sub = use_ts(m) ? 64 : m->tc_sub, want_rg = use_ts(m) ? 0 : ((m->tc_flags >> 1) & 1);
  const int want_order = (m->tc_flags >> 7) & 1;
  if (t.ready && t.sub == want_sub && t.rg == want_rg && t.order == want_order) return 0;
  t.sub = want_sub;
  t.rg = want_rg;
  t.order = want_order;
  const int64_t n = m->n;
  const int d = m->d;
  t.K = tc::padK(d);
  t.tiles = (n + tc::TILE - 1) / tc::TILE;  // 128-column tiles = pairs of 64-row sub-tile images
  const int64_t npad = t.tiles * tc::TILE;
  B200_TRY(t.mu.reserve((size_t)d * 8 + 64));
  B200_TRY(t.img.reserve((size_t)npad * t.K * 2));
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
  // image position -> fitted row, by increasing centred norm (tc_flags bit 7: fitted order, for A/B runs)
  B200_TRY(t.perm.reserve((size_t)n * 4 + 64));
  const unsigned nb = (unsigned)((n + 255) / 256);
  if (m->tc_flags & 128) {
    tc::iota_kernel<<<nb, 256, 0, st>>>(t.perm.as<int32_t>(), n);
  } else {
    DevBuf keys, vals, tmp;
    B200_TRY(keys.reserve((size_t)n * 8 + 64));
    B200_TRY(vals.reserve((size_t)n * 4 + 64));
    uint32_t* k_in = keys.as<uint32_t>();
    uint32_t* k_out = k_in + n;
    tc::norm_key_kernel<<<nb, 256, 0, st>>>(m->x, n, d, t.mu.as<double>(), t.scale, k_in, vals.as<int32_t>());
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
  if ((size_t)n * d * 8 <= ((size_t)16 << 30)) {
    B200_TRY(t.xs.reserve((size_t)n * d * 8));
    const int64_t tot = n * d;
    tc::gather_rows_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(m->x, n, d, t.perm.as<int32_t>(), t.xs.as<double>());
    B200_CUDA(cudaGetLastError());
    m->st.launches += 1;
  }
  B200_CUDA(cudaMemsetAsync(bits, 0, 8, st));
  tc::data_img_kernel<<<(unsigned)((npad + 127) / 128), 128, 0, st>>>(m->x, n, d, t.K, t.sub, t.rg, t.mu.as<double>(), t.scale,
                                                                       t.perm.as<int32_t>(), t.img.as<__half>(), npad, bits);
  B200_CUDA(cudaGetLastError());
  B200_CUDA(cudaMemcpyAsync(hb, bits, 8, cudaMemcpyDeviceToHost, st));
  B200_CUDA(cudaStreamSynchronize(st));
  memcpy(&t.ymax, &hb[0], 8);
  m->st.launches += 3;
  t.ready = true;
  return 0;
}

struct TcConfig {
  int cap, halves, slots, keep, stages;
  size_t smem;
};

// Candidate capacity and CTA shape for k (+1 when the query's own row is dropped later):
// a compaction keeps `keep` = k + margin entries of a CAP-slot buffer; CAP leaves room for
// at least 16 appends between compactions; the B ring gets whatever shared memory is left.
bool tc_config(int d, int k, int margin, bool self, int subrows, int force_halves, TcConfig* c) {
  const int K = tc::padK(d);
  const int base = k + (self ? 1 : 0);
  int keep = base + (margin < 0 ? std::max(4, std::min(k / 2, 16)) : margin);
  const int caps[4] = {32, 48, 64, 128};
  int cap = 128;
  for (int c0 : caps)
    if (c0 >= keep + 20) {
      cap = c0;
      break;
    }
  if (keep > cap - 16) keep = cap - 16;
  if (keep < base) return false;
  const size_t tile = (size_t)tc::TILE * K * 2, sub = (size_t)subrows * K * 2;
  const size_t limit = 227 * 1024 - 512;
  for (int halves = 2; halves >= 1; --halves) {
    if (force_halves && halves != force_halves) continue;
    const size_t fixed = halves * tile + (size_t)cap * (halves * tc::TILE + 1) * 8 + 256;
    if (fixed + (force_halves ? 2 : 3) * sub > limit) continue;
    int slots = (int)std::min<size_t>(8, (limit - fixed) / sub);
    slots = slots >= 8 ? 8 : (slots >= 4 ? 4 : 2);  // power of two: ring index by mask/shift
    c->cap = cap;
    c->halves = halves;
    c->slots = slots;
    c->keep = keep;
    c->smem = fixed + slots * sub;
    return true;
  }
  return false;
}

// TS kernel: A lives in TMEM, so shared memory holds only the B ring (64-row sub-tiles) and the
// candidate buffers; TMEM holds `stages` 64-column accumulators per half next to A (K/2 columns).
bool tc_config_ts(int d, int k, int margin, bool self, int force_halves, TcConfig* c) {
  const int K = tc::padK(d);
  const int base = k + (self ? 1 : 0);
  int keep = base + (margin < 0 ? std::max(4, std::min(k / 2, 16)) : margin);
  const int caps[4] = {32, 48, 64, 128};
  int cap = 128;
  for (int c0 : caps)
    if (c0 >= keep + 20) {
      cap = c0;
      break;
    }
  if (keep > cap - 16) keep = cap - 16;
  if (keep < base) return false;
  const size_t sub = (size_t)64 * K * 2;
  const size_t limit = 227 * 1024 - 512;
  for (int halves = 2; halves >= 1; --halves) {
    if (force_halves && halves != force_halves) continue;
    const int stages = std::min(6, (512 / halves - K / 2) / 64);
    if (stages < 2) continue;
    const size_t fixed = (size_t)cap * (halves * tc::TILE + 1) * 8 + 1024;
    if (fixed + 4 * sub > limit) continue;
    int slots = (int)std::min<size_t>(16, (limit - fixed) / sub);
    slots = slots >= 16 ? 16 : (slots >= 8 ? 8 : 4);
    c->cap = cap;
    c->halves = halves;
    c->slots = slots;
    c->keep = keep;
    c->stages = stages;
    c->smem = fixed + slots * sub;
    return true;
  }
  return false;
}

template <int CAP, int HALVES>
int launch_ts(b200_model* m, const tc::KParams& kp, const TcConfig& cfg, bool dbg, cudaStream_t st) {
  const int threads = (HALVES * 6 + 1) * 32;  // 4 epilogue + 2 issuer warps per half, one producer
  const int grid = std::min(kp.qblocks, m->sm_count);
  if (dbg) {
    auto kern = tc::knn_ts_kernel<CAP, HALVES, true>;
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg.smem));
    kern<<<grid, threads, cfg.smem, st>>>(kp);
  } else {
    auto kern = tc::knn_ts_kernel<CAP, HALVES, false>;
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg.smem));
    kern<<<grid, threads, cfg.smem, st>>>(kp);
  }
  B200_CUDA(cudaGetLastError());
  return 0;
}

template <int CAP, int HALVES>
int launch_tc(b200_model* m, const tc::KParams& kp, const TcConfig& cfg, bool dbg, cudaStream_t st) {
  const int threads = (HALVES * 4 + 3) * 32;
  const int grid = std::min(kp.qblocks, m->sm_count);
  if (dbg) {
    auto kern = tc::knn_tc_kernel<CAP, HALVES, true>;
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg.smem));
    kern<<<grid, threads, cfg.smem, st>>>(kp);
  } else {
    auto kern = tc::knn_tc_kernel<CAP, HALVES, false>;
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg.smem));
    kern<<<grid, threads, cfg.smem, st>>>(kp);
  }
  B200_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace

bool tc_supported(const b200_model* m, int k) {
  TcConfig c;
  if (m->d + tc::NORM_SLOTS > 256) return false;
  return use_ts(m) ? tc_config_ts(m->d, k, 0, true, 0, &c) : tc_config(m->d, k, 0, true, 128, 0, &c);
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
  const bool ts = use_ts(m);
  cfg.stages = 0;
  if (!(ts ? tc_config_ts(m->d, k, m->margin, self_offset >= 0, m->tc_halves, &cfg)
           : tc_config(m->d, k, m->margin, self_offset >= 0, m->tc.sub, m->tc_halves, &cfg))) {
    set_error("tensor kNN path unsupported for d=%d k=%d", m->d, k);
    return 1;
  }
  // Wave quantisation: the persistent grid takes sm_count 256-query blocks per round.  When the last
  // round would keep at most half the SMs busy, those queries go to a second launch with 128-query
  // blocks (twice as many, one round, ~0.64 of a full round's time).  8 GPUs x 125 k queries: 3.64
  // instead of 4 rounds.
  if (cfg.halves == 2 && m->tc_halves == 0 && m->tc_split == 0 && !ts && m->tc_mode == 0 && !m->dbg_scores && !m->dbg_trace &&
      !(m->tc_flags & 8192)) {
    const int64_t blocks = (q + 2 * tc::TILE - 1) / (2 * tc::TILE);
    const int64_t rounds = blocks / m->sm_count, tail = blocks - rounds * m->sm_count;
    TcConfig c1;
    if (rounds >= 1 && tail > 0 && 2 * tail <= m->sm_count &&
        tc_config(m->d, k, m->margin, self_offset >= 0, m->tc.sub, 1, &c1)) {
      const int64_t q_main = rounds * m->sm_count * 2 * tc::TILE;
      B200_TRY(knn_tc(m, z_dev, q_main, k, self_offset, idx_dev, d2_dev, st));
      // the tail: either 128-query blocks (tc_flags bit 15), or -- default -- the same 256-query blocks with
      // the data stream split over S concurrent launches on the otherwise idle SMs, S candidate lists per
      // query merged by the re-rank
      const int split = (int)std::min<int64_t>(4, m->sm_count / tail);
      const bool use_split = split >= 2 && cfg.keep <= 32 && t_tiles_ok(m, split) && !(m->tc_flags & 32768);
      if (use_split) m->tc_split = split;
      else m->tc_halves = 1;
      m->defer_base += q_main;
      const int rc = knn_tc(m, z_dev + (size_t)q_main * m->d, q - q_main, k,
                            self_offset >= 0 ? self_offset + q_main : -1, idx_dev + (size_t)q_main * k,
                            d2_dev + (size_t)q_main * k, st);
      m->defer_base -= q_main;
      m->tc_halves = 0;
      m->tc_split = 0;
      m->st.last_halves = 2;
      return rc;
    }
  }
  const TcData& t = m->tc;
  const int rows = cfg.halves * tc::TILE;
  const int qblocks = (int)((q + rows - 1) / rows);
  const int64_t qpad = (int64_t)qblocks * rows;
  const int S = m->tc_split > 1 ? m->tc_split : 1;  // concurrent launches over sub-ranges of the data stream
  B200_TRY(m->zimg.reserve((size_t)qpad * t.K * 2));
  B200_TRY(m->cand.reserve((size_t)S * qpad * cfg.cap * 4));
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
  tc::query_img_kernel<<<(unsigned)((qpad + 127) / 128), 128, 0, st>>>(z_dev, q, m->d, t.K, t.rg, t.mu.as<double>(), t.scale,
                                                                        t.ymax, m->zimg.as<__half>(), qpad, meta);
  B200_CUDA(cudaGetLastError());
  cudaEventRecord(m->ev0, st);
  tc::KParams kp;
  kp.a_img = m->zimg.as<__half>();
  kp.b_img = t.img.as<__half>();
  kp.K = t.K;
  kp.ntiles = (int)t.tiles;
  kp.qblocks = qblocks;
  kp.slots = cfg.slots;
  kp.stages = cfg.stages;
  kp.sub = t.sub;
  kp.flags = m->tc_flags;
  kp.keep = cfg.keep;
  kp.mode = m->tc_mode;
  kp.cand_idx = m->cand.as<int32_t>();
  kp.cand_cnt = cand_cnt;
  kp.cand_tau = cand_tau;
  kp.dbg = dbg;
  kp.trace = m->dbg_trace;
  kp.trace_start = m->tc_trace_start;
  const bool dg = dbg != nullptr || kp.trace != nullptr || (m->tc_mode & 12);  // trace and MMA/TMA knock-outs live in the debug build
  auto launch_one = [&](const tc::KParams& kq, cudaStream_t s_) -> int {
    int rc = 1;
    if (ts) {
      if (cfg.cap == 32 && cfg.halves == 2) rc = launch_ts<32, 2>(m, kq, cfg, dg, s_);
      else if (cfg.cap == 48 && cfg.halves == 2) rc = launch_ts<48, 2>(m, kq, cfg, dg, s_);
      else if (cfg.cap == 64 && cfg.halves == 2) rc = launch_ts<64, 2>(m, kq, cfg, dg, s_);
      else if (cfg.cap == 128 && cfg.halves == 2) rc = launch_ts<128, 2>(m, kq, cfg, dg, s_);
      else if (cfg.cap == 32 && cfg.halves == 1) rc = launch_ts<32, 1>(m, kq, cfg, dg, s_);
      else if (cfg.cap == 48 && cfg.halves == 1) rc = launch_ts<48, 1>(m, kq, cfg, dg, s_);
      else if (cfg.cap == 64 && cfg.halves == 1) rc = launch_ts<64, 1>(m, kq, cfg, dg, s_);
      else if (cfg.cap == 128 && cfg.halves == 1) rc = launch_ts<128, 1>(m, kq, cfg, dg, s_);
      else set_error("tensor kNN: no kernel for cap=%d halves=%d", cfg.cap, cfg.halves);
    } else if (cfg.cap == 32 && cfg.halves == 2) rc = launch_tc<32, 2>(m, kq, cfg, dg, s_);
    else if (cfg.cap == 48 && cfg.halves == 2) rc = launch_tc<48, 2>(m, kq, cfg, dg, s_);
    else if (cfg.cap == 48 && cfg.halves == 1) rc = launch_tc<48, 1>(m, kq, cfg, dg, s_);
    else if (cfg.cap == 64 && cfg.halves == 2) rc = launch_tc<64, 2>(m, kq, cfg, dg, s_);
    else if (cfg.cap == 128 && cfg.halves == 1) rc = launch_tc<128, 1>(m, kq, cfg, dg, s_);
    else if (cfg.cap == 32 && cfg.halves == 1) rc = launch_tc<32, 1>(m, kq, cfg, dg, s_);
    else if (cfg.cap == 64 && cfg.halves == 1) rc = launch_tc<64, 1>(m, kq, cfg, dg, s_);
    else set_error("tensor kNN: no kernel for cap=%d halves=%d", cfg.cap, cfg.halves);

    return rc;
  };
  int rc = 1;
  tc::RerankLists L;
  L.nlists = S;
  L.list_elems = (int64_t)qpad * cfg.cap;
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
      kq.ntiles = (int)(hi - lo);
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
  m->st.last_cap = cfg.cap;
  m->st.last_keep = cfg.keep;
  m->st.last_slots = cfg.slots;
  m->st.last_halves = cfg.halves;
  if (m->tc_mode != 0) {  // debug ceilings: outputs are meaningless, only stage 1 is timed
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
  tc::rerank_kernel<CAPV><<<rb, 256, (size_t)8 * ((m->d + 1) & ~1) * 8, st>>>(m->x, m->n, m->d, z_dev, q, k, self_offset, kp.cand_idx, cfg.cap, \
                                              cand_cnt, cand_tau, meta, t.scale, t.cmax, t.perm.as<int32_t>(),                  \
                                              (m->tc_flags & 16384) ? nullptr : t.xs.as<double>(), idx_dev, d2_dev, \
                                              fail_list, fail_cnt, L)
  // the SS kernel's final trim leaves at most 32 candidates per query when keep <= 32
  const bool one_pass = !ts && cfg.keep <= 32 && !(m->tc_flags & (1024 | 2048));
  if (S > 1) RERANK(128);  // pass r of the 128-slot instance takes list r
  else if (cfg.cap == 32 || one_pass) RERANK(32);
  else if (cfg.cap <= 64) RERANK(64);
  else RERANK(128);
#undef RERANK
  B200_CUDA(cudaGetLastError());
  cudaEventRecord(m->ev2, st);
  int h_fail = 0;
  B200_CUDA(cudaMemcpyAsync(&h_fail, fail_cnt, 4, cudaMemcpyDeviceToHost, st));
  B200_CUDA(cudaStreamSynchronize(st));
  float ms1 = 0, ms2 = 0;
  cudaEventElapsedTime(&ms1, m->ev0, m->ev1);
  cudaEventElapsedTime(&ms2, m->ev1, m->ev2);
  float ms0 = 0;
  cudaEventElapsedTime(&ms0, m->ev3, m->ev0);
  m->st.ms_prep += ms0;
  m->st.ms_stage1 += ms1;
  m->st.ms_rerank += ms2;
  m->st.launches += 3 + (S - 1);  // query image, S stage-1 launches, re-rank
  m->st.tc_queries += q;
  m->st.uncertified += h_fail;
  m->st.last_cap = cfg.cap;
  m->st.last_keep = cfg.keep;
  m->st.last_slots = cfg.slots;
  if (h_fail > 0 && m->defer_fallback && h_fail <= 64 && m->deferred.size() + h_fail <= 512) {
    // the caller resolves these after its last chunk (one pass over the fitted points instead of one per chunk)
    std::vector<int32_t> loc(h_fail);
    B200_CUDA(cudaMemcpy(loc.data(), fail_list, (size_t)h_fail * 4, cudaMemcpyDeviceToHost));
    for (int32_t v : loc) m->deferred.push_back((int32_t)(v + m->defer_base));
    return 0;
  }
  if (h_fail > 0) {
    cudaEventRecord(m->ev0, st);
    if (m->tc_flags & 4096) B200_TRY(knn_exact(m, z_dev, q, k, self_offset, idx_dev, d2_dev, fail_list, h_fail, st));
    else B200_TRY(knn_rescan(m, z_dev, q, k, self_offset, idx_dev, d2_dev, fail_list, h_fail, st));
    cudaEventRecord(m->ev1, st);
    B200_CUDA(cudaEventSynchronize(m->ev1));
    cudaEventElapsedTime(&ms1, m->ev0, m->ev1);
    m->st.ms_exact += ms1;
  }
  return 0;
}

}  // namespace b200
