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
//  3. Queries that fail the certificate are re-run by the exact kernel
//     (knn_exact.cu) -- so the result is always the exact answer.
//
// Operand layout in HBM: per 128-row tile a ready-to-use UMMA image
// [K/8][128][8 halfs] (K-major, no swizzle: 8x16-byte core matrices contiguous),
// so one cp.async.bulk (TMA engine) brings a whole tile and the smem descriptor
// is fixed: LBO = 2048 B (next K chunk), SBO = 128 B (next 8 rows).
#include "common.cuh"
#include "ptx.cuh"

namespace b200 {

namespace tc {

constexpr int TILE = 128;          // query rows per accumulator half = data columns per epilogue tile
constexpr int SUBROWS = 64;        // data rows per operand sub-tile image (one bulk copy, one UMMA N)
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

// ------------------------------------------------------------- operand images
// Data side: one thread per (padded) point; writes its row of the tile image and
// folds max(||yc||, ||fp16(yc)||) into ymax.
__global__ void data_img_kernel(const double* __restrict__ x, int64_t n, int d, int K,
                                const double* __restrict__ mu, double scale, __half* __restrict__ img,
                                int64_t npad, unsigned long long* __restrict__ ymax_bits) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  double nm = 0.0;
  if (i < npad) {
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
          const double yc = (x[i * d + j] - mu[j]) * scale;
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
      dst[(size_t)c * SUBROWS + r] = *reinterpret_cast<uint4*>(h);
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
__global__ void query_img_kernel(const double* __restrict__ z, int64_t q, int d, int K,
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
    dst[(size_t)c * TILE + r] = *reinterpret_cast<uint4*>(h);
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
  int slots;             // depth of the B ring, in 64-row sub-tiles
  int keep;              // C: candidates a compaction keeps (<= CAP - 8)
  int mode;              // 0 normal; debug ceilings: 1 = epilogue reads TMEM only, 2 = epilogue only releases
  int32_t* cand_idx;     // [qpad][CAP]
  int32_t* cand_cnt;     // [qpad]
  float* cand_tau;       // [qpad]
  float* dbg;            // optional [qpad][ntiles*128] raw scores (tests only)
};

constexpr int SUB = 64;  // data rows per B sub-tile (one TMA copy, one UMMA N)

__device__ __forceinline__ float min3(float a, float b, float c) { return fminf(fminf(a, b), c); }

// Trim buffer column `col` (one query row) to its `keep` best entries; returns the new
// threshold (the keep-th smallest score).  Warp-cooperative, all lanes must call.
// bsc/bid are [CAP][RS] (slot-major, RS = rows + 1 keeps slot walks conflict-free).
template <int CAP>
__device__ __noinline__ float compact_row(float* bsc, int* bid, int RS, int col, int n_valid, int keep,
                                          int lane) {
  constexpr int U = CAP / 32;
  float sc[U];
  int id[U], rank[U], eq[U];
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const int j = lane + 32 * u;
    sc[u] = j < n_valid ? bsc[j * RS + col] : CUDART_INF_F;
    id[u] = bid[j * RS + col];
    rank[u] = 0;
    eq[u] = 0;
  }
  const float* ps = bsc + col;
  int i = 0;
  for (; i + 4 <= n_valid; i += 4) {
    const float s0 = ps[(i + 0) * RS], s1 = ps[(i + 1) * RS], s2 = ps[(i + 2) * RS], s3 = ps[(i + 3) * RS];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      rank[u] += (s0 < sc[u]) + (s1 < sc[u]) + (s2 < sc[u]) + (s3 < sc[u]);
      eq[u] += (s0 == sc[u]) + (s1 == sc[u]) + (s2 == sc[u]) + (s3 == sc[u]);
    }
  }
  for (; i < n_valid; ++i) {
    const float s0 = ps[i * RS];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      rank[u] += (s0 < sc[u]);
      eq[u] += (s0 == sc[u]);
    }
  }
  bool tie = false;
#pragma unroll
  for (int u = 0; u < U; ++u) tie |= (lane + 32 * u < n_valid) && eq[u] > 1;
  if (__any_sync(0xffffffffu, tie)) {  // equal scores: redo with the slot number as tie-break
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

template <int CAP, int HALVES, bool DBG>
__global__ void __launch_bounds__((HALVES * 4 + 2) * 32, 1) knn_tc_kernel(const KParams p) {
  constexpr int ROWS = HALVES * TILE;
  constexpr int RS = ROWS + 1;
  constexpr int EPI_WARPS = HALVES * 4;
  constexpr uint32_t ACC_COLS = HALVES * TILE;  // TMEM columns per accumulator stage
  extern __shared__ __align__(1024) uint8_t smem[];
  const int K = p.K;
  const uint32_t tile_bytes = (uint32_t)TILE * K * 2;  // one 128-row query tile
  const uint32_t sub_bytes = (uint32_t)SUB * K * 2;    // one 64-row data sub-tile
  uint8_t* sA = smem;                                  // HALVES query tiles
  uint8_t* sB = sA + (size_t)HALVES * tile_bytes;      // ring of p.slots sub-tiles
  float* bsc = reinterpret_cast<float*>(sB + (size_t)p.slots * sub_bytes);  // [CAP][RS] scores
  int* bid = reinterpret_cast<int*>(bsc + (size_t)CAP * RS);                // [CAP][RS] data indices
  uint64_t* bars = reinterpret_cast<uint64_t*>((reinterpret_cast<uintptr_t>(bid + (size_t)CAP * RS) + 15) & ~(uintptr_t)15);
  // bars: full[8], empty[8], accf[2], acce[2], afull, aempty
  uint64_t* bar_full = bars;
  uint64_t* bar_empty = bars + 8;
  uint64_t* bar_accf = bars + 16;
  uint64_t* bar_acce = bars + 18;
  uint64_t* bar_afull = bars + 20;
  uint64_t* bar_aempty = bars + 21;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 22);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.slots; ++s) {
      ptx::mbar_init(ptx::smem_u32(&bar_full[s]), 1);
      ptx::mbar_init(ptx::smem_u32(&bar_empty[s]), 1);
    }
    for (int a = 0; a < 2; ++a) {
      ptx::mbar_init(ptx::smem_u32(&bar_accf[a]), 1);
      ptx::mbar_init(ptx::smem_u32(&bar_acce[a]), EPI_WARPS);
    }
    ptx::mbar_init(ptx::smem_u32(bar_afull), 1);
    ptx::mbar_init(ptx::smem_u32(bar_aempty), 1);
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

  const int nqb = (p.qblocks - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;  // my query blocks
  const int nsub = 2 * p.ntiles;

  if (warp == EPI_WARPS) {
    // ===================== producer: bulk copies on the TMA engine =====================
    if (lane == 0) {
      uint32_t jt = 0;
      for (int b = 0; b < nqb; ++b) {
        const int qb = (int)blockIdx.x + b * (int)gridDim.x;
        ptx::mbar_wait(ptx::smem_u32(bar_aempty), (b & 1) ^ 1);
        ptx::mbar_arrive_expect_tx(ptx::smem_u32(bar_afull), HALVES * tile_bytes);
        ptx::bulk_g2s(ptx::smem_u32(sA), reinterpret_cast<const uint8_t*>(p.a_img) + (size_t)qb * HALVES * tile_bytes,
                      HALVES * tile_bytes, ptx::smem_u32(bar_afull));
        for (int st = 0; st < nsub; ++st, ++jt) {
          const uint32_t s = jt % p.slots, ph = (jt / p.slots) & 1;
          ptx::mbar_wait(ptx::smem_u32(&bar_empty[s]), ph ^ 1);
          ptx::mbar_arrive_expect_tx(ptx::smem_u32(&bar_full[s]), sub_bytes);
          ptx::bulk_g2s(ptx::smem_u32(sB + (size_t)s * sub_bytes),
                        reinterpret_cast<const uint8_t*>(p.b_img) + (size_t)st * sub_bytes, sub_bytes,
                        ptx::smem_u32(&bar_full[s]));
        }
      }
    }
  } else if (warp == EPI_WARPS + 1) {
    // ===================== MMA issuer: one thread drives the tensor core =====================
    if (lane == 0) {
      constexpr uint32_t idesc = ptx::umma_idesc_f16(TILE, SUB);
      const int ksteps = K / 16;
      uint32_t it = 0, jt = 0;
      for (int b = 0; b < nqb; ++b) {
        ptx::mbar_wait(ptx::smem_u32(bar_afull), b & 1);
        for (int t = 0; t < p.ntiles; ++t, ++it) {
          const uint32_t a = it & 1, aph = (it >> 1) & 1;
          ptx::mbar_wait(ptx::smem_u32(&bar_acce[a]), aph ^ 1);
#pragma unroll
          for (int hs = 0; hs < 2; ++hs, ++jt) {
            const uint32_t s = jt % p.slots, ph = (jt / p.slots) & 1;
            ptx::mbar_wait(ptx::smem_u32(&bar_full[s]), ph);
            ptx::tc_fence_after();
            const uint32_t b_addr = ptx::smem_u32(sB + (size_t)s * sub_bytes);
#pragma unroll
            for (int h = 0; h < HALVES; ++h) {
              const uint32_t a_addr = ptx::smem_u32(sA + (size_t)h * tile_bytes);
              const uint32_t d_tmem = tmem_base + a * ACC_COLS + h * TILE + hs * SUB;
              for (int ks = 0; ks < ksteps; ++ks) {
                const uint64_t ad = ptx::umma_desc(a_addr + ks * 4096, 2048, 128);
                const uint64_t bd = ptx::umma_desc(b_addr + ks * 2048, 1024, 128);
                ptx::umma_f16(d_tmem, ad, bd, idesc, ks > 0 ? 1u : 0u);
              }
            }
            ptx::umma_commit(ptx::smem_u32(&bar_empty[s]));  // sub-tile slot free once these MMAs retire
          }
          ptx::umma_commit(ptx::smem_u32(&bar_accf[a]));     // 128 columns ready for the epilogue
        }
        ptx::umma_commit(ptx::smem_u32(bar_aempty));         // A tiles may be replaced
      }
    }
  } else {
    // ===================== epilogue: threshold filter + candidate buffers =====================
    const int half = warp >> 2, quad = warp & 3;
    const int row0 = half * TILE + quad * 32;
    const int row = row0 + lane;  // my query row inside the block
    const uint32_t t_lane = tmem_base + ((uint32_t)(quad * 32) << 16) + half * TILE;
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
        ptx::mbar_wait(ptx::smem_u32(&bar_accf[a]), aph);
        ptx::tc_fence_after();
        if (p.mode != 2) {
          const uint32_t t_acc = t_lane + a * ACC_COLS;
          float va[32], vb[32];
          ptx::tmem_ld32(t_acc, va);
#pragma unroll 1
          for (int c = 0; c < TILE / 32; c += 2) {
            // software pipeline: the next 32 columns are in flight while these are filtered
#define B200_PROCESS(v, cc)                                                                        \
  do {                                                                                             \
    if (DBG) {                                                                                     \
      if (p.dbg) {                                                                                 \
        float* o = p.dbg + ((size_t)qb * ROWS + row) * ((size_t)p.ntiles * TILE) + (size_t)t * TILE + (cc) * 32; \
        _Pragma("unroll") for (int i = 0; i < 32; ++i) o[i] = v[i];                                \
      }                                                                                            \
    }                                                                                              \
    float g[4];                                                                                    \
    _Pragma("unroll") for (int gi = 0; gi < 4; ++gi) {                                             \
      const float* w = v + gi * 8;                                                                 \
      g[gi] = fminf(min3(w[0], w[1], w[2]), min3(min3(w[3], w[4], w[5]), w[6], w[7]));             \
    }                                                                                              \
    const float m = fminf(fminf(g[0], g[1]), fminf(g[2], g[3]));                                   \
    if (p.mode == 0 && __any_sync(0xffffffffu, m < tau)) {                                         \
      const int base = t * TILE + (cc) * 32;                                                       \
      _Pragma("unroll") for (int gi = 0; gi < 4; ++gi) {                                           \
        if (g[gi] < tau) {                                                                         \
          _Pragma("unroll") for (int i = 0; i < 8; ++i) {                                          \
            const float s = v[gi * 8 + i];                                                         \
            if (s < tau) {                                                                         \
              my_sc[cnt * RS] = s;                                                                 \
              my_id[cnt * RS] = base + gi * 8 + i;                                                 \
              ++cnt;                                                                               \
            }                                                                                      \
          }                                                                                        \
        }                                                                                          \
        unsigned need = __ballot_sync(0xffffffffu, cnt > CAP - 8);                                 \
        while (need) {                                                                             \
          const int r = __ffs(need) - 1;                                                           \
          need &= need - 1;                                                                        \
          const int nv = __shfl_sync(0xffffffffu, cnt, r);                                         \
          const float nt = compact_row<CAP>(bsc, bid, RS, row0 + r, nv, keep, lane);               \
          if (lane == r) {                                                                         \
            tau = nt;                                                                              \
            cnt = keep;                                                                            \
          }                                                                                        \
        }                                                                                          \
      }                                                                                            \
    }                                                                                              \
  } while (0)
            ptx::tmem_ld_wait();
            ptx::tmem_ld32(t_acc + (c + 1) * 32, vb);
            B200_PROCESS(va, c);
            ptx::tmem_ld_wait();
            if (c + 2 < TILE / 32) ptx::tmem_ld32(t_acc + (c + 2) * 32, va);
            B200_PROCESS(vb, c + 1);
#undef B200_PROCESS
          }
        }
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&bar_acce[a]));
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

// ------------------------------------------------------------------ fp64 re-rank
// One warp per query: exact squared distances of its candidates (reference
// arithmetic), rank by (d2, index), emit the k best, test the certificate.
template <int CAP>
__global__ void __launch_bounds__(256)
rerank_kernel(const double* __restrict__ x, int64_t n, int d, const double* __restrict__ z, int64_t q,
              int k, int64_t self_offset, const int32_t* __restrict__ cand_idx,
              const int32_t* __restrict__ cand_cnt, const float* __restrict__ cand_tau,
              const QMeta* __restrict__ meta, double scale, double data_cmax,
              int32_t* __restrict__ out_idx, double* __restrict__ out_d2, int32_t* __restrict__ fail_list,
              int32_t* __restrict__ fail_cnt) {
  constexpr int U = CAP / 32;
  __shared__ double s_d2[8][CAP];
  __shared__ int s_id[8][CAP];
  __shared__ double s_dk[8];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t p = (int64_t)blockIdx.x * 8 + w;
  if (p >= q) return;
  const int cnt = min(cand_cnt[p], CAP);
  const int self = self_offset >= 0 ? (int)(self_offset + p) : -1;
  const double* zp = z + p * d;
  double dd[U];
  int id[U];
  int nvalid = 0;
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const int j = lane + 32 * u;
    int c = j < cnt ? cand_idx[p * CAP + j] : -1;
    const bool ok = c >= 0 && c < n && c != self;
    dd[u] = ok ? sqdist_seq(x + (int64_t)c * d, zp, d) : CUDART_INF;
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
    const float tau = cand_tau[p];
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

int tc_prepare(b200_model* m, cudaStream_t st) {
  TcData& t = m->tc;
  if (t.ready) return 0;
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
  B200_CUDA(cudaMemsetAsync(bits, 0, 8, st));
  tc::data_img_kernel<<<(unsigned)((npad + 127) / 128), 128, 0, st>>>(m->x, n, d, t.K, t.mu.as<double>(), t.scale,
                                                                       t.img.as<__half>(), npad, bits);
  B200_CUDA(cudaGetLastError());
  B200_CUDA(cudaMemcpyAsync(hb, bits, 8, cudaMemcpyDeviceToHost, st));
  B200_CUDA(cudaStreamSynchronize(st));
  memcpy(&t.ymax, &hb[0], 8);
  m->st.launches += 3;
  t.ready = true;
  return 0;
}

struct TcConfig {
  int cap, halves, slots, keep;
  size_t smem;
};

// Candidate capacity and CTA shape for k (+1 when the query's own row is dropped later):
// a compaction keeps `keep` = k + margin entries of a CAP-slot buffer; CAP leaves room for
// at least 16 appends between compactions; the B ring gets whatever shared memory is left.
bool tc_config(int d, int k, int margin, bool self, TcConfig* c) {
  const int K = tc::padK(d);
  const int base = k + (self ? 1 : 0);
  int keep = base + (margin < 0 ? std::min(k, 16) : margin);
  int cap = 32;
  while (cap < keep + 24 && cap < 128) cap *= 2;
  if (keep > cap - 24) keep = cap - 24;
  if (keep < base) return false;
  const size_t tile = (size_t)tc::TILE * K * 2, sub = (size_t)tc::SUBROWS * K * 2;
  const size_t limit = 227 * 1024 - 512;
  for (int halves = 2; halves >= 1; --halves) {
    const size_t fixed = halves * tile + (size_t)cap * (halves * tc::TILE + 1) * 8 + 256;
    if (fixed + 3 * sub > limit) continue;
    const int slots = (int)std::min<size_t>(8, (limit - fixed) / sub);
    c->cap = cap;
    c->halves = halves;
    c->slots = slots;
    c->keep = keep;
    c->smem = fixed + slots * sub;
    return true;
  }
  return false;
}

template <int CAP, int HALVES>
int launch_tc(b200_model* m, const tc::KParams& kp, const TcConfig& cfg, bool dbg, cudaStream_t st) {
  const int threads = (HALVES * 4 + 2) * 32;
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
  return m->d + tc::NORM_SLOTS <= 256 && tc_config(m->d, k, 0, true, &c);
}

int knn_tc(b200_model* m, const double* z_dev, int64_t q, int k, int64_t self_offset,
           int32_t* idx_dev, double* d2_dev, cudaStream_t st) {
  if (q == 0) return 0;
  if (k < 1 || k > m->n - (self_offset >= 0 ? 1 : 0)) {
    set_error("knn: k=%d out of range for n=%lld", k, (long long)m->n);
    return 1;
  }
  B200_TRY(tc_prepare(m, st));
  TcConfig cfg;
  if (!tc_config(m->d, k, m->margin, self_offset >= 0, &cfg)) {
    set_error("tensor kNN path unsupported for d=%d k=%d", m->d, k);
    return 1;
  }
  const TcData& t = m->tc;
  const int rows = cfg.halves * tc::TILE;
  const int qblocks = (int)((q + rows - 1) / rows);
  const int64_t qpad = (int64_t)qblocks * rows;
  B200_TRY(m->zimg.reserve((size_t)qpad * t.K * 2));
  B200_TRY(m->cand.reserve((size_t)qpad * cfg.cap * 4));
  B200_TRY(m->cand_meta.reserve((size_t)qpad * (8 + sizeof(tc::QMeta)) + 64));
  B200_TRY(m->flags.reserve(256));
  B200_TRY(m->misc.reserve((size_t)q * 4 + 64));
  int32_t* cand_cnt = m->cand_meta.as<int32_t>();
  float* cand_tau = reinterpret_cast<float*>(cand_cnt + qpad);
  tc::QMeta* meta = reinterpret_cast<tc::QMeta*>(cand_tau + qpad);
  int32_t* fail_cnt = m->flags.as<int32_t>() + 8;
  int32_t* fail_list = m->misc.as<int32_t>();
  float* dbg = m->dbg_scores;

  cudaEventRecord(m->ev3, st);
  tc::query_img_kernel<<<(unsigned)((qpad + 127) / 128), 128, 0, st>>>(z_dev, q, m->d, t.K, t.mu.as<double>(), t.scale,
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
  kp.keep = cfg.keep;
  kp.mode = m->tc_mode;
  kp.cand_idx = m->cand.as<int32_t>();
  kp.cand_cnt = cand_cnt;
  kp.cand_tau = cand_tau;
  kp.dbg = dbg;
  int rc = 1;
  const bool dg = dbg != nullptr;
  if (cfg.cap == 32 && cfg.halves == 2) rc = launch_tc<32, 2>(m, kp, cfg, dg, st);
  else if (cfg.cap == 64 && cfg.halves == 2) rc = launch_tc<64, 2>(m, kp, cfg, dg, st);
  else if (cfg.cap == 128 && cfg.halves == 1) rc = launch_tc<128, 1>(m, kp, cfg, dg, st);
  else if (cfg.cap == 32 && cfg.halves == 1) rc = launch_tc<32, 1>(m, kp, cfg, dg, st);
  else if (cfg.cap == 64 && cfg.halves == 1) rc = launch_tc<64, 1>(m, kp, cfg, dg, st);
  else set_error("tensor kNN: no kernel for cap=%d halves=%d", cfg.cap, cfg.halves);
  B200_TRY(rc);
  cudaEventRecord(m->ev1, st);
  B200_CUDA(cudaMemsetAsync(fail_cnt, 0, 4, st));
  const unsigned rb = (unsigned)((q + 7) / 8);
#define RERANK(CAPV)                                                                                     \
  tc::rerank_kernel<CAPV><<<rb, 256, 0, st>>>(m->x, m->n, m->d, z_dev, q, k, self_offset, kp.cand_idx,   \
                                              cand_cnt, cand_tau, meta, t.scale, t.cmax, idx_dev, d2_dev, \
                                              fail_list, fail_cnt)
  if (cfg.cap == 32) RERANK(32);
  else if (cfg.cap == 64) RERANK(64);
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
  m->st.launches += 3;
  m->st.tc_queries += q;
  m->st.uncertified += h_fail;
  m->st.last_cap = cfg.cap;
  m->st.last_keep = cfg.keep;
  m->st.last_slots = cfg.slots;
  if (h_fail > 0) {
    cudaEventRecord(m->ev0, st);
    B200_TRY(knn_exact(m, z_dev, q, k, self_offset, idx_dev, d2_dev, fail_list, h_fail, st));
    cudaEventRecord(m->ev1, st);
    B200_CUDA(cudaEventSynchronize(m->ev1));
    cudaEventElapsedTime(&ms1, m->ev0, m->ev1);
    m->st.ms_exact += ms1;
  }
  return 0;
}

}  // namespace b200
