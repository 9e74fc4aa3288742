// knn_exact.cu -- exact fp64 brute-force kNN on CUDA cores.
//
// Replaces sklearn BallTree.query as NearestNeighbor calls it
// (nearest_neighbor/__init__.py:75-76) and the all-pairs scans of SHEPMOD / LSHEP
// (modified_shepard.f95:45-55, linear_shepard.f95:105-115).  Squared distances
// are formed exactly as the reference does -- direct differences summed over
// ascending dimension with separate multiply and add roundings
// (sklearn metrics/_dist_metrics.pxd rdist) -- so (d2, index) ordering and the
// returned d2 are bit-identical.  This is the small-problem path, the fallback
// for queries the tensor path cannot certify, and the scan used by the fits.
//
// Layout: one thread owns one query (its running top-k in local memory, its
// coordinates in a transposed shared tile), data rows stream through shared
// memory in tiles and are broadcast-read by the whole block.  The data range is
// split across blockIdx.y so small query batches still fill 148 SMs; partial
// lists land in scratch and are merged by merge_topk_kernel.
#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"

namespace b200 {

namespace {

template <int KMAX>
__device__ __forceinline__ void topk_insert(double* bd, int* bi, int k, double d2, int gi) {
  int j = k - 1;
  while (j > 0 && bd[j - 1] > d2) {  // strict: equal distances keep the earlier (lower) index
    bd[j] = bd[j - 1];
    bi[j] = bi[j - 1];
    --j;
  }
  bd[j] = d2;
  bi[j] = gi;
}

// PRE: an fp32 prefilter (common.cuh) in front of the fp64 arithmetic.  Four points at a time get
// fp32 squared distances from float copies of the tile and of the queries; the group is skipped when
// all four are certainly no nearer than the current k-th best (rigorous rounding bound, so the result
// is unchanged bit for bit); otherwise it takes the fp64 path below.  Needs the model's padded float
// image x32 ([n][dsf], dsf = 4*ceil(d/4)) and max |coordinate|.
template <int KMAX, int QB, bool PRE = false>
__global__ void __launch_bounds__(QB)
knn_exact_kernel(const double* __restrict__ x, int64_t n, int d, const double* __restrict__ z,
                 int64_t q, int k, int64_t self_offset, const int32_t* __restrict__ qsel,
                 int64_t split_len, int tile_pts, int32_t* __restrict__ out_idx,
                 double* __restrict__ out_d2, int direct, const float* __restrict__ x32 = nullptr, int dsf = 0,
                 double x_absmax = 0.0) {
  extern __shared__ double sm[];
  double* zs = sm;                    // [d][QB] transposed queries
  double* xs = sm + (size_t)d * QB;   // [tile_pts][d] data tile
  float* xsf = reinterpret_cast<float*>(xs + (size_t)tile_pts * d);  // PRE: [tile_pts][dsf]
  float* zsf = xsf + (size_t)tile_pts * dsf;                         // PRE: [dsf][QB]
  const int t = threadIdx.x;
  const int64_t slot = (int64_t)blockIdx.x * QB + t;  // position in the (selected) query list
  const bool live = slot < q;
  const int64_t p = live ? (qsel ? (int64_t)qsel[slot] : slot) : 0;

  for (int e = t; e < d * QB; e += QB) {
    int qq = e / d, j = e - qq * d;
    int64_t s2 = (int64_t)blockIdx.x * QB + qq;
    double v = 0.0;
    if (s2 < q) {
      int64_t pp = qsel ? (int64_t)qsel[s2] : s2;
      v = z[pp * d + j];
    }
    zs[(size_t)j * QB + qq] = v;
    if (PRE) zsf[(size_t)j * QB + qq] = (float)v;
  }
  float g0 = 0.0f, eps1 = 1.0f, thrw = CUDART_INF_F;
  if (PRE) {
    for (int e = d * QB + t; e < dsf * QB; e += QB) zsf[e] = 0.0f;  // padding dimensions
    __syncthreads();
    double cm = x_absmax;
    for (int j = 0; j < d; ++j) cm = fmax(cm, fabs(zs[(size_t)j * QB + t]));
    g0 = pre_gap(d, cm);
    eps1 = pre_eps1(d);
  }

  double bd[KMAX];
  int bi[KMAX];
#pragma unroll 1
  for (int j = 0; j < k; ++j) {
    bd[j] = CUDART_INF;
    bi[j] = -1;
  }
  double worst = CUDART_INF;
  const int64_t self = (self_offset >= 0) ? self_offset + p : -1;

  const int64_t lo = (int64_t)blockIdx.y * split_len;
  const int64_t hi = min(n, lo + split_len);
  for (int64_t base = lo; base < hi; base += tile_pts) {
    const int cnt = (int)imin64(tile_pts, hi - base);
    __syncthreads();
    {
      const double* src = x + base * d;
      const int tot = cnt * d;
      for (int e = t; e < tot; e += QB) xs[e] = src[e];
      if (PRE) {
        const float4* s4 = reinterpret_cast<const float4*>(x32 + base * dsf);
        float4* d4 = reinterpret_cast<float4*>(xsf);
        for (int e = t; e < cnt * (dsf >> 2); e += QB) d4[e] = s4[e];
      }
    }
    __syncthreads();
    if (live) {
      int i = 0;
      for (; i + 4 <= cnt; i += 4) {
        if (PRE) {
          const int n4 = dsf >> 2;
          const float4* q0 = reinterpret_cast<const float4*>(xsf + (size_t)i * dsf);
          const float4 *q1 = q0 + n4, *q2 = q1 + n4, *q3 = q2 + n4;
          float f0 = 0.0f, f1 = 0.0f, f2 = 0.0f, f3 = 0.0f;
#pragma unroll 2
          for (int j4 = 0; j4 < n4; ++j4) {
            const float* zc = zsf + (size_t)(j4 * 4) * QB + t;
            const float za = zc[0], zb = zc[QB], zcc = zc[2 * QB], zd = zc[3 * QB];
            const float4 v0 = q0[j4], v1 = q1[j4], v2 = q2[j4], v3 = q3[j4];
            float u;
            u = za - v0.x; f0 = fmaf(u, u, f0); u = zb - v0.y; f0 = fmaf(u, u, f0);
            u = zcc - v0.z; f0 = fmaf(u, u, f0); u = zd - v0.w; f0 = fmaf(u, u, f0);
            u = za - v1.x; f1 = fmaf(u, u, f1); u = zb - v1.y; f1 = fmaf(u, u, f1);
            u = zcc - v1.z; f1 = fmaf(u, u, f1); u = zd - v1.w; f1 = fmaf(u, u, f1);
            u = za - v2.x; f2 = fmaf(u, u, f2); u = zb - v2.y; f2 = fmaf(u, u, f2);
            u = zcc - v2.z; f2 = fmaf(u, u, f2); u = zd - v2.w; f2 = fmaf(u, u, f2);
            u = za - v3.x; f3 = fmaf(u, u, f3); u = zb - v3.y; f3 = fmaf(u, u, f3);
            u = zcc - v3.z; f3 = fmaf(u, u, f3); u = zd - v3.w; f3 = fmaf(u, u, f3);
          }
          if (f0 >= thrw && f1 >= thrw && f2 >= thrw && f3 >= thrw) continue;  // none can beat the k-th best
        }
        double a0 = 0, a1 = 0, a2 = 0, a3 = 0;
        const double* r0 = xs + (size_t)i * d;
        const double* r1 = r0 + d;
        const double* r2 = r1 + d;
        const double* r3 = r2 + d;
#pragma unroll 4
        for (int j = 0; j < d; ++j) {
          const double zj = zs[(size_t)j * QB + t];
          double t0 = __dsub_rn(zj, r0[j]), t1 = __dsub_rn(zj, r1[j]);
          double t2 = __dsub_rn(zj, r2[j]), t3 = __dsub_rn(zj, r3[j]);
          a0 = __dadd_rn(a0, __dmul_rn(t0, t0));
          a1 = __dadd_rn(a1, __dmul_rn(t1, t1));
          a2 = __dadd_rn(a2, __dmul_rn(t2, t2));
          a3 = __dadd_rn(a3, __dmul_rn(t3, t3));
        }
        const int g = (int)(base + i);
        if (a0 < worst && g != self) { topk_insert<KMAX>(bd, bi, k, a0, g); worst = bd[k - 1]; }
        if (a1 < worst && g + 1 != self) { topk_insert<KMAX>(bd, bi, k, a1, g + 1); worst = bd[k - 1]; }
        if (a2 < worst && g + 2 != self) { topk_insert<KMAX>(bd, bi, k, a2, g + 2); worst = bd[k - 1]; }
        if (a3 < worst && g + 3 != self) { topk_insert<KMAX>(bd, bi, k, a3, g + 3); worst = bd[k - 1]; }
        if (PRE && worst < CUDART_INF) thrw = pre_thr(pre_up(sqrt(worst)), g0, eps1);
      }
      for (; i < cnt; ++i) {
        double a0 = 0;
        const double* r0 = xs + (size_t)i * d;
        for (int j = 0; j < d; ++j) {
          double t0 = __dsub_rn(zs[(size_t)j * QB + t], r0[j]);
          a0 = __dadd_rn(a0, __dmul_rn(t0, t0));
        }
        const int g = (int)(base + i);
        if (a0 < worst && g != self) { topk_insert<KMAX>(bd, bi, k, a0, g); worst = bd[k - 1]; }
      }
    }
  }
  if (live) {
    const int64_t o = direct ? p * k : ((int64_t)blockIdx.y * q + slot) * k;
    for (int j = 0; j < k; ++j) {
      out_idx[o + j] = bi[j];
      if (out_d2) out_d2[o + j] = bd[j];
    }
  }
}

// k-way merge of `parts` ascending (d2, idx) lists per query.  Lists are
// (parts, q, k); idx -1 marks an empty slot (d2 = +inf).  Output position is
// qsel[slot] when qsel != NULL.
__global__ void merge_topk_kernel(const double* __restrict__ pd2, const int32_t* __restrict__ pidx,
                                  int parts, int64_t q, int k, const int32_t* __restrict__ qsel,
                                  double* __restrict__ od2, int32_t* __restrict__ oidx) {
  const int64_t slot = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (slot >= q) return;
  const int64_t p = qsel ? (int64_t)qsel[slot] : slot;
  unsigned short head[512];
  for (int s = 0; s < parts; ++s) head[s] = 0;
  for (int j = 0; j < k; ++j) {
    double best = CUDART_INF;
    int bidx = 0x7fffffff, bs = -1;
    for (int s = 0; s < parts; ++s) {
      const int h = head[s];
      if (h >= k) continue;
      const int64_t o = ((int64_t)s * q + slot) * k + h;
      const int id = pidx[o];
      if (id < 0) continue;
      const double v = pd2[o];
      if (v < best || (v == best && id < bidx)) {
        best = v;
        bidx = id;
        bs = s;
      }
    }
    if (bs >= 0) head[bs]++;
    oidx[p * k + j] = bs >= 0 ? bidx : -1;
    if (od2) od2[p * k + j] = best;
  }
}

// Same merge, one WARP per query, for many parts (the split scan of a few queries): lane l owns
// lists l, l+32, ... and keeps their current heads in local arrays; each of the k rounds is a
// lane-local minimum, a warp arg-min by (d2, idx), and one load by the winning lane.
__global__ void __launch_bounds__(128)
merge_topk_warp_kernel(const double* __restrict__ pd2, const int32_t* __restrict__ pidx, int parts,
                       int64_t q, int k, const int32_t* __restrict__ qsel, double* __restrict__ od2,
                       int32_t* __restrict__ oidx) {
  const int lane = threadIdx.x & 31;
  const int64_t slot = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (slot >= q) return;
  const int64_t p = qsel ? (int64_t)qsel[slot] : slot;
  constexpr int L = 32;  // parts <= 1024
  double hd[L];
  int hi[L], hp[L];
#pragma unroll 1
  for (int l = 0; l < L; ++l) {
    const int s = lane + 32 * l;
    hd[l] = CUDART_INF;
    hi[l] = 0x7fffffff;
    hp[l] = 0;
    if (s < parts) {
      const int64_t o = ((int64_t)s * q + slot) * k;
      const int id = pidx[o];
      if (id >= 0) {
        hd[l] = pd2[o];
        hi[l] = id;
      }
    }
  }
  const int nl = (parts + 31) / 32;
  for (int j = 0; j < k; ++j) {
    double bd = CUDART_INF;
    int bi = 0x7fffffff, bl = 0;
#pragma unroll 1
    for (int l = 0; l < nl; ++l)
      if (hd[l] < bd || (hd[l] == bd && hi[l] < bi)) {
        bd = hd[l];
        bi = hi[l];
        bl = l;
      }
    double wd = bd;
    int wi = bi, wl = lane;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const double od = __shfl_xor_sync(0xffffffffu, wd, o);
      const int oi = __shfl_xor_sync(0xffffffffu, wi, o);
      const int ol = __shfl_xor_sync(0xffffffffu, wl, o);
      if (od < wd || (od == wd && oi < wi)) {
        wd = od;
        wi = oi;
        wl = ol;
      }
    }
    if (lane == 0) {
      oidx[p * k + j] = wi == 0x7fffffff ? -1 : wi;
      if (od2) od2[p * k + j] = wd;
    }
    if (lane == wl && wi != 0x7fffffff) {  // advance the list that supplied the winner
      const int s = lane + 32 * bl;
      const int h = ++hp[bl];
      hd[bl] = CUDART_INF;
      hi[bl] = 0x7fffffff;
      if (h < k) {
        const int64_t o = ((int64_t)s * q + slot) * k + h;
        const int id = pidx[o];
        if (id >= 0) {
          hd[bl] = pd2[o];
          hi[bl] = id;
        }
      }
    }
  }
}

template <int KMAX>
int launch_exact(b200_model* m, const double* z, int64_t q, int k, int64_t self_offset,
                 const int32_t* qsel, int32_t* idx, double* d2, cudaStream_t st) {
  const int d = m->d;
  const int64_t n = m->n;
  // block size / tile size from the shared-memory budget
  int QB = d <= 96 ? 128 : (d <= 256 ? 64 : 32);
  // few queries (the fallback for uncertified tensor-path queries): one warp of queries per block and
  // small tiles, so that many blocks share an SM and the data range can be split ~1000 ways
  const bool few = q <= 64;
  if (few && q <= 32) QB = 32;
  else if (few && QB > 64) QB = 64;
  const int dsf = ((d + 3) / 4) * 4;
  double absmax = 0.0;
  const float* x32 = nullptr;
  B200_TRY(data_absmax(m, st, &absmax));
  B200_TRY(data_f32(m, st, &x32));
  size_t zbytes = (size_t)d * QB * 8 + (size_t)dsf * QB * 4;
  const size_t per_pt = (size_t)d * 8 + (size_t)dsf * 4;
  size_t budget = few ? 56 * 1024 : 160 * 1024;
  if (budget < zbytes + 16 * per_pt) budget = std::min<size_t>(200 * 1024, zbytes + 16 * per_pt);
  if (budget <= zbytes) {
    set_error("knn_exact: dimension %d too large for the shared-memory tiling", d);
    return 3;
  }
  int tile_pts = (int)((budget - zbytes) / per_pt);
  if (tile_pts > 256) tile_pts = 256;
  tile_pts &= ~3;
  if (tile_pts < 4) {
    set_error("knn_exact: dimension %d too large for the shared-memory tiling", d);
    return 3;
  }
  const size_t smem = zbytes + (size_t)tile_pts * per_pt;
  const int64_t qblocks = (q + QB - 1) / QB;
  // split the data range so that the grid covers >= 2 waves of the SMs
  int64_t S = 1;
  const int64_t want = (few ? 8 : 2) * (int64_t)m->sm_count;
  if (qblocks < want) S = (want + qblocks - 1) / qblocks;
  const int64_t max_s = (n + (int64_t)4 * tile_pts - 1) / ((int64_t)4 * tile_pts);
  if (S > max_s) S = max_s;
  if (S > 1024) S = 1024;
  if (S < 1) S = 1;
  int64_t split_len = (n + S - 1) / S;
  split_len = (split_len + 3) & ~(int64_t)3;
  S = (n + split_len - 1) / split_len;
  const int direct = (S == 1);
  int32_t* oi = idx;
  double* od = d2;
  if (!direct) {
    B200_TRY(m->part_idx.reserve((size_t)S * q * k * sizeof(int32_t)));
    B200_TRY(m->part_d2.reserve((size_t)S * q * k * sizeof(double)));
    oi = m->part_idx.as<int32_t>();
    od = m->part_d2.as<double>();
  } else if (!od) {
    od = nullptr;
  }
  dim3 grid((unsigned)qblocks, (unsigned)S);
#define LAUNCH_QB(QBV)                                                                          \
  do {                                                                                          \
    auto kern = knn_exact_kernel<KMAX, QBV, true>;                                              \
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    kern<<<grid, QBV, smem, st>>>(m->x, n, d, z, q, k, self_offset, qsel, split_len, tile_pts, oi, \
                                  od, direct, x32, dsf, absmax);                                \
  } while (0)
  if (QB == 128) LAUNCH_QB(128);
  else if (QB == 64) LAUNCH_QB(64);
  else LAUNCH_QB(32);
#undef LAUNCH_QB
  B200_CUDA(cudaGetLastError());
  m->st.launches += 1;
  if (!direct) {
    if (S > 32)
      merge_topk_warp_kernel<<<(unsigned)((q * 32 + 127) / 128), 128, 0, st>>>(od, oi, (int)S, q, k, qsel, d2, idx);
    else
      merge_topk_kernel<<<(unsigned)((q + 127) / 128), 128, 0, st>>>(od, oi, (int)S, q, k, qsel, d2, idx);
    B200_CUDA(cudaGetLastError());
    m->st.launches += 1;
  }
  return 0;
}

// ---- k > 1024 (NearestNeighbor's automatic k selection asks for up to n/2 neighbours of ~100 sampled
// points, nearest_neighbor/__init__.py:98-106): all n squared distances of one query at a time, then a
// stable radix sort of (distance bits, row) -- non-negative doubles order like their bit patterns, and
// rows enter in ascending order, so ties go to the lowest index as everywhere else.
__global__ void all_dist_kernel(const double* __restrict__ x, int64_t n, int d, const double* __restrict__ zrow,
                                int64_t self, unsigned long long* __restrict__ key, int32_t* __restrict__ val) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double v = i == self ? CUDART_INF : sqdist_seq(x + i * d, zrow, d);
  key[i] = (unsigned long long)__double_as_longlong(v);
  val[i] = (int32_t)i;
}
__global__ void take_first_kernel(const unsigned long long* __restrict__ key, const int32_t* __restrict__ val, int k,
                                  int32_t* __restrict__ idx, double* __restrict__ d2) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= k) return;
  idx[i] = val[i];
  if (d2) d2[i] = __longlong_as_double((long long)key[i]);
}

int knn_full_sort(b200_model* m, const double* z, int64_t q, int k, int64_t self_offset, const int32_t* qsel,
                  int32_t* idx, double* d2, cudaStream_t st) {
  const int64_t n = m->n;
  if (qsel) {
    set_error("knn: k=%d > 1024 is not available as a fallback path", k);
    return 1;
  }
  DevBuf keys, vals, tmp;
  int rc = keys.reserve((size_t)n * 16 + 64);
  if (!rc) rc = vals.reserve((size_t)n * 8 + 64);
  unsigned long long* k_in = keys.as<unsigned long long>();
  unsigned long long* k_out = k_in + n;
  int32_t* v_in = vals.as<int32_t>();
  int32_t* v_out = v_in + n;
  size_t tmp_bytes = 0;
  if (!rc && cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, k_in, k_out, v_in, v_out, (int)n, 0, 64, st) != cudaSuccess) {
    set_error("knn: radix sort set-up failed");
    rc = 2;
  }
  if (!rc) rc = tmp.reserve(tmp_bytes + 64);
  for (int64_t p = 0; p < q && !rc; ++p) {
    all_dist_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(m->x, n, m->d, z + p * m->d,
                                                                 self_offset >= 0 ? self_offset + p : -1, k_in, v_in);
    if (cub::DeviceRadixSort::SortPairs(tmp.p, tmp_bytes, k_in, k_out, v_in, v_out, (int)n, 0, 64, st) != cudaSuccess) {
      set_error("knn: radix sort failed");
      rc = 2;
      break;
    }
    take_first_kernel<<<(k + 255) / 256, 256, 0, st>>>(k_out, v_out, k, idx + p * k, d2 ? d2 + p * k : nullptr);
    m->st.launches += 3;
  }
  if (!rc && cudaStreamSynchronize(st) != cudaSuccess) {
    set_error("knn: full-sort path failed: %s", cudaGetErrorString(cudaGetLastError()));
    rc = 2;
  }
  keys.release();
  vals.release();
  tmp.release();
  return rc;
}

}  // namespace

// z_dev (>= max query index, d).  qsel == NULL: queries 0..q-1 -> rows 0..q-1 of the
// outputs.  qsel != NULL: nsel selected query indices; results are written to
// the rows they name (fallback for uncertified tensor-path queries).
int knn_exact(b200_model* m, const double* z_dev, int64_t q, int k, int64_t self_offset,
              int32_t* idx_dev, double* d2_dev, const int32_t* qsel, int64_t nsel,
              cudaStream_t st) {
  const int64_t cntq = qsel ? nsel : q;
  if (cntq == 0) return 0;
  if (k < 1 || k > m->n - (self_offset >= 0 ? 1 : 0)) {
    set_error("knn: k=%d out of range for n=%lld", k, (long long)m->n);
    return 1;
  }
  if (k <= 8) return launch_exact<8>(m, z_dev, cntq, k, self_offset, qsel, idx_dev, d2_dev, st);
  if (k <= 32) return launch_exact<32>(m, z_dev, cntq, k, self_offset, qsel, idx_dev, d2_dev, st);
  if (k <= 128) return launch_exact<128>(m, z_dev, cntq, k, self_offset, qsel, idx_dev, d2_dev, st);
  if (k <= 1024) return launch_exact<1024>(m, z_dev, cntq, k, self_offset, qsel, idx_dev, d2_dev, st);
  return knn_full_sort(m, z_dev, cntq, k, self_offset, qsel, idx_dev, d2_dev, st);
}

int merge_topk(const double* d2_parts, const int32_t* idx_parts, int parts, int64_t q, int k,
               double* d2_out, int32_t* idx_out, cudaStream_t st) {
  if (parts > 512) {
    set_error("merge_topk: parts=%d > 512", parts);
    return 1;
  }
  if (q == 0) return 0;
  merge_topk_kernel<<<(unsigned)((q + 127) / 128), 128, 0, st>>>(d2_parts, idx_parts, parts, q, k,
                                                                  nullptr, d2_out, idx_out);
  B200_CUDA(cudaGetLastError());
  return 0;
}


// ------------------------------------------------------------------ rescan of a few queries
// The tensor path's fallback.  A query that failed its certificate already has an exact, sorted list
// of k candidates; its k-th distance B is an upper bound on the true k-th distance, so the true k
// nearest are exactly the k smallest among { j : d2(q, j) <= B }.  One pass over the fitted points
// (thread = point, the few queries broadcast from shared memory) lists those pairs -- a handful per
// query -- and a second kernel ranks each query's list by (d2, index).  Same arithmetic as everywhere
// (direct differences in dimension order, no FMA).  ~0.5 ms for a few queries against 10^6 points,
// where the split brute-force scan takes 2-3 ms; up to 512 queries go through in batches.
namespace {

struct RescanEntry {
  double d2;
  int32_t idx;
  int32_t f;
};

__global__ void rescan_scan_kernel(const double* __restrict__ x, int64_t n, int d, const double* __restrict__ z,
                                   const int32_t* __restrict__ qsel, int nf, int k, int64_t self_offset,
                                   const double* __restrict__ d2_out, int tile_pts, RescanEntry* __restrict__ list,
                                   int cap, int* __restrict__ count) {
  extern __shared__ double sm[];
  const int ds = d | 1;                         // odd row stride: lanes walking different rows hit different banks
  double* zs = sm;                              // [nf][d]
  double* bound = zs + (size_t)nf * d;          // [nf]
  double* xs = bound + ((nf + 1) & ~1);         // [tile_pts][ds]
  const int t = threadIdx.x;
  for (int e = t; e < nf * d; e += blockDim.x) zs[e] = z[(int64_t)qsel[e / d] * d + (e % d)];
  for (int f = t; f < nf; f += blockDim.x) bound[f] = d2_out[(int64_t)qsel[f] * k + (k - 1)];
  const int64_t ntiles = (n + tile_pts - 1) / tile_pts;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t base = tile * tile_pts;
    const int tc = (int)imin64(tile_pts, n - base);
    __syncthreads();
    {  // coalesced copy of the tile's rows into the padded layout, (row, column) advanced without divisions
      const int step_r = (int)blockDim.x / d, step_c = (int)blockDim.x % d;
      int r = t / d, c = t % d;
      const double* src = x + base * d;
      const int tot = tc * d, B = (int)blockDim.x;
      int e = t;
      for (; e + 3 * B < tot; e += 4 * B) {  // four loads in flight per thread
        double v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) v[u] = src[e + u * B];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          xs[(size_t)r * ds + c] = v[u];
          r += step_r;
          c += step_c;
          if (c >= d) {
            c -= d;
            ++r;
          }
        }
      }
      for (; e < tot; e += B) {
        xs[(size_t)r * ds + c] = src[e];
        r += step_r;
        c += step_c;
        if (c >= d) {
          c -= d;
          ++r;
        }
      }
    }
    __syncthreads();
    if (t >= tc) continue;
    const double* xr = xs + (size_t)t * ds;
    const int64_t i = base + t;
    for (int f = 0; f < nf; ++f) {
      const double* zf = zs + (size_t)f * d;
      double s = 0.0;
      for (int j = 0; j < d; ++j) {
        const double df = __dsub_rn(xr[j], zf[j]);
        s = __dadd_rn(s, __dmul_rn(df, df));
      }
      if (s <= bound[f] && !(self_offset >= 0 && i == self_offset + qsel[f])) {
        const int pos = atomicAdd(count, 1);
        if (pos < cap) list[pos] = RescanEntry{s, (int32_t)i, f};
      }
    }
  }
}

constexpr int RESCAN_PER_QUERY = 1024;

// one block per failed query: gather its entries, rank by (d2, index), write the k best
__global__ void __launch_bounds__(128)
rescan_select_kernel(const RescanEntry* __restrict__ list, const int* __restrict__ count, int cap,
                     const int32_t* __restrict__ qsel, int k, int32_t* __restrict__ out_idx,
                     double* __restrict__ out_d2, int* __restrict__ overflow) {
  __shared__ double s_d2[RESCAN_PER_QUERY];
  __shared__ int s_id[RESCAN_PER_QUERY];
  __shared__ int s_n;
  const int f = blockIdx.x, t = threadIdx.x;
  const int total = *count;
  if (total > cap) {  // the list overflowed: the caller falls back to the full scan
    if (t == 0) atomicExch(overflow, 1);
    return;
  }
  if (t == 0) s_n = 0;
  __syncthreads();
  for (int e = t; e < total; e += blockDim.x) {
    const RescanEntry en = list[e];
    if (en.f == f) {
      const int pos = atomicAdd(&s_n, 1);
      if (pos < RESCAN_PER_QUERY) {
        s_d2[pos] = en.d2;
        s_id[pos] = en.idx;
      }
    }
  }
  __syncthreads();
  const int m = s_n;
  if (m > RESCAN_PER_QUERY || m < k) {
    if (t == 0) atomicExch(overflow, 1);
    return;
  }
  const int64_t row = (int64_t)qsel[f] * k;
  for (int e = t; e < m; e += blockDim.x) {
    const double de = s_d2[e];
    const int ie = s_id[e];
    int rank = 0;
    for (int o = 0; o < m; ++o) rank += (s_d2[o] < de || (s_d2[o] == de && s_id[o] < ie)) ? 1 : 0;
    if (rank < k) {
      out_idx[row + rank] = ie;
      if (out_d2) out_d2[row + rank] = de;
    }
  }
}

}  // namespace

int knn_rescan(b200_model* m, const double* z_dev, int64_t q, int k, int64_t self_offset, int32_t* idx_dev,
               double* d2_dev, const int32_t* fail_list, int nf, cudaStream_t st) {
  if (nf == 0) return 0;
  const int d = m->d;
  const int cap = 65536;
  int tile_pts = (int)((64 * 1024) / ((size_t)(d | 1) * 8));
  tile_pts = std::min(256, (tile_pts / 32) * 32);
  // failed queries go through in batches whose coordinates fit next to the tile in shared memory
  const int batch = (int)std::min<size_t>(256, (48 * 1024) / ((size_t)d * 8));
  if (nf > 512 || batch < 1 || !d2_dev || tile_pts < 32)
    return knn_exact(m, z_dev, q, k, self_offset, idx_dev, d2_dev, fail_list, nf, st);
  B200_TRY(m->part_d2.reserve((size_t)cap * sizeof(RescanEntry)));
  B200_TRY(m->flags.reserve(256));
  int* count = m->flags.as<int>() + 24;  // [0] entries listed, [1] overflow flag
  const int64_t ntiles = (m->n + tile_pts - 1) / tile_pts;
  const unsigned grid = (unsigned)std::min<int64_t>(ntiles, (int64_t)m->sm_count * 3);
  for (int lo = 0; lo < nf; lo += batch) {
    const int nb = std::min(batch, nf - lo);
    const int32_t* fl = fail_list + lo;
    const size_t smem = ((size_t)nb * d + ((nb + 1) & ~1) + (size_t)tile_pts * (d | 1)) * 8;
    B200_CUDA(cudaMemsetAsync(count, 0, 8, st));
    B200_CUDA(cudaFuncSetAttribute(rescan_scan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    rescan_scan_kernel<<<grid, tile_pts, smem, st>>>(m->x, m->n, d, z_dev, fl, nb, k, self_offset, d2_dev, tile_pts,
                                                     m->part_d2.as<RescanEntry>(), cap, count);
    B200_CUDA(cudaGetLastError());
    rescan_select_kernel<<<nb, 128, 0, st>>>(m->part_d2.as<RescanEntry>(), count, cap, fl, k, idx_dev, d2_dev,
                                             count + 1);
    B200_CUDA(cudaGetLastError());
    int overflow = 0;
    B200_CUDA(cudaMemcpyAsync(&overflow, count + 1, 4, cudaMemcpyDeviceToHost, st));
    B200_CUDA(cudaStreamSynchronize(st));
    m->st.launches += 2;
    if (overflow) B200_TRY(knn_exact(m, z_dev, q, k, self_offset, idx_dev, d2_dev, fl, nb, st));
  }
  return 0;
}

}  // namespace b200

// ===========================================================================
// Allocation-free exact kNN.
//
// Some callers must not trigger device allocation while they run: serving loops
// with a latency budget, for which the implicit device-wide synchronisation of
// cudaMalloc/cudaFree is a stall; and hosts whose device memory is already
// handed out in full to a pool or another framework, where a late cudaMalloc
// from a library can simply fail, or where every allocation is audited.
// For them this entry point runs the same split-scan + merge kernels as
// b200_knn_dev's exact path, but the partial top-k lists live in the file-scope
// array below instead of a cudaMalloc'ed buffer.  Every GPU that loads the
// library carries the array (12 MiB) until the process exits.  The first call
// on a device looks up the array's device address (do that first call before
// any stream capture begins).  Nothing in the array needs zeroing -- a call
// writes every slot it later reads -- but all calls share it, so calls whose
// kernels may overlap on the GPU race and must be serialised by the caller.
// ===========================================================================
namespace b200 {
namespace {

constexpr int64_t NA_MAX_ENTRIES = 1 << 20;  // partial (d2, idx) pairs: splits * q * k must fit
// [0, 8 MiB): double d2 parts, [8 MiB, 12 MiB): int32 idx parts
static __device__ __align__(256) unsigned char g_noalloc_scratch[NA_MAX_ENTRIES * 12];

// Device address of g_noalloc_scratch per device.  In host code the symbol name is only a
// handle: the address a kernel can use must be asked for, once per device, with that
// device current.
void* g_na_ptr[64] = {nullptr};

int noalloc_scratch(void** out) {
  int dev = 0;
  B200_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) {
    set_error("b200_knn_exact_noalloc: device ordinal %d out of range", dev);
    return 1;
  }
  if (!g_na_ptr[dev]) {
    void* p = nullptr;
    B200_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    g_na_ptr[dev] = p;
  }
  *out = g_na_ptr[dev];
  return 0;
}

template <int KMAX>
int launch_noalloc(const double* x, int64_t n, int d, const double* z, int64_t q, int k,
                   int32_t* idx, double* d2, cudaStream_t st) {
  void* scratch = nullptr;
  B200_TRY(noalloc_scratch(&scratch));
  int sms = 148, dev = 0;
  B200_CUDA(cudaGetDevice(&dev));
  B200_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int QB = d <= 96 ? 128 : (d <= 256 ? 64 : 32);
  const size_t zbytes = (size_t)d * QB * 8;
  const size_t budget = 160 * 1024;
  if (zbytes + (size_t)4 * d * 8 > budget) {  // the query block and a 4-point tile must fit
    set_error("b200_knn_exact_noalloc: dimension %d too large for the shared-memory tiling", d);
    return 3;
  }
  int tile_pts = (int)((budget - zbytes) / ((size_t)d * 8));
  if (tile_pts > 256) tile_pts = 256;
  tile_pts &= ~3;
  if (tile_pts < 4) {
    set_error("b200_knn_exact_noalloc: dimension %d too large for the shared-memory tiling", d);
    return 3;
  }
  const size_t smem = zbytes + (size_t)tile_pts * d * 8;
  const int64_t qblocks = (q + QB - 1) / QB;
  int64_t S = (2 * (int64_t)sms + qblocks - 1) / qblocks;               // aim at two waves of CTAs
  S = std::min<int64_t>(S, (n + 4 * tile_pts - 1) / (4 * (int64_t)tile_pts));
  S = std::min<int64_t>(S, NA_MAX_ENTRIES / (q * k));                    // what the static scratch holds
  S = std::max<int64_t>(1, std::min<int64_t>(S, 512));
  int64_t split_len = (n + S - 1) / S;
  split_len = (split_len + 3) & ~(int64_t)3;
  S = (n + split_len - 1) / split_len;
  const int direct = (S == 1);
  double* pd2 = reinterpret_cast<double*>(scratch);
  int32_t* pidx = reinterpret_cast<int32_t*>(reinterpret_cast<unsigned char*>(scratch) + NA_MAX_ENTRIES * 8);
  dim3 grid((unsigned)qblocks, (unsigned)S);
#define NA_LAUNCH(QBV)                                                                             \
  do {                                                                                             \
    auto kern = knn_exact_kernel<KMAX, QBV, false>; /* no model here: no float image, no prefilter */ \
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    kern<<<grid, QBV, smem, st>>>(x, n, d, z, q, k, -1, nullptr, split_len, tile_pts,             \
                                  direct ? idx : pidx, direct ? d2 : pd2, direct, nullptr, 0, 0.0); \
  } while (0)
  if (QB == 128) NA_LAUNCH(128);
  else if (QB == 64) NA_LAUNCH(64);
  else NA_LAUNCH(32);
#undef NA_LAUNCH
  B200_CUDA(cudaGetLastError());
  if (!direct) {
    merge_topk_kernel<<<(unsigned)((q + 127) / 128), 128, 0, st>>>(pd2, pidx, (int)S, q, k, nullptr, d2, idx);
    B200_CUDA(cudaGetLastError());
  }
  return 0;
}

}  // namespace
}  // namespace b200

extern "C" int b200_knn_exact_noalloc(const double* x_dev, int64_t n, int32_t d, const double* z_dev,
                                      int64_t q, int32_t k, int32_t* idx_dev, double* d2_dev,
                                      void* stream) {
  using namespace b200;
  if (!x_dev || !z_dev || !idx_dev || !d2_dev || n < 1 || d < 1 || q < 0) {
    set_error("b200_knn_exact_noalloc: bad arguments");
    return 1;
  }
  if (q == 0) return 0;
  if (k < 1 || k > n || k > 128) {
    set_error("b200_knn_exact_noalloc: k=%d must be in [1, min(n, 128)]", k);
    return 1;
  }
  if (n > 0x7fffffffLL - 1024) {
    set_error("b200_knn_exact_noalloc: n=%lld exceeds int32 row indices", (long long)n);
    return 1;
  }
  if (q * (int64_t)k * 4 > NA_MAX_ENTRIES) {
    set_error("b200_knn_exact_noalloc: q*k = %lld exceeds the static scratch limit of %lld "
              "(four partial lists of q*k entries must fit in %lld); split the query batch",
              (long long)(q * k), (long long)(NA_MAX_ENTRIES / 4), (long long)NA_MAX_ENTRIES);
    return 1;
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (k <= 8) return launch_noalloc<8>(x_dev, n, d, z_dev, q, k, idx_dev, d2_dev, st);
  if (k <= 32) return launch_noalloc<32>(x_dev, n, d, z_dev, q, k, idx_dev, d2_dev, st);
  return launch_noalloc<128>(x_dev, n, d, z_dev, q, k, idx_dev, d2_dev, st);
}
