// knn_tc_ts.cuh -- stage 1 of the tensor-core kNN, TS form (included by knn_tc.cu inside b200::tc).
//
// Same contraction, threshold filter and candidate buffers as knn_tc_kernel, with a deeper and finer
// accumulator pipeline.  The clock trace of the SS kernel (profiles/r01) shows what bounds it: with
// two 128-column accumulator stages per half, the loop MMA -> commit -> epilogue -> release -> next
// MMA on the same stage (~240 clk pipe latency + ~220 clk per mbarrier hand-off + the epilogue) is
// longer than the tensor time of the other stage, so the pipe idles and every survivor handled while
// the accumulator is held adds to the period.  Here:
//   * the query tile (A operand) lives in TMEM for the whole data stream, written once per query
//     block by the epilogue warps (tcgen05.st); the MMAs are the TS form, which leaves the whole
//     shared-memory port to B and runs N=64 at the full 32 clk per K step;
//   * the accumulator unit is 64 columns: TMEM holds S = 3..6 stages per half next to A, so the
//     stage loop has S-1 units of slack instead of one;
//   * an epilogue warp copies its 64 values to registers and hands the stage back at once; the
//     min trees, the vote, appends and compactions all run after the release;
//   * two issuer threads per half on alternate units; the B ring holds 64-row sub-tile images
//     (one bulk copy each).
#pragma once

// One thread issues every MMA of one half.  KS: compile-time K/16 (0: run-time loop).
template <int HALVES, int KS, bool DBG>
__device__ __forceinline__ void ts_issuer(const KParams& p, int h, int j, int S, int nqb, int nunits, const uint8_t* sB,
                                          uint64_t* bar_full, uint64_t* bar_empty, uint64_t* bar_accf,
                                          uint64_t* bar_acce, uint64_t* bar_afull, uint32_t a_col0) {
  constexpr int UNIT = 64;
  const int K = p.K;
  const int ksteps = K / 16;
  const uint32_t sub_bytes = (uint32_t)UNIT * K * 2;
  const uint32_t idesc = ptx::umma_idesc_f16(TILE, UNIT);
  const uint64_t b_desc0 = ptx::umma_desc(ptx::smem_u32(sB), UNIT * 16, 128);  // [K/8][64][16 B]
  constexpr uint32_t b_step = (UNIT * 32) >> 4;
  const uint32_t slot_step = sub_bytes >> 4;
  const uint32_t a_col = a_col0 + (uint32_t)h * (uint32_t)(K >> 1);
  const uint32_t full0 = ptx::smem_u32(&bar_full[0]), empty0 = ptx::smem_u32(&bar_empty[0]);
  const uint32_t accf0 = ptx::smem_u32(&bar_accf[h * 8]), acce0 = ptx::smem_u32(&bar_acce[h * 8]);
  const uint32_t afull = ptx::smem_u32(&bar_afull[h]);
  const uint32_t d0 = (uint32_t)(h * S) * UNIT;
  const uint32_t nslots = (uint32_t)p.slots;
  const bool skip_mma = DBG && (p.mode & 8);
  // this thread takes every second unit (j = 0: even, 1: odd; the unit count per block is even)
  uint32_t s = (uint32_t)j, sph = 1, sl = (uint32_t)j, fph = 0;  // sph: parity the acce wait looks for
  uint32_t d_tmem = d0 + s * UNIT, accf = accf0 + s * 8, acce = acce0 + s * 8, full = full0 + sl * 8,
           empty = empty0 + sl * 8;
  uint64_t bd = b_desc0 + (uint64_t)(sl * slot_step);
  uint32_t ju = (uint32_t)j;
  for (int b = 0; b < nqb; ++b) {
    while (!ptx::mbar_try_wait(afull, b & 1)) {}
    for (int u = j; u < nunits; u += 2) {
      while (!ptx::mbar_try_wait(acce, sph)) {}
      while (!ptx::mbar_try_wait(full, fph)) {}
      ptx::tc_fence_after();
      if (DBG && h == 0 && j == 0) B200_TRACE(1, (int)ju, 0);
      if (!skip_mma) {
        if (KS > 0) {
#pragma unroll
          for (int ks = 0; ks < KS; ++ks)
            ptx::umma_f16_ts(d_tmem, a_col + ks * 8, bd + (uint64_t)(ks * b_step), idesc, ks > 0 ? 1u : 0u);
        } else {
          for (int ks = 0; ks < ksteps; ++ks)
            ptx::umma_f16_ts(d_tmem, a_col + ks * 8, bd + (uint64_t)(ks * b_step), idesc, ks > 0 ? 1u : 0u);
        }
      }
      if (DBG && h == 0 && j == 0) B200_TRACE(1, (int)ju, 1);
      ptx::umma_commit(accf);
      ptx::umma_commit(empty);
      if (DBG && h == 0 && j == 0) B200_TRACE(1, (int)ju, 3);
      if (DBG) ju += 2;
      s += 2;
      if (s >= (uint32_t)S) {
        s -= (uint32_t)S;
        sph ^= 1;
      }
      d_tmem = d0 + s * UNIT;
      accf = accf0 + s * 8;
      acce = acce0 + s * 8;
      sl += 2;
      if (sl >= nslots) {
        sl -= nslots;
        fph ^= 1;
      }
      bd = b_desc0 + (uint64_t)(sl * slot_step);
      full = full0 + sl * 8;
      empty = empty0 + sl * 8;
    }
  }
}

template <int CAP, int HALVES, bool DBG>
__global__ void __launch_bounds__((HALVES * 6 + 1) * 32, 1) knn_ts_kernel(const KParams p) {
  constexpr int ROWS = HALVES * TILE;
  constexpr int RS = ROWS + 1;
  constexpr int EPI_WARPS = HALVES * 4;
  constexpr int UNIT = 64;  // data columns per accumulator stage = rows of a B sub-tile = UMMA N
  extern __shared__ __align__(1024) uint8_t smem[];
  const int K = p.K;
  const int S = p.stages;
  const uint32_t sub_bytes = (uint32_t)UNIT * K * 2;
  uint8_t* sB = smem;                                                       // ring of p.slots sub-tiles
  float* bsc = reinterpret_cast<float*>(sB + (size_t)p.slots * sub_bytes);  // [CAP][RS] scores
  int* bid = reinterpret_cast<int*>(bsc + (size_t)CAP * RS);                // [CAP][RS] data indices
  uint64_t* bars = reinterpret_cast<uint64_t*>((reinterpret_cast<uintptr_t>(bid + (size_t)CAP * RS) + 15) & ~(uintptr_t)15);
  uint64_t* bar_full = bars;        // [16]   B slot filled (TMA bytes landed)
  uint64_t* bar_empty = bars + 16;  // [16]   B slot free (every half's MMAs on it retired)
  uint64_t* bar_accf = bars + 32;   // [h][8] accumulator stage written
  uint64_t* bar_acce = bars + 48;   // [h][8] accumulator stage copied out by the half's four warps
  uint64_t* bar_afull = bars + 64;  // [h]    query rows of the half stored to TMEM
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 66);

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.slots; ++s) {
      ptx::mbar_init(ptx::smem_u32(&bar_full[s]), 1);
      ptx::mbar_init(ptx::smem_u32(&bar_empty[s]), HALVES);
    }
    for (int h = 0; h < HALVES; ++h) {
      for (int s = 0; s < S; ++s) {
        ptx::mbar_init(ptx::smem_u32(&bar_accf[h * 8 + s]), 1);
        ptx::mbar_init(ptx::smem_u32(&bar_acce[h * 8 + s]), 4);
      }
      ptx::mbar_init(ptx::smem_u32(&bar_afull[h]), 4);
    }
    ptx::fence_barrier_init();
  }
  if (warp == EPI_WARPS + 1) {
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
  const int nunits = 2 * p.ntiles;  // 64-column units of the data stream
  const bool hint = p.flags & 1;
  const uint32_t bsleep = (p.flags & 16) ? 0 : 64;  // producer only
  const uint32_t a_col0 = (uint32_t)(HALVES * S * UNIT);  // TMEM: [h][s] accumulators, then A of each half

  if (warp == EPI_WARPS) {
    // ===================== producer: one bulk copy per 64-row sub-tile =====================
    if (lane == 0) {
      uint32_t s = 0, ph = 0;
      const uint32_t full0 = ptx::smem_u32(&bar_full[0]), empty0 = ptx::smem_u32(&bar_empty[0]);
      const uint32_t sB0 = ptx::smem_u32(sB);
      uint32_t ju = 0;
      for (int b = 0; b < nqb; ++b) {
        const uint8_t* src = reinterpret_cast<const uint8_t*>(p.b_img);
        for (int u = 0; u < nunits; ++u, ++ju, src += sub_bytes) {
          ptx::mbar_wait(hint, empty0 + s * 8, ph ^ 1, bsleep);
          if (DBG) B200_TRACE(0, (int)ju, 0);
          if (DBG && (p.mode & 4) && ju >= (uint32_t)p.slots) {  // debug: no copy, the slot keeps its old tile
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
    // ===================== MMA issuers: two warps per half, one thread each =====================
    // The per-unit path has to stay well under the 320 clk the two halves' MMAs of one unit take, so
    // it is a single thread with incrementally updated addresses: two barrier probes, the fence, five
    // UTCHMMA, two commits (no elect/reconvergence, no parameter reloads, no trace in this build).
    // Two such threads per half on alternate units: one queues MMAs while the other is in its
    // waits and commits (a lone issuer leaves the pipe idle ~60 % of the time, clock trace).
    const int h = (warp - (EPI_WARPS + 1)) >> 1, j = (warp - (EPI_WARPS + 1)) & 1;
    if (lane == 0) {
      const int ksteps = K / 16;
      switch (ksteps) {
        case 2: ts_issuer<HALVES, 2, DBG>(p, h, j, S, nqb, nunits, sB, bar_full, bar_empty, bar_accf, bar_acce, bar_afull, a_col0); break;
        case 3: ts_issuer<HALVES, 3, DBG>(p, h, j, S, nqb, nunits, sB, bar_full, bar_empty, bar_accf, bar_acce, bar_afull, a_col0); break;
        case 4: ts_issuer<HALVES, 4, DBG>(p, h, j, S, nqb, nunits, sB, bar_full, bar_empty, bar_accf, bar_acce, bar_afull, a_col0); break;
        case 5: ts_issuer<HALVES, 5, DBG>(p, h, j, S, nqb, nunits, sB, bar_full, bar_empty, bar_accf, bar_acce, bar_afull, a_col0); break;
        default: ts_issuer<HALVES, 0, DBG>(p, h, j, S, nqb, nunits, sB, bar_full, bar_empty, bar_accf, bar_acce, bar_afull, a_col0); break;
      }
    }
  } else {
    // ===================== epilogue: copy out, release, then filter =====================
    const int half = warp >> 2, quad = warp & 3;
    const int row0 = half * TILE + quad * 32;
    const int row = row0 + lane;  // my query row inside the block
    const uint32_t t_lane = (uint32_t)(quad * 32) << 16;
    const uint32_t a_col = a_col0 + (uint32_t)half * (uint32_t)(K >> 1);
    const uint32_t accf0 = ptx::smem_u32(&bar_accf[half * 8]), acce0 = ptx::smem_u32(&bar_acce[half * 8]);
    const int keep = p.keep;
    float* my_sc = bsc + row;
    int* my_id = bid + row;
    uint32_t s = 0, sph = 0, ju = 0;
    for (int b = 0; b < nqb; ++b) {
      const int qb = (int)blockIdx.x + b * (int)gridDim.x;
      // ---- my query row -> TMEM.  The previous block's MMAs of this half have all retired: this
      // warp has consumed the last accumulator they wrote.
      {
        const uint4* src = reinterpret_cast<const uint4*>(p.a_img + ((size_t)qb * HALVES + half) * TILE * K) + quad * 32 + lane;
        const int KC = K / 8;
        for (int c = 0; c < KC; ++c) ptx::tmem_st4(t_lane + a_col + c * 4, __ldg(src + (size_t)c * TILE));
        ptx::tmem_st_wait();
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&bar_afull[half]));
      }
      float tau = CUDART_INF_F;
      int cnt = 0;
      for (int u = 0; u < nunits; ++u, ++ju) {
        ptx::mbar_wait(hint, accf0 + s * 8, sph);
        if (DBG && warp == 0 && lane == 0) B200_TRACE(2, (int)ju, 0);
        ptx::tc_fence_after();
        const uint32_t t_acc = t_lane + (uint32_t)(half * S + s) * UNIT;
        float va[32], vb[32];
        if ((p.mode & 3) != 2) {
          ptx::tmem_ld32(t_acc, va);
          ptx::tmem_ld32(t_acc + 32, vb);
          ptx::tmem_ld_wait();
        }
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(acce0 + s * 8);  // stage handed back before any filtering
        if (DBG && warp == 0 && lane == 0) B200_TRACE(2, (int)ju, 1);
        if (++s == (uint32_t)S) {
          s = 0;
          sph ^= 1;
        }
        if ((p.mode & 3) == 2) continue;
        if (DBG) {
          if (p.dbg) {
            float* o = p.dbg + ((size_t)qb * ROWS + row) * ((size_t)p.ntiles * TILE) + (size_t)u * UNIT;
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              o[i] = va[i];
              o[32 + i] = vb[i];
            }
          }
        }
        float mc0, mc1;
#define B200_CHUNKMIN(v, out)                                                              \
  do {                                                                                     \
    float g_[4];                                                                           \
    _Pragma("unroll") for (int gi = 0; gi < 4; ++gi) {                                     \
      const float* w = v + gi * 8;                                                         \
      g_[gi] = fminf(min3(w[0], w[1], w[2]), min3(min3(w[3], w[4], w[5]), w[6], w[7]));    \
    }                                                                                      \
    out = fminf(fminf(g_[0], g_[1]), fminf(g_[2], g_[3]));                                 \
  } while (0)
        B200_CHUNKMIN(va, mc0);
        B200_CHUNKMIN(vb, mc1);
#undef B200_CHUNKMIN
        unsigned hitc = (mc0 < tau ? 1u : 0u) | (mc1 < tau ? 2u : 0u);
        hitc = (p.mode & 3) == 0 ? __reduce_or_sync(0xffffffffu, hitc) : 0u;
        if (!hitc) continue;
        // ---- slow path: the values are still in registers
        if (hitc & 1u) B200_APPEND_CHUNK(va, u * UNIT, mc0);
        if (hitc & 2u) B200_APPEND_CHUNK(vb, u * UNIT + 32, mc1);
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
