// Decode-shaped (M <= 64) Marlin-layout W4A16 / W8A16 GEMM for sm_100a: small CTAs, several per SM.
//
// Same stored layouts and the same operand path as marlin_gemm_kernel.cuh (packed tiles by TMA -> dequant warps expand
// them in registers -> tcgen05.st straight into TMEM as the A operand of a swap-AB tcgen05.mma; activations are the
// 128B-swizzled B operand from shared memory), but a different machine around it.  Round-1 measurements
// (profiles/r1_gemm_*) showed the one-fat-CTA-per-SM kernel to be bound by its own hand-offs (~950 clk per 256x128 unit
// with the dequant ALU removed, ~4.5 us fixed per launch with HBM idle between launches) -- so here:
//
//   * a CTA is 8 warps (w0 weight TMA | w1 activation + scale TMA | w2 TMEM alloc + MMA issuer | w3 spare | w4-7
//     dequant, one per TMEM lane quarter) with <= 75 KB of shared memory and 128 TMEM columns: THREE fit on an SM.
//     Latency is hidden by independent CTAs instead of by choreography inside one, and the grid is sized BELOW the
//     number of resident slots so that the next GEMM of the stream (PDL) starts streaming its weights while this one
//     is still in its main loop or tail: HBM stays busy across launches;
//   * one accumulator per CTA lives in TMEM for the whole K range of a tile segment; group scales go onto the weights in
//     the activation dtype (exactly what the reference does, gptq_marlin.cu:270-278), so nothing is read back from TMEM
//     per quantisation group and the dequant warps double as the epilogue (they own the TMEM lane quarters anyway);
//   * work is split stream-K style over 128-deep k-stages; a tile shared by several CTAs is reduced by the LAST CTA to
//     arrive (ticket counter; nobody spins, so there is no co-residency requirement), in CTA order -> deterministic.
//
// Reference being replaced: csrc/quantization/gptq_marlin/gptq_marlin.cu:410-1363 (kernel), :1577-1731 (launcher).
#pragma once
#include "marlin_gemm_kernel.cuh"

namespace b200 {

constexpr int kDecStageK = 128;            // one pipeline stage / scheduling unit = 128 (k) x 128 (n) weights
constexpr int kDecSlabs = kDecStageK / 16;
constexpr int kDecCtrlWarps = 4;

#ifndef B200_DEC_OCC
#define B200_DEC_OCC 3
#endif
#ifndef B200_DEC_DQW64
#define B200_DEC_DQW64 1
#endif
#ifndef B200_DEC_DQW
#define B200_DEC_DQW 1
#endif

template <int MPAD, int BITS>
struct DecCfg {
  static constexpr int DQW = MPAD >= 64 ? B200_DEC_DQW64 : B200_DEC_DQW;   // dequant warps per TMEM lane quarter
  static constexpr int NDQ = 4 * DQW;
  static constexpr int THREADS = 32 * (kDecCtrlWarps + NDQ);
  static constexpr int EW = MPAD / DQW;                                      // accumulator columns per epilogue thread
  static constexpr int W_ROW_BYTES = kTileN * 16 * BITS / 8;                 // one 16(k) x 128(n) slab
  static constexpr int W_STAGE_BYTES = kDecSlabs * W_ROW_BYTES;              // 8 KB (4-bit) / 16 KB (8-bit)
  static constexpr int A_BOX_BYTES = MPAD * 128;                             // one 64(k) x MPAD chunk, 128-byte swizzle
  static constexpr int A_STAGE_BYTES = 2 * A_BOX_BYTES;
  static constexpr int S_STAGE_BYTES = 1024;                                 // up to four 128-column scale rows (group 32)
  static constexpr int OCC_WANT = (MPAD >= 64 && BITS == 8) ? 2 : B200_DEC_OCC;
  static constexpr int NSLOT = MPAD >= 64 ? 2 * DQW : 3;                     // 64-k operand slots in TMEM
  static constexpr int TMEM_NEED = NSLOT * 32 + MPAD;
  static constexpr int TMEM_COLS = TMEM_NEED <= 128 ? 128 : (TMEM_NEED <= 256 ? 256 : 512);
  static constexpr int OCC = OCC_WANT * TMEM_COLS > 512 ? 512 / TMEM_COLS : OCC_WANT;
  static constexpr int NA = MPAD <= 16 ? 4 : (MPAD <= 32 ? 3 : 2);
  static constexpr int NBARS_FIXED = 2 * NA + 2 * NSLOT + 2;
  static constexpr int BUDGET = (227 * 1024) / OCC - 1024 /*driver reserve*/ - 1024 /*align slack*/ - 512 /*barriers*/;
  static constexpr int NW_RAW = (BUDGET - NA * (A_STAGE_BYTES + S_STAGE_BYTES)) / W_STAGE_BYTES;
  static constexpr int NW = NW_RAW > 8 ? 8 : NW_RAW;
  static constexpr int NBARS = 2 * NW + NBARS_FIXED;
  static constexpr int SMEM_BYTES = 1024 + NA * (A_STAGE_BYTES + S_STAGE_BYTES) + NW * W_STAGE_BYTES + NBARS * 8 + 16;
  static_assert(NW >= 2, "weight ring needs at least two stages");
  static_assert(NBARS * 8 + 16 <= 512, "barrier area");
  static_assert(EW % 16 == 0, "epilogue reads 16 columns at a time");
  static_assert(OCC * TMEM_COLS <= 512, "TMEM budget per SM");
};

// partial-sum slots and ticket counters of the decode kernel: fixed size, allocated once per device, never regrown
constexpr int kDecMaxGrid = 148 * 4 + 48;
constexpr int kDecMaxTiles = 1 << 16;
int dec_scratch(int dev, cudaStream_t st, float** partials, int** counters);   // marlin_gemm.cu

template <typename T, int MPAD, int BITS, int GPS>   // GPS: scale rows per stage -- 0 channelwise, 1 (g128), 2 (g64), 4 (g32)
__global__ void __launch_bounds__(DecCfg<MPAD, BITS>::THREADS, DecCfg<MPAD, BITS>::OCC)
marlin_decode_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_w,
                     const __grid_constant__ CUtensorMap tmap_s, const GemmArgs args) {
  using Cfg = DecCfg<MPAD, BITS>;
  constexpr int NW = Cfg::NW, NA = Cfg::NA, NSLOT = Cfg::NSLOT, DQW = Cfg::DQW;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  // layout: [activation stages (1024-aligned)] [weight stages] [scale stages] [barriers]
  const uint32_t act_base = smem_base;
  const uint32_t w_base = act_base + NA * Cfg::A_STAGE_BYTES;
  const uint32_t sc_base = w_base + NW * Cfg::W_STAGE_BYTES;
  const uint32_t bar_base = sc_base + NA * Cfg::S_STAGE_BYTES;
  auto wfull_bar = [&](int s) { return bar_base + 8u * s; };
  auto wempty_bar = [&](int s) { return bar_base + 8u * (NW + s); };
  auto actfull_bar = [&](int s) { return bar_base + 8u * (2 * NW + s); };
  auto actempty_bar = [&](int s) { return bar_base + 8u * (2 * NW + NA + s); };
  auto slotfull_bar = [&](int s) { return bar_base + 8u * (2 * NW + 2 * NA + s); };
  auto slotempty_bar = [&](int s) { return bar_base + 8u * (2 * NW + 2 * NA + NSLOT + s); };
  const uint32_t accfull_bar = bar_base + 8u * (2 * NW + 2 * NA + 2 * NSLOT);
  const uint32_t accempty_bar = accfull_bar + 8u;
  const uint32_t tmem_slot = bar_base + 8u * Cfg::NBARS;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  volatile uint32_t* tmem_slot_p = reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_slot - smem_base));
  volatile int* ticket_p = reinterpret_cast<volatile int*>(smem_gen + (tmem_slot - smem_base) + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmap_w);
    prefetch_tmap(&tmap_a);
    if (GPS != 0) prefetch_tmap(&tmap_s);
  }
  if (threadIdx.x >= 32 && threadIdx.x < 32 + Cfg::NBARS) {
    const int i = threadIdx.x - 32;
    uint32_t cnt;
    if (i < NW) cnt = 1;                                         // wfull: producer's expect_tx arrive
    else if (i < 2 * NW) cnt = Cfg::NDQ;                         // wempty: every dequant warp
    else if (i < 2 * NW + NA) cnt = 1;                           // actfull
    else if (i < 2 * NW + 2 * NA) cnt = 1 + (GPS != 0 ? Cfg::NDQ : 0);   // actempty: issuer commit (+ scale readers)
    else if (i < 2 * NW + 2 * NA + NSLOT) cnt = 4;               // slotfull: the four lane-quarter warps of the slot
    else if (i < 2 * NW + 2 * NA + 2 * NSLOT) cnt = 1;           // slotempty: issuer commit
    else if (i == 2 * NW + 2 * NA + 2 * NSLOT) cnt = 1;          // accfull: issuer commit
    else cnt = Cfg::NDQ;                                         // accempty: every dequant (= epilogue) warp
    mbar_init(bar_base + 8u * i, cnt);
    fence_barrier_init();
  }
  __syncthreads();
  pdl_launch_dependents();

  const int KS = args.KS;
  const Partition part(args.total_units, (int)gridDim.x);
  const int u_begin = part.start((int)blockIdx.x);
  const int u_end = part.start((int)blockIdx.x + 1);
  const int n_units = u_end - u_begin;

  auto get_tmem_base = [&]() -> uint32_t {
    tc_fence_before();
    asm volatile("bar.sync 2, %0;" ::"n"(Cfg::THREADS - 64) : "memory");
    tc_fence_after();
    return *tmem_slot_p;
  };

  if (warp == 0) {
    // ================= weight producer: no dependence on the previous kernel of the stream (no PDL wait) ==============
    const uint64_t pol_w = policy_evict_first();
    int tile = u_begin / KS, ks = u_begin - tile * KS;
    int s = 0, ph = 1;
    for (int it = 0; it < n_units; ++it) {
      mbar_wait(wempty_bar(s), ph);
      if (elect_one()) {
        mbar_arrive_expect_tx(wfull_bar(s), Cfg::W_STAGE_BYTES);
        tma_2d_g2s_hint(w_base + s * Cfg::W_STAGE_BYTES, &tmap_w, tile * (Cfg::W_ROW_BYTES / 8), ks * kDecSlabs,
                        wfull_bar(s), pol_w);
      }
      __syncwarp();
      if (++s == NW) { s = 0; ph ^= 1; }
      if (++ks == KS) { ks = 0; ++tile; }
    }
  } else if (warp == 1) {
    // ================= activation (+ group scale) producer ==========================================================
    const uint64_t pol_a = policy_evict_last();
    constexpr uint32_t tx = Cfg::A_STAGE_BYTES + GPS * 256;
    int tile = u_begin / KS, ks = u_begin - tile * KS;
    int s = 0, ph = 1;
    pdl_wait();
    for (int it = 0; it < n_units; ++it) {
      mbar_wait(actempty_bar(s), ph);
      if (elect_one()) {
        mbar_arrive_expect_tx(actfull_bar(s), tx);
        tma_3d_g2s_hint(act_base + s * Cfg::A_STAGE_BYTES, &tmap_a, 0, 0, ks * 2, actfull_bar(s), pol_a);
        if (GPS != 0)
          tma_2d_g2s_hint(sc_base + s * Cfg::S_STAGE_BYTES, &tmap_s, tile * kTileN, ks * GPS, actfull_bar(s), pol_a);
      }
      __syncwarp();
      if (++s == NA) { s = 0; ph ^= 1; }
      if (++ks == KS) { ks = 0; ++tile; }
    }
  } else if (warp == 2) {
    // ================= TMEM allocator + MMA issuer ====================================================================
    tmem_alloc<Cfg::TMEM_COLS>(tmem_slot);
    const uint32_t tmem_base = get_tmem_base();
    constexpr uint32_t idesc = make_idesc(1 /*F32*/, Pack2<T>::FMT, Pack2<T>::FMT, 128, MPAD);
    const uint32_t d_tmem = tmem_base + NSLOT * 32;
    int ks = u_begin % KS;
    int sa = 0, pha = 0, slot = 0, phs = 0, seg = 0;
    for (int it = 0; it < n_units; ++it) {
      const bool first = (it == 0 || ks == 0);
      const bool last = (it == n_units - 1 || ks == KS - 1);
      mbar_wait(actfull_bar(sa), pha);
      if (first && seg > 0) mbar_wait(accempty_bar, (seg - 1) & 1);   // the epilogue has read the previous segment
      const uint32_t b_smem = act_base + sa * Cfg::A_STAGE_BYTES;
#pragma unroll
      for (int sub = 0; sub < 2; ++sub) {
        mbar_wait(slotfull_bar(slot), phs);
        tc_fence_after();
        if (elect_one()) {
#pragma unroll
          for (int rr = 0; rr < 4; ++rr)
            mma_f16_ts(d_tmem, tmem_base + slot * 32 + rr * 8,
                       make_kmajor_sw128_desc(b_smem + sub * Cfg::A_BOX_BYTES + rr * 32), idesc,
                       (first && sub == 0 && rr == 0) ? 0u : 1u);
          tc_commit(slotempty_bar(slot));
          if (sub == 1) {
            tc_commit(actempty_bar(sa));
            if (last) tc_commit(accfull_bar);
          }
        }
        __syncwarp();
        if (++slot == NSLOT) { slot = 0; phs ^= 1; }
      }
      if (++sa == NA) { sa = 0; pha ^= 1; }
      if (last) ++seg;
      if (++ks == KS) ks = 0;
    }
  } else if (warp == 3) {
    (void)get_tmem_base();
  } else {
    // ================= dequant warps (smem packed -> registers -> TMEM), epilogue at the end of every tile segment ======
    const uint32_t tmem_base = get_tmem_base();
    const int qd = warp & 3;                        // TMEM lane quarter: channels 32 qd .. 32 qd + 31 of the tile
    const int p = (warp - kDecCtrlWarps) >> 2;      // which 64-k sub-slots of a stage this warp expands (DQW == 2)
    const int t64 = qd >> 1, jpair = qd & 1;
    const int c = lane >> 2;
    constexpr int WPT = BITS == 4 ? 2 : 4;          // packed words per thread per slab
    const uint32_t a_lane = tmem_base + ((uint32_t)(32 * qd) << 16);
    const uint8_t* wsm = smem_gen + (w_base - smem_base) + t64 * (Cfg::W_ROW_BYTES / 2) + lane * (4 * BITS) + jpair * (2 * BITS);
    const uint8_t* ssm = smem_gen + (sc_base - smem_base) + (64 * t64 + 8 * c + 4 * jpair) * 2;
    // epilogue identity of this thread: one output channel (TMEM lane), EW of its MPAD columns
    const int row = 32 * qd + lane;
    const int et = threadIdx.x - kDecCtrlWarps * 32;
    constexpr int EW = Cfg::EW;
    const int col0 = p * EW;
    const int m_valid = args.M < MPAD ? args.M : MPAD;
    bool waited = false;

    int tile = u_begin / KS, ks = u_begin - tile * KS;
    int s = 0, phw = 0, sa = 0, pha = 0, seg = 0, it_seg0 = 0;
    int slot = p % NSLOT, phs = 1;                  // slot of this warp's next sub-slot, parity of its "empty" wait
    for (int it = 0; it < n_units; ++it) {
      const bool last = (it == n_units - 1 || ks == KS - 1);
      mbar_wait(wfull_bar(s), phw);
      uint32_t s2[4] = {0, 0, 0, 0};                // broadcast scales of this thread's 4 channels: index 2 jj + b
      auto load_scales = [&](int srow) {
        const uint2 raw = *reinterpret_cast<const uint2*>(ssm + sa * Cfg::S_STAGE_BYTES + srow * 256);
        s2[0] = bcast16(raw.x);
        s2[1] = bcast16_hi(raw.x);
        s2[2] = bcast16(raw.y);
        s2[3] = bcast16_hi(raw.y);
      };
      if (GPS != 0) {
        mbar_wait(actfull_bar(sa), pha);            // the scale rows ride with the activations
        if (GPS == 1) load_scales(0);
      }
#pragma unroll
      for (int sub = p; sub < 2; sub += DQW) {
        uint32_t w[4][WPT];
#pragma unroll
        for (int rr = 0; rr < 4; ++rr) {
          const uint8_t* src = wsm + s * Cfg::W_STAGE_BYTES + (4 * sub + rr) * Cfg::W_ROW_BYTES;
          if (BITS == 4) {
            const uint2 v = *reinterpret_cast<const uint2*>(src);
            w[rr][0] = v.x; w[rr][1] = v.y;
          } else {
            const uint4 v = *reinterpret_cast<const uint4*>(src);
            w[rr][0] = v.x; w[rr][1] = v.y; w[rr][2] = v.z; w[rr][3] = v.w;
          }
        }
        if (sub + DQW >= 2) {
          // last read of this stage by this warp: hand it back to the producer (release orders the reads above)
          __syncwarp();
          if (lane == 0) mbar_arrive(wempty_bar(s));
        }
        if (GPS == 2) load_scales(sub);
        uint32_t ov[2][16];
#pragma unroll
        for (int rr = 0; rr < 4; ++rr) {
          if (GPS == 4 && (rr & 1) == 0) load_scales(2 * sub + (rr >> 1));
#pragma unroll
          for (int jj = 0; jj < 2; ++jj) {
            uint32_t* o4 = &ov[jj][4 * rr];
            if (BITS == 4) dequant4<T>(w[rr][jj], o4);
            else dequant8<T>(w[rr][2 * jj], w[rr][2 * jj + 1], o4);
            if (GPS != 0) {
              o4[0] = Pack2<T>::mul(o4[0], s2[2 * jj]);
              o4[1] = Pack2<T>::mul(o4[1], s2[2 * jj + 1]);
              o4[2] = Pack2<T>::mul(o4[2], s2[2 * jj]);
              o4[3] = Pack2<T>::mul(o4[3], s2[2 * jj + 1]);
            }
          }
        }
        mbar_wait(slotempty_bar(slot), phs);
        tc_fence_after();
        tmem_st_16x128b_x8(a_lane + slot * 32, ov[0]);
        tmem_st_16x128b_x8(a_lane + ((uint32_t)16 << 16) + slot * 32, ov[1]);
        tc_wait_st();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(slotfull_bar(slot));
        slot += DQW;
        if (slot >= NSLOT) { slot -= NSLOT; phs ^= 1; }
      }
      if (GPS != 0) {
        __syncwarp();
        if (lane == 0) mbar_arrive(actempty_bar(sa));   // scale rows consumed
      }
      if (++s == NW) { s = 0; phw ^= 1; }
      if (++sa == NA) { sa = 0; pha ^= 1; }

      if (last) {
        // ---------- epilogue of the segment [it_seg0, it] of `tile` ----------
        const int len = it - it_seg0 + 1;
        const int ks0 = ks - (len - 1);
        const bool full_k = (len == KS);
        const int n = tile * kTileN + row;
        const bool n_ok = n < args.N;
        if (!waited) { pdl_wait(); waited = true; }   // C, partial slots, counters may still be in use by the previous kernel
        float ch_scale = 1.f;
        if (GPS == 0 && n_ok)
          ch_scale = to_f32<T>(reinterpret_cast<const T*>(args.scales)[(n & ~31) + scale_single_pos(n & 31)]);
        T* cdst = reinterpret_cast<T*>(args.C) + n;
        float* part_dst = args.scratch + ((size_t)(blockIdx.x * 2 + (it_seg0 == 0 ? 0 : 1)) * MPAD) * kTileN + row;
        mbar_wait(accfull_bar, seg & 1);
        tc_fence_after();
        const uint32_t acc_src = a_lane + NSLOT * 32 + col0;
#pragma unroll
        for (int c0 = 0; c0 < EW; c0 += 16) {
          uint32_t t[16];
          tmem_ld_32x32b_x16(acc_src + c0, t);
          tc_wait_ld();
          if (c0 + 16 == EW) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(accempty_bar);
          }
          if (full_k) {
            if (n_ok) {
#pragma unroll
              for (int j = 0; j < 16; ++j)
                if (col0 + c0 + j < m_valid) cdst[(size_t)(col0 + c0 + j) * args.N] = from_f32<T>(__uint_as_float(t[j]) * ch_scale);
            }
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (col0 + c0 + j < m_valid) part_dst[(col0 + c0 + j) * kTileN] = __uint_as_float(t[j]);
          }
        }
        if (!full_k) {
          // last-arriver reduction: publish the partial, take a ticket; whoever holds the last one sums every piece of
          // the tile in CTA order (its own included, re-read from the slot: same value, fixed order -> deterministic)
          const int u_t0 = tile * KS;
          const int c_first = part.owner(u_t0), c_last = part.owner(u_t0 + KS - 1);
          __threadfence();
          asm volatile("bar.sync 1, %0;" ::"n"(Cfg::NDQ * 32) : "memory");
          if (et == 0) *ticket_p = atomicAdd(&args.counters[tile], 1);
          asm volatile("bar.sync 1, %0;" ::"n"(Cfg::NDQ * 32) : "memory");
          if (*ticket_p == c_last - c_first) {
            __threadfence();
            // slot of CTA cc's piece: its first segment (slot 0) unless cc is the CTA in which the tile begins and that
            // CTA's range started in an earlier tile
            const int first_bit = part.start(c_first) < u_t0 ? 1 : 0;
#pragma unroll
            for (int c0 = 0; c0 < EW; c0 += 16) {
              float acc[16];
#pragma unroll
              for (int j = 0; j < 16; ++j) acc[j] = 0.f;
              for (int cc = c_first; cc <= c_last; ++cc) {
                const float* src = args.scratch + ((size_t)(cc * 2 + (cc == c_first ? first_bit : 0)) * MPAD) * kTileN + row;
#pragma unroll
                for (int j = 0; j < 16; ++j)
                  if (col0 + c0 + j < m_valid) acc[j] += __ldcg(src + (col0 + c0 + j) * kTileN);
              }
              if (n_ok) {
#pragma unroll
                for (int j = 0; j < 16; ++j)
                  if (col0 + c0 + j < m_valid) cdst[(size_t)(col0 + c0 + j) * args.N] = from_f32<T>(acc[j] * ch_scale);
              }
            }
            if (et == 0) args.counters[tile] = 0;   // self-reset for the next launch
          }
          (void)ks0;
        }
        ++seg;
        it_seg0 = it + 1;
      }
      if (++ks == KS) { ks = 0; ++tile; }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<Cfg::TMEM_COLS>(*tmem_slot_p);
}

// =====================================================================================================================
// Fast path (4-bit, group 128 or channelwise): zero point and group scale leave the dequant warps.
//
// Measured on the machine above (tools/dec_trace.py, one CTA per SM): a dequant warp needs ~2100 clk per 128-k stage, of
// which ~560 are mbarrier waits that were already satisfied (~90-190 clk each), ~220 the two tcgen05.st + wait::st round
// trips, ~760 the expansion of 16 packed words (latency-bound: one warp per scheduler) -- per-warp latency, not any
// pipe, is the limit, so the cure is fewer hand-offs per byte, fewer instructions per word and more warps in flight:
//   * hand-offs are per 128-k STAGE (one 64-column TMEM slot), not per 64-k sub-slot, and there are two dequant warps
//     per TMEM lane quarter that take alternate stages: 8 dequant warps per CTA, 2 CTAs per SM;
//   * the weights go into TMEM WITH their bias, w' = 2^e + nibble (bf16: 128 + n, fp16: 1024 + n, both exact): a packed
//     word costs 4 LOP3 + 3 shifts (issued as IMAD.HI on the FMA pipe) and nothing else.  The missing
//     -(2^e + 8) * sum_k a[m,k] comes back in the epilogue from per-(row, group) activation sums, fp32
//     [K/128][stride], written by the kernel that produced the activations (or by act_group_sums_kernel);
//   * a stage is exactly one quantisation group at g128: every stage gets a fresh accumulator (two, ping-pong) and all
//     dequant warps -- they own the TMEM lane quarters -- read it back one stage later, each its half of the
//     columns, and apply scale and correction in fp32:  acc[m] += s[g,n] * (D_g[n,m] - bias * asum[g,m]).
// Rings, stream-K partition and last-arriver reduction are the machine above.
template <typename T> struct ZpcFmt;
template <> struct ZpcFmt<__nv_bfloat16> { static constexpr uint32_t EX = 0x43004300u; static constexpr float BIAS = 136.f; };
template <> struct ZpcFmt<__half> { static constexpr uint32_t EX = 0x64006400u; static constexpr float BIAS = 1032.f; };

#ifndef B200_DEC_NSHF_FMA
#define B200_DEC_NSHF_FMA 3      // how many of the three shifts go to the FMA pipe as IMAD.HI (rest: SHF on the ALU pipe)
#endif

template <typename T>
__device__ __forceinline__ void dequant4_biased(uint32_t q, uint32_t* o, const uint32_t (&shmul)[3]) {
  constexpr uint32_t MASK = 0x000f000f, EX = ZpcFmt<T>::EX;
  const uint32_t q4 = B200_DEC_NSHF_FMA >= 1 ? __umulhi(q, shmul[0]) : (q >> 4);
  const uint32_t q8 = B200_DEC_NSHF_FMA >= 2 ? __umulhi(q, shmul[1]) : (q >> 8);
  const uint32_t q12 = B200_DEC_NSHF_FMA >= 3 ? __umulhi(q, shmul[2]) : (q >> 12);
  asm("lop3.b32 %0, %1, %2, %3, 0xea;" : "=r"(o[0]) : "r"(q), "n"(MASK), "n"(EX));     // p0,p4
  asm("lop3.b32 %0, %1, %2, %3, 0xea;" : "=r"(o[2]) : "r"(q4), "n"(MASK), "n"(EX));    // p1,p5
  asm("lop3.b32 %0, %1, %2, %3, 0xea;" : "=r"(o[1]) : "r"(q8), "n"(MASK), "n"(EX));    // p2,p6
  asm("lop3.b32 %0, %1, %2, %3, 0xea;" : "=r"(o[3]) : "r"(q12), "n"(MASK), "n"(EX));   // p3,p7
}

__device__ __forceinline__ void tmem_ld_32x32b_x8(uint32_t taddr, uint32_t (&r)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr)
               : "memory");
}

#ifdef B200_DEC_TRACE
#define DTRACE(role, it, ev) if (args.trace && (int)blockIdx.x == args.trace_cta && lane == 0 && (it) < 32) { args.trace[((role) * 32 + (it)) * 16 + (ev)] = clock64(); }
#else
#define DTRACE(role, it, ev)
#endif

template <int MPAD>
struct Dec3Cfg {
  static constexpr int NDQ = 8;                                              // dequant warps: 2 per TMEM lane quarter
  static constexpr int THREADS = 32 * (kDecCtrlWarps + NDQ);
  static constexpr int EW = MPAD / 2;                                        // accumulator columns per epilogue thread
  static constexpr int EC = EW < 16 ? EW : 16;                               // columns per TMEM read
  static constexpr int TMEM_COLS = 256;                                      // [2 x 64-column operand slots][2 x MPAD accumulators]
  static constexpr int OCC = 2;
  static constexpr int ACC_COL0 = 128;
  static constexpr int W_ROW_BYTES = kTileN * 16 * 4 / 8;
  static constexpr int W_STAGE_BYTES = kDecSlabs * W_ROW_BYTES;              // 8 KB
  static constexpr int A_BOX_BYTES = MPAD * 128;
  static constexpr int A_STAGE_BYTES = 2 * A_BOX_BYTES;
  static constexpr int X_STAGE_BYTES = 512;                                  // [scale row 256 B][activation sums MPAD x 4 B]
  static constexpr int NA = MPAD <= 32 ? 4 : 3;
  static constexpr int BUDGET = (227 * 1024) / OCC - 1024 - 1024 - 512;
  static constexpr int NW_RAW = (BUDGET - NA * (A_STAGE_BYTES + X_STAGE_BYTES)) / W_STAGE_BYTES;
  static constexpr int NW = NW_RAW > 10 ? 10 : NW_RAW;
  static constexpr int NBARS = 2 * NW + 2 * NA + 8;
  static constexpr int SMEM_BYTES = 1024 + NA * (A_STAGE_BYTES + X_STAGE_BYTES) + NW * W_STAGE_BYTES + NBARS * 8 + 16;
  static_assert(NW >= 3, "weight ring");
  static_assert(NBARS * 8 + 16 <= 512, "barrier area");
  static_assert(ACC_COL0 + 2 * MPAD <= TMEM_COLS, "TMEM budget");
};

template <typename T, int MPAD, int GPS>   // GPS: 0 channelwise, 1 group 128
__global__ void __launch_bounds__(Dec3Cfg<MPAD>::THREADS, Dec3Cfg<MPAD>::OCC)
marlin_decode3_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_w,
                      const __grid_constant__ CUtensorMap tmap_s, const GemmArgs args) {
  using Cfg = Dec3Cfg<MPAD>;
  constexpr int NW = Cfg::NW, NA = Cfg::NA;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  // layout: [activation stages (1024-aligned)] [weight stages] [scale row + activation sums per stage] [barriers]
  const uint32_t act_base = smem_base;
  const uint32_t w_base = act_base + NA * Cfg::A_STAGE_BYTES;
  const uint32_t x_base = w_base + NW * Cfg::W_STAGE_BYTES;
  const uint32_t bar_base = x_base + NA * Cfg::X_STAGE_BYTES;
  constexpr int B_WE = NW, B_AF = 2 * NW, B_AE = B_AF + NA, B_SF = B_AE + NA, B_SE = B_SF + 2, B_ACF = B_SE + 2, B_ACE = B_ACF + 2;
  auto bar = [&](int i) { return bar_base + 8u * i; };
  const uint32_t tmem_slot = bar_base + 8u * Cfg::NBARS;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  volatile uint32_t* tmem_slot_p = reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_slot - smem_base));
  volatile int* ticket_p = reinterpret_cast<volatile int*>(smem_gen + (tmem_slot - smem_base) + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    prefetch_tmap(&tmap_w);
    prefetch_tmap(&tmap_a);
    if (GPS != 0) prefetch_tmap(&tmap_s);
  }
  if (threadIdx.x >= 32 && threadIdx.x < 32 + Cfg::NBARS) {
    const int i = threadIdx.x - 32;
    uint32_t cnt = 1;                                            // wfull, actfull, slotempty, accfull
    if (i >= B_WE && i < B_AF) cnt = 4;                          // wempty: the four lane-quarter warps that own the stage
    else if (i >= B_AE && i < B_SF) cnt = 1 + Cfg::NDQ;          // actempty: issuer commit + every epilogue warp (scale, sums)
    else if (i >= B_SF && i < B_SE) cnt = 4;                     // slotfull: the four lane-quarter warps
    else if (i >= B_ACE) cnt = Cfg::NDQ;                         // accempty: every epilogue warp
    mbar_init(bar(i), cnt);
    fence_barrier_init();
  }
  __syncthreads();
  pdl_launch_dependents();

  const int KS = args.KS;
  const Partition part(args.total_units, (int)gridDim.x);
  const int u_begin = part.start((int)blockIdx.x);
  const int u_end = part.start((int)blockIdx.x + 1);
  const int n_units = u_end - u_begin;

  auto get_tmem_base = [&]() -> uint32_t {
    tc_fence_before();
    asm volatile("bar.sync 2, %0;" ::"n"(Cfg::THREADS - 64) : "memory");
    tc_fence_after();
    return *tmem_slot_p;
  };

  if (warp == 0) {
    // ================= weight producer (no PDL wait: weights are static) ============================================
    const uint64_t pol_w = policy_evict_first();
    int tile = u_begin / KS, ks = u_begin - tile * KS;
    int s = 0, ph = 1;
    for (int it = 0; it < n_units; ++it) {
      mbar_wait(bar(B_WE + s), ph);
      if (elect_one()) {
        mbar_arrive_expect_tx(bar(s), Cfg::W_STAGE_BYTES);
        tma_2d_g2s_hint(w_base + s * Cfg::W_STAGE_BYTES, &tmap_w, tile * (Cfg::W_ROW_BYTES / 8), ks * kDecSlabs, bar(s), pol_w);
      }
      __syncwarp();
      if (++s == NW) { s = 0; ph ^= 1; }
      if (++ks == KS) { ks = 0; ++tile; }
    }
  } else if (warp == 1) {
    // ================= activations + group scale row + activation sums =============================================
    const uint64_t pol_a = policy_evict_last();
    constexpr uint32_t tx = Cfg::A_STAGE_BYTES + (GPS != 0 ? 256 : 0) + MPAD * 4;
    int tile = u_begin / KS, ks = u_begin - tile * KS;
    int s = 0, ph = 1;
    pdl_wait();   // activations and their sums come from the previous kernels of the stream
    for (int it = 0; it < n_units; ++it) {
      mbar_wait(bar(B_AE + s), ph);
      if (elect_one()) {
        mbar_arrive_expect_tx(bar(B_AF + s), tx);
        tma_3d_g2s_hint(act_base + s * Cfg::A_STAGE_BYTES, &tmap_a, 0, 0, ks * 2, bar(B_AF + s), pol_a);
        if (GPS != 0) tma_2d_g2s_hint(x_base + s * Cfg::X_STAGE_BYTES, &tmap_s, tile * kTileN, ks, bar(B_AF + s), pol_a);
        bulk_g2s(x_base + s * Cfg::X_STAGE_BYTES + 256, args.asum + (size_t)ks * args.asum_stride, MPAD * 4, bar(B_AF + s));
      }
      __syncwarp();
      if (++s == NA) { s = 0; ph ^= 1; }
      if (++ks == KS) { ks = 0; ++tile; }
    }
  } else if (warp == 2) {
    // ================= TMEM allocator + MMA issuer: stage `it` -> operand slot it & 1 -> accumulator it & 1 ============
    tmem_alloc<Cfg::TMEM_COLS>(tmem_slot);
    const uint32_t tmem_base = get_tmem_base();
    constexpr uint32_t idesc = make_idesc(1 /*F32*/, Pack2<T>::FMT, Pack2<T>::FMT, 128, MPAD);
    int sa = 0, pha = 0;
    for (int it = 0; it < n_units; ++it) {
      const int b = it & 1, phb = (it >> 1) & 1;
      const uint32_t d_tmem = tmem_base + Cfg::ACC_COL0 + b * MPAD;
      DTRACE(1, it, 0);
      mbar_wait(bar(B_AF + sa), pha);
      DTRACE(1, it, 1);
      mbar_wait(bar(B_ACE + b), phb ^ 1);      // the epilogue has read this accumulator (two stages ago)
      DTRACE(1, it, 2);
      mbar_wait(bar(B_SF + b), phb);
      DTRACE(1, it, 3);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t b_smem = act_base + sa * Cfg::A_STAGE_BYTES;
#pragma unroll
        for (int r = 0; r < 8; ++r)
          mma_f16_ts(d_tmem, tmem_base + b * 64 + r * 8,
                     make_kmajor_sw128_desc(b_smem + (r >> 2) * Cfg::A_BOX_BYTES + (r & 3) * 32), idesc, r == 0 ? 0u : 1u);
        tc_commit(bar(B_SE + b));
        tc_commit(bar(B_AE + sa));
        tc_commit(bar(B_ACF + b));
      }
      __syncwarp();
      DTRACE(1, it, 4);
      if (++sa == NA) { sa = 0; pha ^= 1; }
    }
  } else if (warp == 3) {
    (void)get_tmem_base();
  } else {
    // ================= dequant warps; they also scale / correct / drain the accumulators ================================
    const uint32_t tmem_base = get_tmem_base();
    const int qd = warp & 3;                        // TMEM lane quarter: channels 32 qd .. 32 qd + 31 of the tile
    const int p = (warp - kDecCtrlWarps) >> 2;      // dequantises the stages with (it & 1) == p; epilogue columns p * EW ..
    const int t64 = qd >> 1, jpair = qd & 1;
    const uint32_t a_lane = tmem_base + ((uint32_t)(32 * qd) << 16);
    const uint8_t* wsm = smem_gen + (w_base - smem_base) + t64 * (Cfg::W_ROW_BYTES / 2) + lane * 16 + jpair * 8;
    const uint8_t* xsm = smem_gen + (x_base - smem_base);
    const int row = 32 * qd + lane;
    const int et = threadIdx.x - kDecCtrlWarps * 32;
    constexpr int EW = Cfg::EW, EC = Cfg::EC;
    const int col0 = p * EW;
    const int m_valid = args.M < MPAD ? args.M : MPAD;
    const int sc_pos = 64 * (row >> 6) + 8 * (row & 7) + ((row & 63) >> 3);   // scale_perm: pos 8 c + b' <-> col c + 8 b'
    const uint32_t shmul[3] = {args.shmul[0], args.shmul[1], args.shmul[2]};
    bool waited = false;

    float facc[EW];
#pragma unroll
    for (int j = 0; j < EW; ++j) facc[j] = 0.f;

    // epilogue of stage jt (one quantisation group): facc += scale * (D - bias * asum); releases accumulator and stage
    auto epi = [&](int jt) {
      const int b = jt & 1, sj = jt % NA;
      if (p == 0 && qd == 0) DTRACE(0, jt, 8);
      mbar_wait(bar(B_ACF + b), (jt >> 1) & 1);
      if (p == 0 && qd == 0) DTRACE(0, jt, 9);
      tc_fence_after();
      const uint8_t* x = xsm + sj * Cfg::X_STAGE_BYTES;
      float sc = 1.f;
      if (GPS != 0) sc = to_f32<T>(*reinterpret_cast<const T*>(x + 2 * sc_pos));
#pragma unroll
      for (int c0 = 0; c0 < EW; c0 += EC) {
        uint32_t t[EC];
        if constexpr (EC == 16) tmem_ld_32x32b_x16(a_lane + Cfg::ACC_COL0 + b * MPAD + col0 + c0, t);
        else tmem_ld_32x32b_x8(a_lane + Cfg::ACC_COL0 + b * MPAD + col0 + c0, t);
        float as[EC];
#pragma unroll
        for (int j = 0; j < EC; j += 4) {
          const float4 v = *reinterpret_cast<const float4*>(x + 256 + (col0 + c0 + j) * 4);
          as[j] = v.x; as[j + 1] = v.y; as[j + 2] = v.z; as[j + 3] = v.w;
        }
        tc_wait_ld();
        if (c0 + EC == EW) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) { mbar_arrive(bar(B_ACE + b)); mbar_arrive(bar(B_AE + sj)); }   // accumulator, scale row and sums released
        }
#pragma unroll
        for (int j = 0; j < EC; ++j)
          facc[c0 + j] = fmaf(sc, fmaf(-ZpcFmt<T>::BIAS, as[j], __uint_as_float(t[j])), facc[c0 + j]);
      }
      if (p == 0 && qd == 0) DTRACE(0, jt, 10);
    };

    int tile = u_begin / KS, ks = u_begin - tile * KS;
    int s = 0, phw = 0, it_seg0 = 0;
    for (int it = 0; it < n_units; ++it) {
      const bool last = (it == n_units - 1 || ks == KS - 1);
      if ((it & 1) == p) {
        if (qd == 0) DTRACE(0, it, 0);
        mbar_wait(bar(s), phw);
        if (qd == 0) DTRACE(0, it, 1);
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          uint32_t w[4][2];
#pragma unroll
          for (int rr = 0; rr < 4; ++rr) {
            const uint2 v = *reinterpret_cast<const uint2*>(wsm + s * Cfg::W_STAGE_BYTES + (4 * half + rr) * Cfg::W_ROW_BYTES);
            w[rr][0] = v.x; w[rr][1] = v.y;
          }
          if (half == 1) {
            __syncwarp();
            if (lane == 0) mbar_arrive(bar(B_WE + s));   // packed words are in registers: stage back to the producer
          }
#pragma unroll
          for (int jj = 0; jj < 2; ++jj) {
            uint32_t ov[16];
#pragma unroll
            for (int rr = 0; rr < 4; ++rr) dequant4_biased<T>(w[rr][jj], &ov[4 * rr], shmul);
            if (half == 0 && jj == 0) {
              if (qd == 0) DTRACE(0, it, 2);
              mbar_wait(bar(B_SE + p), ((it >> 1) & 1) ^ 1);   // the MMAs of stage it - 2 have read this slot
              if (qd == 0) DTRACE(0, it, 3);
              tc_fence_after();
            }
            tmem_st_16x128b_x8(a_lane + ((uint32_t)(16 * jj) << 16) + p * 64 + half * 32, ov);
          }
        }
        if (qd == 0) DTRACE(0, it, 4);
        tc_wait_st();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar(B_SF + p));
        if (qd == 0) DTRACE(0, it, 5);
      }
      if (++s == NW) { s = 0; phw ^= 1; }
      if (it > it_seg0) epi(it - 1);

      if (last) {
        // ---------- end of the segment [it_seg0, it] of `tile` ----------
        epi(it);
        const int len = it - it_seg0 + 1;
        const bool full_k = (len == KS);
        const int n = tile * kTileN + row;
        const bool n_ok = n < args.N;
        if (!waited) { pdl_wait(); waited = true; }   // C, partial slots, counters may still be in use by the previous kernel
        float ch_scale = 1.f;
        if (GPS == 0 && n_ok)
          ch_scale = to_f32<T>(reinterpret_cast<const T*>(args.scales)[(n & ~31) + scale_single_pos(n & 31)]);
        T* cdst = reinterpret_cast<T*>(args.C) + n;
        if (full_k) {
          if (n_ok) {
#pragma unroll
            for (int j = 0; j < EW; ++j)
              if (col0 + j < m_valid) cdst[(size_t)(col0 + j) * args.N] = from_f32<T>(facc[j] * ch_scale);
          }
        } else {
          // last-arriver reduction: publish the partial, take a ticket; whoever holds the last one sums every piece of
          // the tile in CTA order (its own included, re-read from the slot: same value, fixed order -> deterministic)
          float* part_dst = args.scratch + ((size_t)(blockIdx.x * 2 + (it_seg0 == 0 ? 0 : 1)) * MPAD) * kTileN + row;
#pragma unroll
          for (int j = 0; j < EW; ++j)
            if (col0 + j < m_valid) part_dst[(col0 + j) * kTileN] = facc[j];
          const int u_t0 = tile * KS;
          const int c_first = part.owner(u_t0), c_last = part.owner(u_t0 + KS - 1);
          __threadfence();
          asm volatile("bar.sync 1, %0;" ::"n"(Cfg::NDQ * 32) : "memory");
          if (et == 0) *ticket_p = atomicAdd(&args.counters[tile], 1);
          asm volatile("bar.sync 1, %0;" ::"n"(Cfg::NDQ * 32) : "memory");
          if (*ticket_p == c_last - c_first) {
            __threadfence();
            // slot of CTA cc's piece: its first segment (slot 0) unless cc is the CTA in which the tile begins and that
            // CTA's range started in an earlier tile
            const int first_bit = part.start(c_first) < u_t0 ? 1 : 0;
            float acc[EW];
#pragma unroll
            for (int j = 0; j < EW; ++j) acc[j] = 0.f;
            for (int cc = c_first; cc <= c_last; ++cc) {
              const float* src = args.scratch + ((size_t)(cc * 2 + (cc == c_first ? first_bit : 0)) * MPAD) * kTileN + row;
#pragma unroll
              for (int j = 0; j < EW; ++j)
                if (col0 + j < m_valid) acc[j] += __ldcg(src + (col0 + j) * kTileN);
            }
            if (n_ok) {
#pragma unroll
              for (int j = 0; j < EW; ++j)
                if (col0 + j < m_valid) cdst[(size_t)(col0 + j) * args.N] = from_f32<T>(acc[j] * ch_scale);
            }
            if (et == 0) args.counters[tile] = 0;   // self-reset for the next launch
          }
        }
#pragma unroll
        for (int j = 0; j < EW; ++j) facc[j] = 0.f;
        it_seg0 = it + 1;
      }
      if (++ks == KS) { ks = 0; ++tile; }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<Cfg::TMEM_COLS>(*tmem_slot_p);
}

// per-(row, group of 128 k) sums of the activations, fp32 [K/128][stride]: one warp per (row, group)
template <typename T>
__global__ void act_group_sums_kernel(float* __restrict__ out, int stride, const T* __restrict__ a, int M, int G, int64_t lda) {
  pdl_launch_dependents();
  pdl_wait();
  const int wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (wid >= M * G) return;
  const int m = wid / G, g = wid - m * G;
  const uint2 v = *reinterpret_cast<const uint2*>(a + (size_t)m * lda + g * 128 + lane * 4);
  const T* e = reinterpret_cast<const T*>(&v);
  float s = (to_f32<T>(e[0]) + to_f32<T>(e[1])) + (to_f32<T>(e[2]) + to_f32<T>(e[3]));
  s = warp_sum(s);
  if (lane == 0) out[(size_t)g * stride + m] = s;
}

// ---- host side --------------------------------------------------------------------------------------------------
template <typename T, int MPAD, int BITS, int GPS>
int launch_decode(const GemmArgs& a_in, const void* A, int64_t lda, cudaStream_t st) {
  using Cfg = DecCfg<MPAD, BITS>;
  GemmArgs a = a_in;
  a.m_chunks = 1;
  a.KS = a.K / kDecStageK;
  a.total_units = a.n_tiles * a.KS;
  auto kern = marlin_decode_kernel<T, MPAD, BITS, GPS>;
  int dev = 0;
  cudaGetDevice(&dev);
  static int occ[64] = {};
  if (occ[dev & 63] == 0) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
    cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    int n = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kern, Cfg::THREADS, Cfg::SMEM_BYTES) != cudaSuccess || n < 1)
      return set_error(B200_ECUDA, "marlin decode gemm: kernel does not fit on an SM (smem %d, threads %d)", Cfg::SMEM_BYTES, Cfg::THREADS);
    occ[dev & 63] = n < Cfg::OCC ? n : Cfg::OCC;
    if (getenv("B200_DEC_VERBOSE"))
      fprintf(stderr, "[dec1] MPAD=%d BITS=%d GPS=%d: occupancy %d (want %d), smem %d, threads %d\n", MPAD, BITS, GPS, n, Cfg::OCC,
              Cfg::SMEM_BYTES, Cfg::THREADS);
  }
  // grid: below the number of resident slots, so that the next kernel of the stream finds a free slot on every SM while
  // this one is still running; at least `min_units` stages per CTA so that split-K partial traffic stays small
  static const int env_cps = getenv("B200_DEC_CTAS_PER_SM") ? atoi(getenv("B200_DEC_CTAS_PER_SM")) : 0;
  static const int env_minu = getenv("B200_DEC_MIN_UNITS") ? atoi(getenv("B200_DEC_MIN_UNITS")) : 0;
  int cps = env_cps > 0 ? env_cps : Cfg::OCC;
  const int min_units = env_minu > 0 ? env_minu : (MPAD >= 64 ? 4 : 2);
  int grid = num_sms() * cps;
  if (grid > kDecMaxGrid) grid = kDecMaxGrid;
  if (grid > a.total_units / min_units) grid = a.total_units / min_units;
  if (grid < 1) grid = 1;
  if (int e = dec_scratch(dev, st, &a.scratch, &a.counters)) return e;
  CUtensorMap tmap_a, tmap_w, tmap_s;
  const CUtensorMapDataType dt = std::is_same<T, __half>::value ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  bool ok = make_tmap_act3d(&tmap_a, A, dt, (uint64_t)a.M, (uint64_t)a.K, (uint64_t)lda * 2, MPAD, kDecStageK / 64);
  const uint64_t wrow_bytes = (uint64_t)a.N * 16 * BITS / 8;
  ok = ok && make_tmap_2d(&tmap_w, a.B, CU_TENSOR_MAP_DATA_TYPE_UINT64, 8, (uint64_t)a.K / 16, wrow_bytes / 8, wrow_bytes,
                          kDecSlabs, Cfg::W_ROW_BYTES / 8, CU_TENSOR_MAP_SWIZZLE_NONE);
  if (GPS != 0)
    ok = ok && make_tmap_2d(&tmap_s, a.scales, dt, 2, (uint64_t)a.num_groups, (uint64_t)a.N, (uint64_t)a.N * 2,
                            (uint32_t)GPS, kTileN, CU_TENSOR_MAP_SWIZZLE_NONE);
  else
    tmap_s = tmap_a;
  if (!ok)
    return set_error(B200_ECUDA, "marlin decode gemm: cuTensorMapEncodeTiled failed (A=%p M=%d K=%d N=%d lda=%lld)", A, a.M,
                     a.K, a.N, (long long)lda);
  a.prof = nullptr; a.trace = nullptr; a.trace_cta = 0; a.exp = 0; a.cluster_s = 0;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(Cfg::THREADS);
  cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  int na = 0;
  if (pdl_enabled()) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, tmap_a, tmap_w, tmap_s, a);
  if (e != cudaSuccess) return set_error(B200_ECUDA, "marlin decode gemm launch: %s", cudaGetErrorString(e));
  return B200_OK;
}

int dec_asum_scratch(int dev, cudaStream_t st, float** asum);   // marlin_gemm.cu: fp32 [kDecMaxGroups][64]
constexpr int kDecMaxGroups = 8192;

template <typename T, int MPAD, int GPS>
int launch_decode3(const GemmArgs& a_in, const void* A, int64_t lda, cudaStream_t st) {
  using Cfg = Dec3Cfg<MPAD>;
  GemmArgs a = a_in;
  a.m_chunks = 1;
  a.KS = a.K / kDecStageK;
  a.total_units = a.n_tiles * a.KS;
  auto kern = marlin_decode3_kernel<T, MPAD, GPS>;
  int dev = 0;
  cudaGetDevice(&dev);
  static bool attr_set[64] = {};
  if (!attr_set[dev & 63]) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
    cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    attr_set[dev & 63] = true;
    if (getenv("B200_DEC_VERBOSE")) {
      int n = 0;
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kern, Cfg::THREADS, Cfg::SMEM_BYTES);
      fprintf(stderr, "[dec3] MPAD=%d GPS=%d: occupancy query %d (sized for %d), smem %d, threads %d, NW=%d NA=%d\n", MPAD, GPS, n,
              Cfg::OCC, Cfg::SMEM_BYTES, Cfg::THREADS, Cfg::NW, Cfg::NA);
    }
  }
  static const int env_cps = getenv("B200_DEC_CTAS_PER_SM") ? atoi(getenv("B200_DEC_CTAS_PER_SM")) : 0;
  static const int env_minu = getenv("B200_DEC_MIN_UNITS") ? atoi(getenv("B200_DEC_MIN_UNITS")) : 0;
  const int cps = env_cps > 0 ? env_cps : Cfg::OCC;
  const int min_units = env_minu > 0 ? env_minu : (MPAD >= 64 ? 4 : 2);
  int grid = num_sms() * cps;
  if (grid > kDecMaxGrid) grid = kDecMaxGrid;
  if (grid > a.total_units / min_units) grid = a.total_units / min_units;
  if (grid < 1) grid = 1;
  if (int e = dec_scratch(dev, st, &a.scratch, &a.counters)) return e;
  if (a.asum == nullptr) {
    // the caller did not bring the activation group sums: one small kernel ahead of the GEMM computes them
    float* asum = nullptr;
    if (int e = dec_asum_scratch(dev, st, &asum)) return e;
    const int G = a.K / 128;
    const int warps = a.M * G;
    cudaError_t e = launch_pdl(act_group_sums_kernel<T>, dim3((unsigned)((warps + 7) / 8)), dim3(256), 0, st, asum, 64,
                               (const T*)A, a.M, G, lda);
    if (e != cudaSuccess) return set_error(B200_ECUDA, "act_group_sums launch: %s", cudaGetErrorString(e));
    a.asum = asum;
    a.asum_stride = 64;
  }
  a.shmul[0] = 1u << 28; a.shmul[1] = 1u << 24; a.shmul[2] = 1u << 20;
  CUtensorMap tmap_a, tmap_w, tmap_s;
  const CUtensorMapDataType dt = std::is_same<T, __half>::value ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  bool ok = make_tmap_act3d(&tmap_a, A, dt, (uint64_t)a.M, (uint64_t)a.K, (uint64_t)lda * 2, MPAD, kDecStageK / 64);
  const uint64_t wrow_bytes = (uint64_t)a.N * 8;
  ok = ok && make_tmap_2d(&tmap_w, a.B, CU_TENSOR_MAP_DATA_TYPE_UINT64, 8, (uint64_t)a.K / 16, wrow_bytes / 8, wrow_bytes,
                          kDecSlabs, Cfg::W_ROW_BYTES / 8, CU_TENSOR_MAP_SWIZZLE_NONE);
  if (GPS != 0)
    ok = ok && make_tmap_2d(&tmap_s, a.scales, dt, 2, (uint64_t)a.num_groups, (uint64_t)a.N, (uint64_t)a.N * 2, 1, kTileN,
                            CU_TENSOR_MAP_SWIZZLE_NONE);
  else
    tmap_s = tmap_a;
  if (!ok)
    return set_error(B200_ECUDA, "marlin decode gemm: cuTensorMapEncodeTiled failed (A=%p M=%d K=%d N=%d lda=%lld)", A, a.M,
                     a.K, a.N, (long long)lda);
  a.prof = nullptr; a.trace = nullptr; a.trace_cta = 0; a.exp = 0; a.cluster_s = 0;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(Cfg::THREADS);
  cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  int na = 0;
  if (pdl_enabled()) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
#ifdef B200_DEC_TRACE
  if (getenv("B200_DEC_TRACE")) {
    a.trace_cta = atoi(getenv("B200_DEC_TRACE"));
    cudaMallocManaged(&a.trace, sizeof(long long) * 2 * 32 * 16);
    cudaMemset(a.trace, 0, sizeof(long long) * 2 * 32 * 16);
  }
#endif
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, tmap_a, tmap_w, tmap_s, a);
  if (e != cudaSuccess) return set_error(B200_ECUDA, "marlin decode gemm launch: %s", cudaGetErrorString(e));
#ifdef B200_DEC_TRACE
  if (a.trace) {
    cudaDeviceSynchronize();
    long long t0 = 0;
    for (int i = 0; i < 2 * 32 * 16; ++i) if (a.trace[i] && (t0 == 0 || a.trace[i] < t0)) t0 = a.trace[i];
    for (int r = 0; r < 2; ++r)
      for (int i = 0; i < 32; ++i) {
        bool any = false;
        for (int ev = 0; ev < 16; ++ev) any = any || a.trace[(r * 32 + i) * 16 + ev];
        if (!any) continue;
        fprintf(stderr, "[dec trace] %s it=%2d:", r ? "issuer " : "dequant", i);
        for (int ev = 0; ev < 12; ++ev) fprintf(stderr, " %6lld", a.trace[(r * 32 + i) * 16 + ev] ? a.trace[(r * 32 + i) * 16 + ev] - t0 : -1);
        fprintf(stderr, "\n");
      }
    cudaFree(a.trace);
  }
#endif
  return B200_OK;
}

template <typename T>
int dispatch_decode3(const GemmArgs& a, const void* A, int64_t lda, cudaStream_t st) {
  const bool g128 = a.num_groups > 1;
  if (a.M <= 16) return g128 ? launch_decode3<T, 16, 1>(a, A, lda, st) : launch_decode3<T, 16, 0>(a, A, lda, st);
  if (a.M <= 32) return g128 ? launch_decode3<T, 32, 1>(a, A, lda, st) : launch_decode3<T, 32, 0>(a, A, lda, st);
  return g128 ? launch_decode3<T, 64, 1>(a, A, lda, st) : launch_decode3<T, 64, 0>(a, A, lda, st);
}

// M <= 64, no fused all-reduce.  gps_code: scale rows per 128-k stage (0 channelwise, 1, 2, 4)
template <typename T, int BITS>
int dispatch_decode(const GemmArgs& a, const void* A, int64_t lda, cudaStream_t st) {
  const int gps = a.num_groups > 1 ? kDecStageK / a.group_size : 0;
  static const int force_v1 = getenv("B200_DEC_V1") ? atoi(getenv("B200_DEC_V1")) : 0;
  if (BITS == 4 && gps <= 1 && !force_v1 && a.K / 128 <= kDecMaxGroups) return dispatch_decode3<T>(a, A, lda, st);
#define B200_DEC_CASE(G)                                                        \
  if (gps == G) {                                                               \
    if (a.M <= 16) return launch_decode<T, 16, BITS, G>(a, A, lda, st);         \
    if (a.M <= 32) return launch_decode<T, 32, BITS, G>(a, A, lda, st);         \
    return launch_decode<T, 64, BITS, G>(a, A, lda, st);                        \
  }
  B200_DEC_CASE(0)
  B200_DEC_CASE(1)
  B200_DEC_CASE(2)
  B200_DEC_CASE(4)
#undef B200_DEC_CASE
  return set_error(B200_EINVAL, "marlin decode gemm: unexpected group size %d", a.group_size);
}
int dispatch_decode_f16_4(const GemmArgs& a, const void* A, int64_t lda, cudaStream_t st);
int dispatch_decode_f16_8(const GemmArgs& a, const void* A, int64_t lda, cudaStream_t st);
int dispatch_decode_bf16_4(const GemmArgs& a, const void* A, int64_t lda, cudaStream_t st);
int dispatch_decode_bf16_8(const GemmArgs& a, const void* A, int64_t lda, cudaStream_t st);

}  // namespace b200
