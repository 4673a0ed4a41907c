// Marlin-layout W4A16 / W8A16 GEMM for sm_100a:  C[M,N] = A[M,K] . (s (.) (Q - 2^(bits-1)))
//
// Replaces gptq_marlin_gemm / marlin_gemm of the reference (csrc/quantization/gptq_marlin/gptq_marlin.cu:410-1363
// kernel, :1577-1731 launcher, :1735-1868 op; csrc/quantization/marlin/dense/marlin_cuda_kernel.cu).  The STORED
// layouts are the reference's (qweight [K/16, N*16/pf] int32 in 16x64 mma-fragment order, scales permuted by
// scale_perm / scale_perm_single); everything else is a different machine:
//
//   * swap-AB: the 128 weight rows (output channels) are the M=128 side of a tcgen05.mma, the skinny activation
//     batch (padded to MPAD in {16,32,64,128}) is the N side.  Decode M=1..64 therefore fills the tensor core.
//   * the dequantised weight operand never touches shared memory: packed int4 tiles arrive by tensor-map TMA
//     (one 16 KB request = 16 k16-slabs x 1 KB of a 128-column tile), 16 dequant warps expand them in registers
//     (lop3 magic-number trick, exact zero-point subtraction, group scale in the activation dtype like the
//     reference :270-278) and write them with tcgen05.st.16x128b.x2 straight into TMEM.  The Marlin fragment order
//     (thread l holds k = {2q,2q+1,2q+8,2q+9}, n = c, c+8 with c = l/4, q = l%4) IS the register layout of that
//     store shape, so no shuffle is needed and K stays in natural order in TMEM.
//   * activations come by one 3-D tensor-map request per stage (128-byte swizzle, zero-filled rows beyond M) and are
//     the B operand from shared memory; accumulators live in TMEM (double buffered), read back by 4 epilogue warps.
//   * one thread can only issue a TMA request every ~200-500 cycles (measured, tools/tma_bw2.cu), so THREE producer
//     warps take the pipeline stages round-robin and a stage is 256 k deep (16 KB of weights).
//   * work is split stream-K style: the (tile, 256-deep k-stage) space is cut into one contiguous range per CTA
//     (grid = #SMs).  Tiles whose K range is shared by several CTAs are reduced deterministically: every
//     contributor writes an fp32 partial to its own scratch slot, bumps the tile counter, and the last arriver adds
//     the slots in CTA order with 16-byte loads (no float atomics, no spin locks; counters reset themselves).
//   * programmatic dependent launch: the kernel is launched with the stream-serialization attribute and only waits
//     for its predecessor (griddepcontrol.wait) before touching activations / outputs / scratch, so barrier setup,
//     TMEM allocation and the first ring of weight + scale requests overlap the previous kernel of the decode step.
//
// Warp roles (768 threads): w0-2 TMA producers | w3 TMEM allocator + MMA issuer | w4-19 dequant (TMEM lane quarter =
// warp%4, warp j of a quarter owns half-stages h % 4 == j) | w20-23 epilogue (quarter = warp%4).
#include "common.cuh"
#include "ptx.cuh"
#include "tma_host.cuh"
#include <stdlib.h>

namespace b200 {
using namespace ptx;

constexpr int kTileN = 128;
constexpr int kStageK = 256;                // one pipeline stage / scheduling unit = 256 (k) x 128 (n) weights
constexpr int kHalfK = 128;                 // dequant / TMEM granularity: half a stage
constexpr int kSlabs = kHalfK / 16;         // 16-row k slabs per half stage
constexpr int kGemmThreads = 768;
constexpr int kNumProducers = 3;
constexpr int kMmaWarp = 3;
constexpr int kDequantWarp0 = 4;
constexpr int kNumDequantWarps = 16;        // 4 per TMEM lane quarter
constexpr int kEpiWarp0 = kDequantWarp0 + kNumDequantWarps;  // 20..23 (warp % 4 == quarter)
constexpr int kNA = 4;                      // TMEM ring of dequantised half stages (one per dequant warp of a quarter)
constexpr int kAColsPerHalf = kHalfK / 2;   // 64 columns of packed 16-bit pairs
constexpr int kAccCol0 = kNA * kAColsPerHalf;  // 256

template <int MPAD, int BITS>
struct GemmCfg {
  static constexpr int W_ROW_BYTES = kTileN * 16 * BITS / 8;         // one 16(k) x 128(n) slab: 1024 (4-bit) / 2048 (8-bit)
  static constexpr int W_HALF_BYTES = kSlabs * W_ROW_BYTES;          // 8192 / 16384
  static constexpr int W_STAGE_BYTES = 2 * W_HALF_BYTES;             // 16384 / 32768
  static constexpr int A_BOX_BYTES = MPAD * 128;                     // one 64(k) x MPAD chunk, 128-byte swizzle
  static constexpr int A_STAGE_BYTES = 4 * A_BOX_BYTES;
  static constexpr int S_STAGE_BYTES = 2048;                         // up to eight 128-column scale rows (group 32)
  static constexpr int STAGE_BYTES = W_STAGE_BYTES + A_STAGE_BYTES + S_STAGE_BYTES;
  static constexpr int NSTAGES_RAW = (198 * 1024) / STAGE_BYTES;
  static constexpr int NSTAGES = NSTAGES_RAW > 8 ? 8 : NSTAGES_RAW;
  static constexpr int NACC = 2;
  static constexpr int TMEM_COLS = 512;
  static constexpr int BAR_BYTES = (2 * NSTAGES + 2 * kNA + 4) * 8 + 16;
  static constexpr int SMEM_BYTES = 1024 /*align slack*/ + NSTAGES * STAGE_BYTES + BAR_BYTES;
  static_assert(NSTAGES >= 2, "pipeline needs at least two stages");
  static_assert(kAccCol0 + NACC * MPAD <= 512, "TMEM budget");
};

struct GemmArgs {
  const uint8_t* B;      // packed weights
  const void* scales;    // [num_groups, N]
  void* C;               // [M, N]
  float* scratch;        // [2 * grid][MPAD * 128] fp32 partial slots
  int* counters;         // [tiles], zero between launches
  int M, N, K;
  int num_groups, group_size;  // group_size == K when channelwise
  int n_tiles, m_chunks, KS;   // KS = ceil(K / kStageK)
  int total_units;
  int groups_per_stage;        // scale rows staged per stage: 0 (channelwise), 2, 4 or 8
  int dbg;                     // bring-up switches (B200_GEMM_DBG), 0 in production
  long long* prof;             // bring-up cycle counters [grid][16] (B200_GEMM_PROF), nullptr in production
};

struct Partition {
  int q, r;
  __device__ __forceinline__ Partition(int total, int g) : q(total / g), r(total % g) {}
  __device__ __forceinline__ int start(int b) const { return b * q + min(b, r); }
  __device__ __forceinline__ int owner(int u) const {
    const int big = r * (q + 1);
    return u < big ? u / (q + 1) : r + (u - big) / q;
  }
};


// ---- dequantisation ------------------------------------------------------------------------------------------
// one packed int32 of the 4-bit Marlin layout -> 4 registers in tcgen05.st.16x128b.x2 order:
//   o[0] = (n=c  , k=2q,2q+1)  o[1] = (n=c+8, k=2q,2q+1)  o[2] = (n=c  , k=2q+8,2q+9)  o[3] = (n=c+8, k=2q+8,2q+9)
// nibble positions (marlin_perms.py:29 interleave [0,2,4,6,1,3,5,7]): p0,p4 -> o[0]; p1,p5 -> o[2]; p2,p6 -> o[1]; p3,p7 -> o[3]
template <typename T>
__device__ __forceinline__ void dequant4(uint32_t q, uint32_t (&o)[4]);

template <>
__device__ __forceinline__ void dequant4<__half>(uint32_t q, uint32_t (&o)[4]) {
  // (q & 0x000f000f) | 0x64006400 = half2(1024 + nib); (q & 0x00f000f0) | 0x6400.. = half2(1024 + 16 nib)
  constexpr uint32_t LO = 0x000f000f, HI = 0x00f000f0, EX = 0x64006400;
  constexpr uint32_t SUB = 0x64086408;   // 1032 = 1024 + 8
  constexpr uint32_t MUL = 0x2c002c00;   // 1/16
  constexpr uint32_t ADD = 0xd480d480;   // -72 = -(1024/16 + 8)
  uint32_t lo0, hi0, lo1, hi1;
  asm("lop3.b32 %0, %1, %2, %3, 0xea;" : "=r"(lo0) : "r"(q), "n"(LO), "n"(EX));
  asm("lop3.b32 %0, %1, %2, %3, 0xea;" : "=r"(hi0) : "r"(q), "n"(HI), "n"(EX));
  q >>= 8;
  asm("lop3.b32 %0, %1, %2, %3, 0xea;" : "=r"(lo1) : "r"(q), "n"(LO), "n"(EX));
  asm("lop3.b32 %0, %1, %2, %3, 0xea;" : "=r"(hi1) : "r"(q), "n"(HI), "n"(EX));
  __half2 a = __hsub2(*reinterpret_cast<__half2*>(&lo0), *reinterpret_cast<const __half2*>(&SUB));
  __half2 b = __hsub2(*reinterpret_cast<__half2*>(&lo1), *reinterpret_cast<const __half2*>(&SUB));
  __half2 c = __hfma2(*reinterpret_cast<__half2*>(&hi0), *reinterpret_cast<const __half2*>(&MUL),
                      *reinterpret_cast<const __half2*>(&ADD));
  __half2 d = __hfma2(*reinterpret_cast<__half2*>(&hi1), *reinterpret_cast<const __half2*>(&MUL),
                      *reinterpret_cast<const __half2*>(&ADD));
  o[0] = *reinterpret_cast<uint32_t*>(&a);
  o[1] = *reinterpret_cast<uint32_t*>(&b);
  o[2] = *reinterpret_cast<uint32_t*>(&c);
  o[3] = *reinterpret_cast<uint32_t*>(&d);
}

template <>
__device__ __forceinline__ void dequant4<__nv_bfloat16>(uint32_t q, uint32_t (&o)[4]) {
  // bf16 has 7 mantissa bits: (nib | 0x4300) = 128 + nib exactly; each nibble pair needs its own shift
  constexpr uint32_t MASK = 0x000f000f, EX = 0x43004300;
  constexpr uint32_t ONE = 0x3f803f80;   // 1.0
  constexpr uint32_t ADD = 0xc308c308;   // -136 = -(128 + 8)
  uint32_t v[4];
  asm("lop3.b32 %0, %1, %2, %3, 0xea;" : "=r"(v[0]) : "r"(q), "n"(MASK), "n"(EX));  // p0,p4 -> o[0]
  q >>= 4;
  asm("lop3.b32 %0, %1, %2, %3, 0xea;" : "=r"(v[2]) : "r"(q), "n"(MASK), "n"(EX));  // p1,p5 -> o[2]
  q >>= 4;
  asm("lop3.b32 %0, %1, %2, %3, 0xea;" : "=r"(v[1]) : "r"(q), "n"(MASK), "n"(EX));  // p2,p6 -> o[1]
  q >>= 4;
  asm("lop3.b32 %0, %1, %2, %3, 0xea;" : "=r"(v[3]) : "r"(q), "n"(MASK), "n"(EX));  // p3,p7 -> o[3]
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    __nv_bfloat162 t = __hfma2(*reinterpret_cast<__nv_bfloat162*>(&v[i]), *reinterpret_cast<const __nv_bfloat162*>(&ONE),
                               *reinterpret_cast<const __nv_bfloat162*>(&ADD));
    o[i] = *reinterpret_cast<uint32_t*>(&t);
  }
}

// 8-bit: two int32 per fragment (half b = 0/1), byte order [k=2q, 2q+8, 2q+1, 2q+9] (marlin_perms.py:31)
template <typename T>
__device__ __forceinline__ void dequant8(uint32_t q0, uint32_t q1, uint32_t (&o)[4]);

template <>
__device__ __forceinline__ void dequant8<__half>(uint32_t q0, uint32_t q1, uint32_t (&o)[4]) {
  constexpr uint32_t EX = 0x64646464;
  constexpr uint32_t SUB = 0x64806480;  // 1152 = 1024 + 128
  uint32_t t[4];
  t[0] = __byte_perm(q0, EX, 0x4240);  // bytes 0,2 -> (k=2q, 2q+1), n = c
  t[1] = __byte_perm(q1, EX, 0x4240);  //                             n = c + 8
  t[2] = __byte_perm(q0, EX, 0x4341);  // bytes 1,3 -> (k=2q+8, 2q+9)
  t[3] = __byte_perm(q1, EX, 0x4341);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    __half2 v = __hsub2(*reinterpret_cast<__half2*>(&t[i]), *reinterpret_cast<const __half2*>(&SUB));
    o[i] = *reinterpret_cast<uint32_t*>(&v);
  }
}

template <>
__device__ __forceinline__ void dequant8<__nv_bfloat16>(uint32_t q0, uint32_t q1, uint32_t (&o)[4]) {
  // through fp32: (0x4B000000 | byte) = 8388608 + byte exactly; subtract 8388736 = 8388608 + 128
  constexpr uint32_t EX = 0x4B000000;
  const float bias = 8388736.0f;
  auto one = [&](uint32_t q, int sel_lo, int sel_hi) -> uint32_t {
    float f0 = __uint_as_float(__byte_perm(q, EX, sel_lo)) - bias;
    float f1 = __uint_as_float(__byte_perm(q, EX, sel_hi)) - bias;
    __nv_bfloat162 v = __floats2bfloat162_rn(f0, f1);
    return *reinterpret_cast<uint32_t*>(&v);
  };
  o[0] = one(q0, 0x7650, 0x7652);
  o[1] = one(q1, 0x7650, 0x7652);
  o[2] = one(q0, 0x7651, 0x7653);
  o[3] = one(q1, 0x7651, 0x7653);
}

template <typename T> struct Pack2;
template <> struct Pack2<__half> {
  using type = __half2;
  static __device__ __forceinline__ uint32_t mul(uint32_t v, uint32_t s) {
    __half2 r = __hmul2(*reinterpret_cast<__half2*>(&v), *reinterpret_cast<__half2*>(&s));
    return *reinterpret_cast<uint32_t*>(&r);
  }
  static __device__ __forceinline__ uint32_t bcast(uint16_t s) { return (uint32_t)s | ((uint32_t)s << 16); }
  static constexpr uint32_t FMT = 0;  // F16
};
template <> struct Pack2<__nv_bfloat16> {
  using type = __nv_bfloat162;
  static __device__ __forceinline__ uint32_t mul(uint32_t v, uint32_t s) {
    __nv_bfloat162 r = __hmul2(*reinterpret_cast<__nv_bfloat162*>(&v), *reinterpret_cast<__nv_bfloat162*>(&s));
    return *reinterpret_cast<uint32_t*>(&r);
  }
  static __device__ __forceinline__ uint32_t bcast(uint16_t s) { return (uint32_t)s | ((uint32_t)s << 16); }
  static constexpr uint32_t FMT = 1;  // BF16
};

// position of output column n (0..31 within a 32-chunk) in a scale row permuted by scale_perm_single
// (gptq_marlin.py:34-38: perm[8 i + jj] = 2 i + {0,1,8,9,16,17,24,25}[jj])
__device__ __forceinline__ int scale_single_pos(int n32) {
  const int rem = n32 & 7;
  return 8 * (rem >> 1) + 2 * (n32 >> 3) + (rem & 1);
}

#define PROF_T0() const long long p_t0 = args.prof ? clock64() : 0
#define PROF_ACC(var) if (args.prof) { const long long p_t1 = clock64(); var += p_t1 - p_mark; p_mark = p_t1; }
// ---- the kernel --------------------------------------------------------------------------------------------------
template <typename T, int MPAD, int BITS>
__global__ void __launch_bounds__(kGemmThreads, 1)
marlin_gemm_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_w,
                   const __grid_constant__ CUtensorMap tmap_s, const GemmArgs args) {
  using Cfg = GemmCfg<MPAD, BITS>;
  constexpr int NS = Cfg::NSTAGES;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  // layout: [activation stages (1024-aligned)] [weight stages] [scale stages] [barriers]
  const uint32_t act_base = smem_base;
  const uint32_t w_base = act_base + NS * Cfg::A_STAGE_BYTES;
  const uint32_t sc_base = w_base + NS * Cfg::W_STAGE_BYTES;
  const uint32_t bar_base = sc_base + NS * Cfg::S_STAGE_BYTES;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (NS + s); };
  auto afull_bar = [&](int s) { return bar_base + 8u * (2 * NS + s); };
  auto aempty_bar = [&](int s) { return bar_base + 8u * (2 * NS + kNA + s); };
  auto accfull_bar = [&](int s) { return bar_base + 8u * (2 * NS + 2 * kNA + s); };
  auto accempty_bar = [&](int s) { return bar_base + 8u * (2 * NS + 2 * kNA + 2 + s); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * NS + 2 * kNA + 4);
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));  // generic pointer to the aligned base

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long k_t0 = args.prof ? clock64() : 0;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmap_a);
    prefetch_tmap(&tmap_w);
    prefetch_tmap(&tmap_s);
    for (int s = 0; s < NS; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 8 + 1);   // 2 halves x 4 quarter warps + the MMA commit
    }
    for (int s = 0; s < kNA; ++s) {
      mbar_init(afull_bar(s), 4);
      mbar_init(aempty_bar(s), 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(accfull_bar(s), 1);
      mbar_init(accempty_bar(s), 4);
    }
    fence_barrier_init();
  }
  if (warp == kMmaWarp) tmem_alloc<Cfg::TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_launch_dependents();   // the next kernel of the stream may start its own prologue now
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_slot - smem_base));

  const Partition part(args.total_units, gridDim.x);
  const int u_begin = part.start(blockIdx.x), u_end = part.start(blockIdx.x + 1);
  const int n_units = u_end - u_begin;
  const int KS = args.KS;
  const int gps = args.groups_per_stage;

  if (warp < kNumProducers) {
    // ================= TMA producers: stage `it` belongs to warp it % 3; 3 tensor copies per 256 x 128 stage ==========
    // (the whole warp runs the loops so that coordinates / descriptors stay warp-uniform; one elected lane issues)
    {
      const uint64_t pol_w = policy_evict_first();   // weights are streamed exactly once
      const uint64_t pol_a = policy_evict_last();    // activations / scales are re-read by every CTA
      const uint32_t tx = Cfg::W_STAGE_BYTES + Cfg::A_STAGE_BYTES + (uint32_t)gps * 256u;
      auto coords = [&](int it, int& nt, int& mc, int& ks) {
        const int u = u_begin + it;
        const int tile = u / KS;
        ks = u - tile * KS;
        nt = tile % args.n_tiles;
        mc = tile / args.n_tiles;
      };
      auto issue_ws = [&](int it) {
        int nt, mc, ks;
        coords(it, nt, mc, ks);
        const int s = it % NS;
        if (elect_one()) {
          mbar_arrive_expect_tx(full_bar(s), tx);
          tma_2d_g2s_hint(w_base + s * Cfg::W_STAGE_BYTES, &tmap_w, nt * (Cfg::W_ROW_BYTES / 8), ks * (kStageK / 16),
                          full_bar(s), pol_w);
          if (gps) tma_2d_g2s_hint(sc_base + s * Cfg::S_STAGE_BYTES, &tmap_s, nt * kTileN, ks * gps, full_bar(s), pol_a);
        }
        __syncwarp();
      };
      auto issue_a = [&](int it) {
        int nt, mc, ks;
        coords(it, nt, mc, ks);
        const int s = it % NS;
        if (elect_one())
          tma_3d_g2s_hint(act_base + s * Cfg::A_STAGE_BYTES, &tmap_a, 0, mc * MPAD, ks * (kStageK / 64), full_bar(s), pol_a);
        __syncwarp();
      };
      // first ring pass: weights + scales do not depend on the previous kernel -> issue them before the PDL wait
      for (int it = warp; it < n_units && it < NS; it += kNumProducers) issue_ws(it);
      pdl_wait();
      for (int it = warp; it < n_units && it < NS; it += kNumProducers) issue_a(it);
      int it = warp;
      while (it < NS) it += kNumProducers;
      long long p_mark = args.prof ? clock64() : 0, p_wait = 0, p_issue = 0;
      for (; it < n_units; it += kNumProducers) {
        mbar_wait(empty_bar(it % NS), ((it / NS) & 1) ^ 1);
        PROF_ACC(p_wait);
        issue_ws(it);
        issue_a(it);
        PROF_ACC(p_issue);
      }
      if (args.prof && warp == 0 && lane == 0) { args.prof[blockIdx.x * 16 + 0] = p_wait; args.prof[blockIdx.x * 16 + 1] = p_issue; }
    }
  } else if (warp == kMmaWarp) {
    // ================= MMA issuer (warp-converged loops, one elected lane issues MMAs and commits) =================
    {
      constexpr uint32_t idesc = make_idesc(1 /*F32*/, Pack2<T>::FMT, Pack2<T>::FMT, 128, MPAD);
      int it = 0, seg = 0;
      long long p_mark = args.prof ? clock64() : 0, p_wacc = 0, p_wfull = 0, p_wafull = 0, p_issue = 0;
      for (int u = u_begin; u < u_end; ++seg) {
        const int ks0 = u % KS;
        const int ks1 = min(KS, ks0 + (u_end - u));
        const int ab = seg % Cfg::NACC;
        mbar_wait(accempty_bar(ab), ((seg / Cfg::NACC) & 1) ^ 1);
        PROF_ACC(p_wacc);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + kAccCol0 + ab * MPAD;
        for (int ks = ks0; ks < ks1; ++ks, ++it) {
          const int s = it % NS;
          mbar_wait(full_bar(s), (it / NS) & 1);
          PROF_ACC(p_wfull);
          const uint32_t b_smem = act_base + s * Cfg::A_STAGE_BYTES;
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const int hu = 2 * it + h, sa = hu % kNA;
            mbar_wait(afull_bar(sa), (hu / kNA) & 1);
            PROF_ACC(p_wafull);
            tc_fence_after();
            const uint32_t a_tmem = tmem_base + sa * kAColsPerHalf;
            if (elect_one()) {
#pragma unroll
              for (int r = 0; r < kSlabs; ++r) {
                const int R = h * kSlabs + r;   // k16 slab within the stage
                mma_f16_ts(d_tmem, a_tmem + r * 8,
                           make_kmajor_sw128_desc(b_smem + (R >> 2) * Cfg::A_BOX_BYTES + (R & 3) * 32), idesc,
                           (ks > ks0 || h > 0 || r > 0) ? 1u : 0u);
              }
              tc_commit(aempty_bar(sa));
              if (h == 1) tc_commit(empty_bar(s));
              if (h == 1 && ks == ks1 - 1) tc_commit(accfull_bar(ab));
            }
            __syncwarp();
            PROF_ACC(p_issue);
          }
        }
        u += ks1 - ks0;
      }
      if (args.prof && lane == 0) { long long* q = args.prof + blockIdx.x * 16; q[2] = p_wacc; q[3] = p_wfull; q[4] = p_wafull; q[5] = p_issue; }
    }
  } else if (warp >= kDequantWarp0 && warp < kEpiWarp0) {
    // ================= dequant warps: smem (packed) -> registers -> TMEM =================
    const int qd = warp & 3;                      // TMEM lane quarter: rows 32 qd .. 32 qd + 31 of the tile
    const int own = (warp - kDequantWarp0) >> 2;  // this warp handles half stages hu % 4 == own (TMEM slot `own`)
    const int h = own & 1;                        // ... which are always the same half of their stage
    const int t64 = qd >> 1, jpair = qd & 1;
    const int c = lane >> 2;
    const bool grouped = gps > 0;
    const int gph = gps >> 1;                                        // scale rows per half stage
    const int slabs_per_group = grouped ? (kSlabs / gph) : kSlabs;   // k16 slabs sharing one scale row
    const uint32_t a_col = tmem_base + ((uint32_t)(32 * qd) << 16) + own * kAColsPerHalf;
    const uint8_t* wsm = smem_gen + (w_base - smem_base) + h * Cfg::W_HALF_BYTES + t64 * (Cfg::W_ROW_BYTES / 2) +
                         lane * (4 * BITS) + jpair * (2 * BITS);
    const uint8_t* ssm = smem_gen + (sc_base - smem_base) + h * gph * 256 + (64 * t64 + 8 * c + 4 * jpair) * 2;
    long long p_mark = args.prof ? clock64() : 0, p_wfull = 0, p_lds = 0, p_waempty = 0, p_deq = 0, p_st = 0;
    for (int hu = own; hu < 2 * n_units; hu += kNA) {
      const int it = hu >> 1;
      const int s = it % NS;
      mbar_wait(full_bar(s), (it / NS) & 1);
      PROF_ACC(p_wfull);
      // packed words of this thread for all 8 slabs: 64-chunk t64, fragment lane, n16-tiles (2 jpair, 2 jpair + 1)
      uint32_t w[kSlabs][BITS == 4 ? 2 : 4];
#pragma unroll
      for (int r = 0; r < kSlabs; ++r) {
        if (BITS == 4) {
          const uint2 v = *reinterpret_cast<const uint2*>(wsm + s * Cfg::W_STAGE_BYTES + r * Cfg::W_ROW_BYTES);
          w[r][0] = v.x; w[r][1] = v.y;
        } else {
          const uint4 v = *reinterpret_cast<const uint4*>(wsm + s * Cfg::W_STAGE_BYTES + r * Cfg::W_ROW_BYTES);
          w[r][0] = v.x; w[r][1] = v.y; w[r][2] = v.z; w[r][3] = v.w;
        }
      }
      PROF_ACC(p_lds);
      mbar_wait(aempty_bar(own), ((hu / kNA) & 1) ^ 1);
      PROF_ACC(p_waempty);
      tc_fence_after();
      uint32_t s2[4] = {0, 0, 0, 0};  // broadcast scales for (jj, b): index 2 jj + b
#pragma unroll
      for (int r = 0; r < kSlabs; ++r) {
        if (grouped && (r % slabs_per_group) == 0) {
          // scale_perm (gptq_marlin.py:30-33): position 8 c + b' of a 64-chunk holds column c + 8 b'; this thread needs
          // b' = 4 jpair .. 4 jpair + 3  ->  one 8-byte read of the scale row the producer staged with the weights
          const uint2 raw = *reinterpret_cast<const uint2*>(ssm + s * Cfg::S_STAGE_BYTES + (r / slabs_per_group) * 256);
          s2[0] = Pack2<T>::bcast((uint16_t)(raw.x & 0xffff));
          s2[1] = Pack2<T>::bcast((uint16_t)(raw.x >> 16));
          s2[2] = Pack2<T>::bcast((uint16_t)(raw.y & 0xffff));
          s2[3] = Pack2<T>::bcast((uint16_t)(raw.y >> 16));
        }
#pragma unroll
        for (int jj = 0; jj < 2; ++jj) {
          uint32_t o[4];
          if (BITS == 4) dequant4<T>(w[r][jj], o);
          else dequant8<T>(w[r][2 * jj], w[r][2 * jj + 1], o);
          if (grouped) {
            o[0] = Pack2<T>::mul(o[0], s2[2 * jj]);
            o[1] = Pack2<T>::mul(o[1], s2[2 * jj + 1]);
            o[2] = Pack2<T>::mul(o[2], s2[2 * jj]);
            o[3] = Pack2<T>::mul(o[3], s2[2 * jj + 1]);
          }
          tmem_st_16x128b_x2(a_col + ((uint32_t)(16 * jj) << 16) + r * 8, o[0], o[1], o[2], o[3]);
        }
      }
      __syncwarp();
      PROF_ACC(p_deq);
      if (lane == 0) mbar_arrive(empty_bar(s));  // smem slot consumed (values are in registers and have been used)
      tc_wait_st();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(afull_bar(own));
      PROF_ACC(p_st);
    }
    if (args.prof && warp == kDequantWarp0 && lane == 0) {
      long long* q = args.prof + blockIdx.x * 16; q[6] = p_wfull; q[7] = p_lds; q[8] = p_waempty; q[9] = p_deq; q[10] = p_st;
    }
  }
  else if (warp >= kEpiWarp0) {
    // ================= epilogue: TMEM -> registers -> global (direct, or split-K partial + last-arriver reduce) ==
    const int qd = warp & 3;
    const int row = 32 * qd + lane;  // output channel within the tile == TMEM lane
    const int et = threadIdx.x - kEpiWarp0 * 32;  // 0..127
    __shared__ int s_last;
    const bool grouped = gps > 0;
    pdl_wait();   // C, the partial slots and the tile counters may still be in use by the previous kernel
    long long p_mark = args.prof ? clock64() : 0, p_wacc = 0, p_epi = 0;
    int seg = 0;
    for (int u = u_begin; u < u_end; ++seg) {
      const int tile = u / KS, ks0 = u - tile * KS;
      const int ks1 = min(KS, ks0 + (u_end - u));
      const int nt = tile % args.n_tiles, mc = tile / args.n_tiles;
      const int n = nt * kTileN + row;
      const int m0 = mc * MPAD;
      const int m_valid = min(MPAD, args.M - m0);
      const bool n_ok = n < args.N;
      const int ab = seg % Cfg::NACC;
      const bool full_k = (ks0 == 0 && ks1 == KS);
      // slot of this segment: 0 when it is the CTA's first segment, else 1
      const int my_slot = 2 * blockIdx.x + ((u == u_begin) ? 0 : 1);
      float* part_dst = args.scratch + (size_t)my_slot * (kTileN * MPAD);
      T* cdst = reinterpret_cast<T*>(args.C) + (size_t)m0 * args.N + n;

      mbar_wait(accfull_bar(ab), (seg / Cfg::NACC) & 1);
      PROF_ACC(p_wacc);
      tc_fence_after();
      const uint32_t acc_addr = tmem_base + ((uint32_t)(32 * qd) << 16) + kAccCol0 + ab * MPAD;
      if (full_k) {
        float ch_scale = 1.f;
        if (!grouped && n_ok)
          ch_scale = to_f32<T>(reinterpret_cast<const T*>(args.scales)[(n & ~31) + scale_single_pos(n & 31)]);
#pragma unroll 1
        for (int j0 = 0; j0 < MPAD; j0 += 16) {
          if (j0 >= m_valid) break;
          uint32_t v[16];
          tmem_ld_32x32b_x16(acc_addr + j0, v);
          tc_wait_ld();
          if (n_ok) {
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (j0 + j < m_valid) cdst[(size_t)(j0 + j) * args.N] = from_f32<T>(__uint_as_float(v[j]) * ch_scale);
          }
        }
      } else {
#pragma unroll 1
        for (int j0 = 0; j0 < MPAD; j0 += 16) {
          if (j0 >= m_valid) break;
          uint32_t v[16];
          tmem_ld_32x32b_x16(acc_addr + j0, v);
          tc_wait_ld();
#pragma unroll
          for (int j = 0; j < 16; ++j)
            if (j0 + j < m_valid) part_dst[(j0 + j) * kTileN + row] = __uint_as_float(v[j]);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(accempty_bar(ab));

      if (!full_k) {
        // contributors of this tile are the CTAs whose unit ranges intersect [tile*KS, (tile+1)*KS)
        const int b_first = part.owner(tile * KS), b_last = part.owner(tile * KS + KS - 1);
        const int nseg = b_last - b_first + 1;
        __threadfence();
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (et == 0) {
          const int old = atomicAdd(&args.counters[tile], 1);
          s_last = (old == nseg - 1);
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (s_last) {
          __threadfence();
          // the tile's partials are [m][128] fp32 per slot: thread et sums float4 number et + 128 j (row m = j * 4 +
          // et / 32, channels 4 (et % 32) .. +3) over the slots in CTA order, 8 independent 16-byte loads in flight
          const int total4 = m_valid * (kTileN / 4);
          const int r4 = (et & 31) * 4;
          const int nn = nt * kTileN + r4;
          float4 chs = make_float4(1.f, 1.f, 1.f, 1.f);
          if (!grouped && nn < args.N) {
            const T* sp = reinterpret_cast<const T*>(args.scales);
            chs.x = to_f32<T>(sp[(nn & ~31) + scale_single_pos((nn + 0) & 31)]);
            chs.y = to_f32<T>(sp[(nn & ~31) + scale_single_pos((nn + 1) & 31)]);
            chs.z = to_f32<T>(sp[(nn & ~31) + scale_single_pos((nn + 2) & 31)]);
            chs.w = to_f32<T>(sp[(nn & ~31) + scale_single_pos((nn + 3) & 31)]);
          }
          constexpr int U = 8;
#pragma unroll 1
          for (int i0 = et; i0 < total4; i0 += 128 * U) {
            float4 acc[U];
#pragma unroll
            for (int j = 0; j < U; ++j) acc[j] = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int b = b_first; b <= b_last; ++b) {
              const int slot = 2 * b + ((part.start(b) / KS == tile) ? 0 : 1);
              const float4* src = reinterpret_cast<const float4*>(args.scratch + (size_t)slot * (kTileN * MPAD));
              float4 v[U];
#pragma unroll
              for (int j = 0; j < U; ++j) {
                const int idx = i0 + 128 * j;
                v[j] = idx < total4 ? __ldcg(src + idx) : make_float4(0.f, 0.f, 0.f, 0.f);
              }
#pragma unroll
              for (int j = 0; j < U; ++j) {
                acc[j].x += v[j].x; acc[j].y += v[j].y; acc[j].z += v[j].z; acc[j].w += v[j].w;
              }
            }
            if (nn < args.N) {   // N % 64 == 0, so a group of 4 channels is either fully inside or fully outside
#pragma unroll
              for (int j = 0; j < U; ++j) {
                const int idx = i0 + 128 * j;
                if (idx < total4) {
                  const int m = idx >> 5;
                  T o[4] = {from_f32<T>(acc[j].x * chs.x), from_f32<T>(acc[j].y * chs.y), from_f32<T>(acc[j].z * chs.z),
                            from_f32<T>(acc[j].w * chs.w)};
                  *reinterpret_cast<uint2*>(reinterpret_cast<T*>(args.C) + (size_t)(m0 + m) * args.N + nn) =
                      *reinterpret_cast<const uint2*>(o);
                }
              }
            }
          }
          if (et == 0) args.counters[tile] = 0;  // self-reset: the next launch starts from zero
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");  // s_last is reused by the next segment
      }
      u += ks1 - ks0;
      PROF_ACC(p_epi);
    }
    if (args.prof && et == 0) { args.prof[blockIdx.x * 16 + 13] = p_wacc; args.prof[blockIdx.x * 16 + 14] = p_epi; }
  }

  tc_fence_before();
  __syncthreads();
  if (args.prof && threadIdx.x == 0) { args.prof[blockIdx.x * 16 + 11] = clock64() - k_t0; args.prof[blockIdx.x * 16 + 12] = n_units; }
  if (warp == kMmaWarp) tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
}

// act-order: the reference permutes the activation columns into a temporary before the GEMM (permute_cols_kernel,
// gptq_marlin.cu:283-339); the repacked weight rows are sorted by group, so with is_k_full the groups are again
// uniform blocks of group_size rows and the regular kernel applies.
template <typename T>
__global__ void permute_cols_kernel(T* __restrict__ out, const T* __restrict__ a, const int* __restrict__ perm, int K,
                                    int64_t lda) {
  pdl_wait();
  pdl_launch_dependents();
  const int64_t m = blockIdx.y;
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < K; k += gridDim.x * blockDim.x)
    out[m * K + k] = a[m * lda + perm[k]];
}

// ---- host side -------------------------------------------------------------------------------------------------------
struct DeviceScratch {
  float* partials = nullptr;
  int* counters = nullptr;
  void* a_tmp = nullptr;
  size_t partial_floats = 0;
  size_t a_tmp_bytes = 0;
  int counter_ints = 0;
};
static DeviceScratch g_scratch[64];

static bool capturing(cudaStream_t st) {
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  cudaStreamIsCapturing(st, &cap);
  return cap != cudaStreamCaptureStatusNone;
}

static int ensure_scratch(int dev, size_t partial_floats, int counter_ints, cudaStream_t st) {
  DeviceScratch& s = g_scratch[dev];
  if (s.partials && s.partial_floats >= partial_floats && s.counter_ints >= counter_ints) return B200_OK;
  if (capturing(st))
    return set_error(B200_EINVAL,
                     "marlin gemm: split-K scratch must be reserved before CUDA-graph capture "
                     "(run one eager call or b200_gemm_scratch_reserve first)");
  // grow generously once: worst case MPAD = 128 on 256 CTAs
  if (partial_floats < (size_t)2 * 256 * kTileN * 128) partial_floats = (size_t)2 * 256 * kTileN * 128;
  if (counter_ints < 1 << 16) counter_ints = 1 << 16;
  cudaDeviceSynchronize();
  if (s.partials) cudaFree(s.partials);
  if (s.counters) cudaFree(s.counters);
  if (cudaMalloc(&s.partials, partial_floats * sizeof(float)) != cudaSuccess ||
      cudaMalloc(&s.counters, counter_ints * sizeof(int)) != cudaSuccess)
    return set_error(B200_ECUDA, "marlin gemm: scratch allocation failed");
  cudaMemset(s.counters, 0, counter_ints * sizeof(int));
  cudaDeviceSynchronize();
  s.partial_floats = partial_floats;
  s.counter_ints = counter_ints;
  return B200_OK;
}

static int ensure_a_tmp(int dev, size_t bytes, cudaStream_t st) {
  DeviceScratch& s = g_scratch[dev];
  if (s.a_tmp && s.a_tmp_bytes >= bytes) return B200_OK;
  if (capturing(st))
    return set_error(B200_EINVAL, "marlin gemm: act-order scratch must be reserved before CUDA-graph capture (run one eager call first)");
  if (bytes < ((size_t)8 << 20)) bytes = (size_t)8 << 20;
  cudaDeviceSynchronize();
  if (s.a_tmp) cudaFree(s.a_tmp);
  if (cudaMalloc(&s.a_tmp, bytes) != cudaSuccess) return set_error(B200_ECUDA, "marlin gemm: act-order scratch allocation failed");
  s.a_tmp_bytes = bytes;
  return B200_OK;
}

template <typename T, int MPAD, int BITS>
static int launch_gemm(const GemmArgs& a_in, const void* A, int64_t lda, cudaStream_t st) {
  using Cfg = GemmCfg<MPAD, BITS>;
  GemmArgs a = a_in;
  a.m_chunks = (a.M + MPAD - 1) / MPAD;
  const int tiles = a.n_tiles * a.m_chunks;
  a.total_units = tiles * a.KS;
  int grid = num_sms();
  if (grid > a.total_units) grid = a.total_units;
  int dev = 0;
  cudaGetDevice(&dev);
  if (int e = ensure_scratch(dev, (size_t)2 * grid * kTileN * MPAD, tiles, st)) return e;
  a.scratch = g_scratch[dev].partials;
  a.counters = g_scratch[dev].counters;
  CUtensorMap tmap_a, tmap_w, tmap_s;
  const CUtensorMapDataType dt = std::is_same<T, __half>::value ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  // activations [M, K] (row pitch lda) as {64, M, K/64}: one box = 4 chunks of 64 (k) x MPAD rows, 128-byte swizzle,
  // rows beyond M and chunks beyond K read as zero
  bool ok = make_tmap_act3d(&tmap_a, A, dt, (uint64_t)a.M, (uint64_t)a.K, (uint64_t)lda * 2, MPAD, kStageK / 64);
  // packed weights [K/16, N*16*BITS/8 bytes] viewed as 8-byte elements: box = 16 slabs x one 128-column tile
  const uint64_t wrow_bytes = (uint64_t)a.N * 16 * BITS / 8;
  ok = ok && make_tmap_2d(&tmap_w, a.B, CU_TENSOR_MAP_DATA_TYPE_UINT64, 8, (uint64_t)a.K / 16, wrow_bytes / 8, wrow_bytes,
                          kStageK / 16, Cfg::W_ROW_BYTES / 8, CU_TENSOR_MAP_SWIZZLE_NONE);
  // group scales [num_groups, N]: box = (scale rows of one stage) x 128 columns
  if (a.groups_per_stage > 0)
    ok = ok && make_tmap_2d(&tmap_s, a.scales, dt, 2, (uint64_t)a.num_groups, (uint64_t)a.N, (uint64_t)a.N * 2,
                            (uint32_t)a.groups_per_stage, kTileN, CU_TENSOR_MAP_SWIZZLE_NONE);
  else
    tmap_s = tmap_a;
  if (!ok)
    return set_error(B200_ECUDA, "marlin gemm: cuTensorMapEncodeTiled failed (A=%p M=%d K=%d N=%d lda=%lld)", A, a.M, a.K,
                     a.N, (long long)lda);
  auto kern = marlin_gemm_kernel<T, MPAD, BITS>;
  static bool attr_set[64] = {};
  if (!attr_set[dev & 63]) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
    attr_set[dev & 63] = true;
  }
  static const bool prof_on = getenv("B200_GEMM_PROF") != nullptr;
  a.prof = nullptr;
  if (prof_on) { cudaMallocManaged(&a.prof, sizeof(long long) * 16 * grid); cudaMemset(a.prof, 0, sizeof(long long) * 16 * grid); }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kGemmThreads);
  cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, tmap_a, tmap_w, tmap_s, a);
  if (e != cudaSuccess) return set_error(B200_ECUDA, "marlin gemm launch: %s", cudaGetErrorString(e));
  if (prof_on) {
    cudaDeviceSynchronize();
    static const char* names[15] = {"prod.wait_empty", "prod.issue", "mma.wait_accempty", "mma.wait_full", "mma.wait_afull", "mma.issue",
                                    "deq.wait_full", "deq.lds", "deq.wait_aempty", "deq.compute", "deq.wait_st", "cta.total", "cta.units",
                                    "epi.wait_accfull", "epi.work"};
    fprintf(stderr, "[gemm prof] M=%d N=%d K=%d MPAD=%d grid=%d stages=%d:", a.M, a.N, a.K, MPAD, grid, Cfg::NSTAGES);
    for (int i = 0; i < 15; ++i) {
      double sum = 0, mx = 0;
      for (int b = 0; b < grid; ++b) { sum += (double)a.prof[b * 16 + i]; if ((double)a.prof[b * 16 + i] > mx) mx = (double)a.prof[b * 16 + i]; }
      fprintf(stderr, " %s=%.0f(max %.0f)", names[i], sum / grid, mx);
    }
    fprintf(stderr, "\n");
    cudaFree(a.prof);
  }
  return B200_OK;
}

template <typename T, int BITS>
static int dispatch_mpad(const GemmArgs& a, const void* A, int64_t lda, cudaStream_t st) {
  if (a.M <= 16) return launch_gemm<T, 16, BITS>(a, A, lda, st);
  if (a.M <= 32) return launch_gemm<T, 32, BITS>(a, A, lda, st);
  if (a.M <= 64) return launch_gemm<T, 64, BITS>(a, A, lda, st);
  return launch_gemm<T, 128, BITS>(a, A, lda, st);
}

}  // namespace b200

extern "C" int b200_gemm_scratch_reserve(int64_t max_m, int64_t max_n) {
  (void)max_m; (void)max_n;
  int dev = 0;
  cudaGetDevice(&dev);
  return b200::ensure_scratch(dev, 1, 1, 0);
}

extern "C" int b200_gptq_marlin_gemm(void* c, const void* a, const int32_t* b_q_weight, const void* b_scales,
                                     const int32_t* g_idx, const int32_t* perm, int32_t* workspace,
                                     int64_t workspace_numel, int num_bits, int64_t size_m, int64_t size_n,
                                     int64_t size_k, int64_t lda, int num_groups, int is_k_full, int dtype,
                                     void* stream) {
  using namespace b200;
  (void)workspace;
  B200_CHECK(num_bits == 4 || num_bits == 8, "num_bits must be 4 or 8. Got = %d", num_bits);
  B200_CHECK(dtype == B200_F16 || dtype == B200_BF16, "gpt_marlin_gemm only supports bfloat16 and float16");
  B200_CHECK(size_n % 64 == 0, "size_n = %lld, is not divisible by min_thread_n = 64", (long long)size_n);
  B200_CHECK(size_k % 128 == 0, "size_k = %lld is not divisible by 128 (gptq_marlin.py:21 min_thread_k)", (long long)size_k);
  B200_CHECK(workspace_numel >= (size_n / 64) * 16, "workspace.numel = %lld is below min_workspace_size = %lld",
             (long long)workspace_numel, (long long)((size_n / 64) * 16));
  B200_CHECK(num_groups >= 1 && size_k % num_groups == 0, "size_k = %lld, is not divisible by b_scales.size(0) = %d",
             (long long)size_k, num_groups);
  B200_CHECK((reinterpret_cast<uintptr_t>(a) & 15) == 0 && (lda * 2) % 16 == 0, "A must be 16-byte aligned with a 16-byte multiple row pitch");
  const bool act_order = (g_idx != nullptr && perm != nullptr);
  B200_CHECK((g_idx == nullptr) == (perm == nullptr), "g_idx and perm must both be given (act-order) or both be empty");
  if (act_order && !is_k_full)
    return set_error(B200_ENOTSUP, "gptq_marlin_gemm: act-order with is_k_full=False (row-parallel shard of an act-order "
                                   "checkpoint) is not covered by the sm_100a kernel");
  if (size_m == 0 || size_n == 0) return B200_OK;
  const int gs = (int)(size_k / num_groups);
  if (num_groups > 1) B200_CHECK(gs == 32 || gs == 64 || gs == 128, "Unexpected group_size = %d", gs);
  cudaStream_t st = (cudaStream_t)stream;
  if (act_order) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (int e = ensure_a_tmp(dev, (size_t)size_m * size_k * 2, st)) return e;
    void* tmp = g_scratch[dev].a_tmp;
    dim3 grid((unsigned)((size_k + 255) / 256), (unsigned)size_m);
    if (dtype == B200_F16)
      launch_pdl(permute_cols_kernel<__half>, grid, dim3(256), 0, st, (__half*)tmp, (const __half*)a, perm, (int)size_k, lda);
    else
      launch_pdl(permute_cols_kernel<__nv_bfloat16>, grid, dim3(256), 0, st, (__nv_bfloat16*)tmp, (const __nv_bfloat16*)a, perm, (int)size_k, lda);
    a = tmp;
    lda = size_k;
  }
  GemmArgs ga;
  ga.B = reinterpret_cast<const uint8_t*>(b_q_weight);
  ga.scales = b_scales;
  ga.C = c;
  ga.scratch = nullptr;
  ga.counters = nullptr;
  ga.M = (int)size_m; ga.N = (int)size_n; ga.K = (int)size_k;
  ga.num_groups = num_groups; ga.group_size = gs;
  ga.n_tiles = (int)((size_n + kTileN - 1) / kTileN);
  ga.m_chunks = 1;
  ga.KS = (int)((size_k + kStageK - 1) / kStageK);
  ga.total_units = 0;
  ga.groups_per_stage = num_groups > 1 ? kStageK / gs : 0;
  { const char* d = getenv("B200_GEMM_DBG"); ga.dbg = d ? atoi(d) : 0; }
  int rc;
  if (dtype == B200_F16) rc = (num_bits == 4) ? dispatch_mpad<__half, 4>(ga, a, lda, st) : dispatch_mpad<__half, 8>(ga, a, lda, st);
  else rc = (num_bits == 4) ? dispatch_mpad<__nv_bfloat16, 4>(ga, a, lda, st) : dispatch_mpad<__nv_bfloat16, 8>(ga, a, lda, st);
  if (rc) return rc;
  B200_LAUNCH_CHECK("gptq_marlin_gemm");
  return B200_OK;
}
