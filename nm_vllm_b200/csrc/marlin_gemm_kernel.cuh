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
//     (one 16 KB request = 16 k16-slabs x 1 KB of a 128-column tile), dequant warps expand them in registers
//     (lop3 magic-number trick, exact zero-point subtraction, group scale in the activation dtype like the
//     reference :270-278) and write them with tcgen05.st.16x128b straight into TMEM.  The Marlin fragment order
//     (thread l holds k = {2q,2q+1,2q+8,2q+9}, n = c, c+8 with c = l/4, q = l%4) IS the register layout of that
//     store shape, so no shuffle is needed and K stays in natural order in TMEM.
//   * one thread can only issue a small tcgen05.mma every ~50 cycles (measured, tools/mma_rate.cu: N=16..64, K=16),
//     which is slower than the weight stream needs, while four issuing warps reach one per ~13 cycles.  The k range
//     of a tile is therefore dealt round-robin, 128 k at a time, to NOWN "own-groups": each has 4 dequant warps (one
//     per TMEM lane quarter), two 64-k TMEM operand slots, one issuer warp and its own fp32 accumulator; the
//     epilogue adds the NOWN accumulators.
//   * activations come by one 3-D tensor-map request per stage (128-byte swizzle, zero-filled rows beyond M) and are
//     the B operand from shared memory.
//   * work is split stream-K style: the (tile, 256-deep k-stage) space is cut into one contiguous range per CTA
//     (grid = #SMs, all CTAs resident).  A tile whose K range is shared by several CTAs is reduced deterministically
//     by the CTA that owns its first k-stage (the one that finishes last): the others write an fp32 partial to their
//     own scratch slot and bump the tile counter; the owner's epilogue warps wait for the counter and add the slots
//     in CTA order WHILE its own MMAs are still running, so the reduction is off the critical path (no float
//     atomics; counters reset themselves).
//   * programmatic dependent launch: the kernel only waits for its predecessor (griddepcontrol.wait) before touching
//     activations / outputs / scratch, so barrier setup, TMEM allocation and the first ring of weight + scale
//     requests overlap the previous kernel of the decode step.
//
// Warp roles: w0-1 TMA producers | w2 TMEM allocator | w4.. issuers (one per own-group) | w8.. dequant (TMEM lane
// quarter = warp%4) | last 4 warps epilogue (quarter = warp%4).
#pragma once
#include "common.cuh"
#include "ptx.cuh"
#include "tma_host.cuh"
#include "streamk.cuh"
#include "ar_fuse.cuh"
#include <stdlib.h>
#include <type_traits>

namespace b200 {
using namespace ptx;

constexpr int kTileN = 128;
constexpr int kStageK = 256;                // one pipeline stage / scheduling unit = 256 (k) x 128 (n) weights
constexpr int kHalfK = 128;                 // one own-group turn: half a stage
constexpr int kSlabs = kHalfK / 16;         // 16-row k slabs per half stage
constexpr int kNumProducers = 2;
constexpr int kAllocWarp = 2;
constexpr int kIssuerWarp0 = 4;
constexpr int kDequantWarp0 = 8;

enum GroupMode { kChannelwise = 0, kGroup128 = 1, kGroupSmall = 2 };

template <int MPAD, int BITS>
struct GemmCfg {
  static constexpr int NOWN = MPAD <= 64 ? 4 : 2;                    // own-groups (issuer warps / accumulators)
  static constexpr int NBUF = MPAD <= 32 ? 2 : 1;                    // accumulator buffers per own-group
  static constexpr int NDQ = 4 * NOWN;                               // dequant warps
  static constexpr int EPI_WARP0 = kDequantWarp0 + NDQ;
  static constexpr int NEG = MPAD >= 64 ? 2 : 1;                     // epilogue warpgroups: each takes MPAD / NEG of the columns
  static constexpr int NEPI = 4 * NEG;                               // epilogue warps
  static constexpr int ACCW = MPAD / NEG;                            // fp32 accumulators per epilogue thread
  static constexpr int THREADS = (EPI_WARP0 + NEPI) * 32;
  // registers per thread: at launch (what ptxas may use under __launch_bounds__(THREADS, 1)), after setmaxnreg.dec in
  // warps 0-7, and after setmaxnreg.inc in the epilogue warps (they take what warps 0-7 gave back)
  static constexpr int REG_LAUNCH = ((65536 / THREADS) / 8) * 8;
  static constexpr int REG_LOW = 40;
  static constexpr int REG_EPI = ((REG_LAUNCH + (256 * (REG_LAUNCH - REG_LOW)) / (NEPI * 32)) / 8) * 8;
  static constexpr bool DEFER_ST = MPAD <= 32;                       // needs 16 more registers in the dequant warps
  static constexpr int ACC_COL0 = 64 * NOWN;                         // TMEM: [own-group operand slots][accumulators]
  static constexpr int TMEM_COLS = 512;
  static constexpr int W_ROW_BYTES = kTileN * 16 * BITS / 8;         // one 16(k) x 128(n) slab: 1024 (4-bit) / 2048 (8-bit)
  static constexpr int W_HALF_BYTES = kSlabs * W_ROW_BYTES;          // 8192 / 16384
  static constexpr int W_STAGE_BYTES = 2 * W_HALF_BYTES;             // 16384 / 32768
  static constexpr int A_BOX_BYTES = MPAD * 128;                     // one 64(k) x MPAD chunk, 128-byte swizzle
  static constexpr int A_STAGE_BYTES = 4 * A_BOX_BYTES;
  static constexpr int S_STAGE_BYTES = 2048;                         // up to eight 128-column scale rows (group 32)
  // two rings: packed weights (HBM latency, released as soon as the dequant warps hold them in registers) and
  // activations + group scales (L2 latency, released when the MMAs that read them have completed)
  static constexpr int NA = MPAD <= 32 ? 4 : (MPAD == 64 ? 3 : 2);
  static constexpr int AS_STAGE_BYTES = A_STAGE_BYTES + S_STAGE_BYTES;
  // a third, small ring: the two g128 scale rows of each stage (512 B), read by the epilogue warps -- deep enough that
  // neither the weight producer (at most NW stages ahead of the dequant warps) nor the epilogue ever waits on it
  static constexpr int NSC = 16;
  static constexpr int SCQ_STAGE_BYTES = 512;
  static constexpr int NW_RAW = (218 * 1024 - NA * AS_STAGE_BYTES - NSC * SCQ_STAGE_BYTES) / W_STAGE_BYTES;
  static constexpr int NW = NW_RAW > 12 ? 12 : NW_RAW;
  static constexpr int NBARS = 2 * NW + 2 * NA + 4 * NOWN + 2 * NOWN * NBUF + 2 * NSC;
  static constexpr int BAR_BYTES = NBARS * 8 + 16;
  // SiLU(gate) * up epilogue (args.silu_inter > 0): the 64 "up" channels of a tile hand their MPAD outputs to the 64 "gate"
  // channels through this buffer, [MPAD][64] in the activation dtype
  static constexpr int XCH_BYTES = MPAD * 64 * 2;
  static constexpr int SMEM_BYTES = 1024 /*align slack*/ + NW * W_STAGE_BYTES + NA * AS_STAGE_BYTES + NSC * SCQ_STAGE_BYTES + BAR_BYTES + XCH_BYTES;
  static_assert(NW >= 2 && NA >= 2, "pipeline needs at least two stages per ring");
  static_assert(ACC_COL0 + NOWN * NBUF * MPAD <= 512, "TMEM budget");
  static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");
};

struct GemmArgs {
  const uint8_t* B;      // packed weights
  const void* scales;    // [num_groups, N]
  void* C;               // [M, N]
  float* scratch;        // [grid][MPAD * 128] fp32 partial slots
  int* counters;         // [tiles], zero between launches
  int M, N, K;
  int num_groups, group_size;  // group_size == K when channelwise
  int n_tiles, m_chunks, KS;   // KS = ceil(K / kStageK)
  int total_units;
  int groups_per_stage;        // scale rows staged per stage: 0 (channelwise), 2, 4 or 8
  int cluster_s;               // > 1: each tile's K range is split over a cluster of this many CTAs (DSMEM reduce)
  long long* prof;             // bring-up cycle counters [grid][16] (B200_GEMM_PROF), nullptr in production
  long long* trace;            // bring-up event trace of CTA trace_cta [11 roles][64][8]
  int trace_cta;
  ArFuse ar;                   // ar.world > 1: the epilogue all-reduces the output over the TP ranks (row-parallel layers)
  const float* asum;           // decode kernel v2: per-(group of 128 k, row) activation sums, fp32 [K/128][asum_stride]
  int asum_stride;
  uint32_t shmul[3];           // decode kernel v2: 2^28, 2^24, 2^20 -- right shifts done as IMAD.HI on the FMA pipe
  int silu_inter;              // > 0: gate_up layer with SiLU(gate) * up fused into the epilogue.  N = 2 * silu_inter, tile t = gate
                               // columns [64 t, 64 t + 64) (TMEM lanes 0..63) + up columns silu_inter + [64 t, 64 t + 64) (lanes
                               // 64..127); C is [M, silu_inter]
  int exp;                     // bring-up ablations (B200_GEMM_EXP, prof / ablate builds only): 1 no dequant ALU, 2 no MMAs, 3 no epilogue reads, 4 no TMEM stores, 5 no activation TMA, 6 no weight TMA
};

// ---- dequantisation ------------------------------------------------------------------------------------------
// one packed int32 of the 4-bit Marlin layout -> 4 registers in tcgen05.st.16x128b order:
//   o[0] = (n=c  , k=2q,2q+1)  o[1] = (n=c+8, k=2q,2q+1)  o[2] = (n=c  , k=2q+8,2q+9)  o[3] = (n=c+8, k=2q+8,2q+9)
// nibble positions (marlin_perms.py:29 interleave [0,2,4,6,1,3,5,7]): p0,p4 -> o[0]; p1,p5 -> o[2]; p2,p6 -> o[1]; p3,p7 -> o[3]
template <typename T>
__device__ __forceinline__ void dequant4(uint32_t q, uint32_t* o);

template <>
__device__ __forceinline__ void dequant4<__half>(uint32_t q, uint32_t* o) {
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
__device__ __forceinline__ void dequant4<__nv_bfloat16>(uint32_t q, uint32_t* o) {
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
__device__ __forceinline__ void dequant8(uint32_t q0, uint32_t q1, uint32_t* o);

template <>
__device__ __forceinline__ void dequant8<__half>(uint32_t q0, uint32_t q1, uint32_t* o) {
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
__device__ __forceinline__ void dequant8<__nv_bfloat16>(uint32_t q0, uint32_t q1, uint32_t* o) {
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
  static __device__ __forceinline__ uint32_t mul(uint32_t v, uint32_t s) {
    __half2 r = __hmul2(*reinterpret_cast<__half2*>(&v), *reinterpret_cast<__half2*>(&s));
    return *reinterpret_cast<uint32_t*>(&r);
  }
  static constexpr uint32_t FMT = 0;  // F16
};
template <> struct Pack2<__nv_bfloat16> {
  static __device__ __forceinline__ uint32_t mul(uint32_t v, uint32_t s) {
    __nv_bfloat162 r = __hmul2(*reinterpret_cast<__nv_bfloat162*>(&v), *reinterpret_cast<__nv_bfloat162*>(&s));
    return *reinterpret_cast<uint32_t*>(&r);
  }
  static constexpr uint32_t FMT = 1;  // BF16
};
__device__ __forceinline__ uint32_t bcast16(uint32_t s) { return __byte_perm(s, 0, 0x1010); }
__device__ __forceinline__ uint32_t bcast16_hi(uint32_t s) { return __byte_perm(s, 0, 0x3232); }

// position of output column n (0..31 within a 32-chunk) in a scale row permuted by scale_perm_single
// (gptq_marlin.py:34-38: perm[8 i + jj] = 2 i + {0,1,8,9,16,17,24,25}[jj])
__device__ __forceinline__ int scale_single_pos(int n32) {
  const int rem = n32 & 7;
  return 8 * (rem >> 1) + 2 * (n32 >> 3) + (rem & 1);
}

// registers -> TMEM, 16 lanes x 128 bit pattern repeated eight times along columns (32 columns = four k16 slabs):
// register 4 i + j of the list is register j of the .x2 pattern (ptx.cuh) for slab i
__device__ __forceinline__ void tmem_st_16x128b_x8(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.16x128b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}

__device__ __forceinline__ int ld_acquire_gpu(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// per-role cycle counters: compiled in only with -DB200_GEMM_PROF (they cost ~4 issue slots per packed word otherwise)
#ifdef B200_GEMM_PROF
#define PROF_ON(x) (x)
#define PROF_ACC(var) if (args.prof) { const long long p_t1 = clock64(); var += p_t1 - p_mark; p_mark = p_t1; }
// event trace of ONE CTA (B200_GEMM_TRACE=<cta>): trace[role][idx][ev] = clock64()
#define EXP_IS(x) (args.exp == (x))
#define TRACE(role, idx, ev) if (args.trace && (int)blockIdx.x == args.trace_cta && lane == 0 && (idx) < 64) { args.trace[((role) * 64 + (idx)) * 8 + (ev)] = clock64(); }
#else
#ifdef B200_GEMM_ABLATE   // ablation switches without the cycle counters (tools/build_variant.py ablate -DB200_GEMM_ABLATE)
#define EXP_IS(x) (args.exp == (x))
#else
#define EXP_IS(x) false
#endif
#define TRACE(role, idx, ev)
#define PROF_ON(x) false
#define PROF_ACC(var)
#endif

// ---- fused all-reduce in the epilogue (row-parallel layers under tensor parallelism) -----------------------------------
// A thread that holds 4 final outputs of one channel rounds them to T (what the unfused GEMM would have stored), pushes
// them as ONE LL cell {2 x T, epoch, 2 x T, epoch} into cell `idx` of its slot in every peer's region, later polls the
// peers' cells of the same index in its own region and sums the W contributions in rank order in fp32: bit-identical on
// every rank and identical to GEMM -> b200_ar_all_reduce.  Every rank runs the same schedule, so `idx` only has to be
// unique per (channel, group of rows) within a launch.
template <typename T>
__device__ __forceinline__ uint32_t ar_pack2(float a, float b) {
  const T x = from_f32<T>(a), y = from_f32<T>(b);
  return (uint32_t)(*reinterpret_cast<const unsigned short*>(&x)) | ((uint32_t)(*reinterpret_cast<const unsigned short*>(&y)) << 16);
}
template <typename T>
__device__ __forceinline__ float ar_unpack(uint32_t w, int hi) {
  const unsigned short b = (unsigned short)(hi ? (w >> 16) : (w & 0xffffu));
  return to_f32<T>(*reinterpret_cast<const T*>(&b));
}
// One-shot (owner < 0): every rank gets every rank's cell and sums.  Two-shot (owner = the rank that owns this thread's
// channels, uniform per warp): the others push to the owner only (phase area 0), the owner sums and pushes the result to
// everyone (phase area 1), the others pick it up -- (W-1)/W of the payload each way instead of (W-1) x.
template <typename T>
__device__ __forceinline__ void ar_push4(const ArFuse& ar, uint32_t epoch, size_t idx, const float (&v)[4], int owner) {
  if (owner == ar.rank) return;   // the owner's own contribution stays in registers
  const uint32_t w0 = ar_pack2<T>(v[0], v[1]), w1 = ar_pack2<T>(v[2], v[3]);
  const size_t off = ar.ll_off + ((size_t)(epoch & 1u) * 2 * ar.world + ar.rank) * ar.ll_slot + idx * 16;
#pragma unroll
  for (int q = 0; q < kArMaxWorld; ++q)
    if (q < ar.world && q != ar.rank && (owner < 0 || q == owner)) st_ll(ar.region[q] + off, w0, w1, epoch);
}
template <typename T>
__device__ __forceinline__ void ar_sum4(const ArFuse& ar, uint32_t epoch, size_t idx, float (&v)[4], int owner) {
  uint8_t* self = nullptr;
#pragma unroll
  for (int q = 0; q < kArMaxWorld; ++q)
    if (q == ar.rank) self = ar.region[q];
  const size_t par_off = ar.ll_off + (size_t)(epoch & 1u) * 2 * ar.world * ar.ll_slot;
  if (owner >= 0 && owner != ar.rank) {
    // two-shot, not the owner: the finished sum arrives in the owner's slot of phase area 1
    const uint8_t* src = self + par_off + (size_t)(ar.world + owner) * ar.ll_slot + idx * 16;
    uint4 x;
    do { x = ld_ll(src); } while (x.y != epoch || x.w != epoch);
    v[0] = ar_unpack<T>(x.x, 0); v[1] = ar_unpack<T>(x.x, 1);
    v[2] = ar_unpack<T>(x.z, 0); v[3] = ar_unpack<T>(x.z, 1);
    return;
  }
  const uint32_t w0 = ar_pack2<T>(v[0], v[1]), w1 = ar_pack2<T>(v[2], v[3]);
  const uint8_t* base = self + par_off + idx * 16;
  float s[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int q = 0; q < kArMaxWorld; ++q) {
    if (q < ar.world) {
      uint32_t a0 = w0, a1 = w1;
      if (q != ar.rank) {
        uint4 x;
        do { x = ld_ll(base + (size_t)q * ar.ll_slot); } while (x.y != epoch || x.w != epoch);
        a0 = x.x; a1 = x.z;
      }
      s[0] += ar_unpack<T>(a0, 0); s[1] += ar_unpack<T>(a0, 1);
      s[2] += ar_unpack<T>(a1, 0); s[3] += ar_unpack<T>(a1, 1);
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) v[i] = s[i];
  if (owner >= 0) {
    // two-shot owner: publish the rounded sums (what every rank will store) in my slot of phase area 1 of every peer
    const uint32_t r0 = ar_pack2<T>(s[0], s[1]), r1 = ar_pack2<T>(s[2], s[3]);
    const size_t off = par_off + (size_t)(ar.world + ar.rank) * ar.ll_slot + idx * 16;
#pragma unroll
    for (int q = 0; q < kArMaxWorld; ++q)
      if (q < ar.world && q != ar.rank) st_ll(ar.region[q] + off, r0, r1, epoch);
  }
}

// ---- the kernel --------------------------------------------------------------------------------------------------
template <typename T, int MPAD, int BITS, int GM, bool FUSE>
__global__ void __launch_bounds__(GemmCfg<MPAD, BITS>::THREADS, 1)
marlin_gemm_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_w,
                   const __grid_constant__ CUtensorMap tmap_s, const GemmArgs args) {
  using Cfg = GemmCfg<MPAD, BITS>;
  // element type of the dequantised weight operand in TMEM: the activation dtype.  (kind::f16 encodes the A and B
  // formats separately, but fp16 weights x bf16 activations traps with an illegal-instruction error on sm_100a --
  // tried, B200 -- so bf16 pays the more expensive bf16 expansion.)
  using WT = T;
  constexpr int NW = Cfg::NW, NA = Cfg::NA, NOWN = Cfg::NOWN, NBUF = Cfg::NBUF;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  // layout: [activation stages (1024-aligned)] [weight stages] [scale stages] [barriers]
  const uint32_t act_base = smem_base;
  const uint32_t w_base = act_base + NA * Cfg::A_STAGE_BYTES;
  const uint32_t sc_base = w_base + NW * Cfg::W_STAGE_BYTES;
  const uint32_t scq_base = sc_base + NA * Cfg::S_STAGE_BYTES;
  const uint32_t bar_base = scq_base + Cfg::NSC * Cfg::SCQ_STAGE_BYTES;
  constexpr int NSC = Cfg::NSC;
  constexpr int NR = 2 * NW + 2 * NA;   // ring barriers
  auto wfull_bar = [&](int s) { return bar_base + 8u * s; };
  auto wempty_bar = [&](int s) { return bar_base + 8u * (NW + s); };
  auto actfull_bar = [&](int s) { return bar_base + 8u * (2 * NW + s); };
  auto actempty_bar = [&](int s) { return bar_base + 8u * (2 * NW + NA + s); };
  auto afull_bar = [&](int o, int sub) { return bar_base + 8u * (NR + 2 * o + sub); };
  auto aempty_bar = [&](int o, int sub) { return bar_base + 8u * (NR + 2 * NOWN + 2 * o + sub); };
  auto accfull_bar = [&](int o, int b) { return bar_base + 8u * (NR + 4 * NOWN + o * NBUF + b); };
  auto accempty_bar = [&](int o, int b) { return bar_base + 8u * (NR + 4 * NOWN + NOWN * NBUF + o * NBUF + b); };
  auto scfull_bar = [&](int s) { return bar_base + 8u * (NR + 4 * NOWN + 2 * NOWN * NBUF + s); };
  auto scempty_bar = [&](int s) { return bar_base + 8u * (NR + 4 * NOWN + 2 * NOWN * NBUF + NSC + s); };
  const uint32_t tmem_slot = bar_base + 8u * Cfg::NBARS;
  const uint32_t xch_base = bar_base + Cfg::BAR_BYTES;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));  // generic pointer to the aligned base

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long k_t0 = PROF_ON(args.prof) ? clock64() : 0;
  if (threadIdx.x == 0) { TRACE(9, 63, 7); }

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmap_w);
    prefetch_tmap(&tmap_a);
    if (GM != kChannelwise) prefetch_tmap(&tmap_s);
  }
  if (threadIdx.x >= 32 && threadIdx.x < 32 + Cfg::NBARS) {
    // one barrier per thread (a serial loop over ~60 barriers costs more than a microsecond of every launch)
    const int i = threadIdx.x - 32;
    uint32_t cnt;
    if (i < NW) cnt = 1;                       // wfull: the producer's expect_tx arrive
    else if (i < 2 * NW) cnt = 8;              // wempty: 2 halves x 4 dequant quarter warps
    else if (i < 2 * NW + NA) cnt = 1;         // actfull
    else if (i < NR) cnt = 2 + (GM == kGroupSmall ? 8 : 0);   // actempty: the two issuers' commits (+ the 8 dequant
                                               // warps of the two halves, which read the staged small-group scales)
    else {
      const int k = i - NR;
      if (k < 2 * NOWN) cnt = 4;                              // afull: 4 dequant quarter warps
      else if (k < 4 * NOWN) cnt = 1;                         // aempty: issuer commit
      else if (k < 4 * NOWN + NOWN * NBUF) cnt = 1;           // accfull: issuer commit
      else if (k < 4 * NOWN + 2 * NOWN * NBUF) cnt = Cfg::NEPI;   // accempty: every epilogue warp
      else if (k < 4 * NOWN + 2 * NOWN * NBUF + NSC) cnt = 1;     // scfull
      else cnt = Cfg::NEPI;                                       // scempty: every epilogue warp
    }
    mbar_init(bar_base + 8u * i, cnt);
    fence_barrier_init();
  }
  __syncthreads();            // barriers are live: the producers start streaming weights right away
  pdl_launch_dependents();    // the next kernel of the stream may start its own prologue now
  // register re-balance (warpgroup-aligned roles): producers / allocator / issuers need few registers, the epilogue
  // warps hold MPAD fp32 accumulators per thread
  // (setmaxnreg is the first instruction of each role branch below, so that ptxas sees which code it governs)

  // two work decompositions:
  //   stream-K (cluster_s <= 1): the (tile, k-stage) space is cut into one contiguous range per CTA; shared tiles are
  //     reduced through global scratch by the CTA that owns the tile's first stage;
  //   cluster split-K (cluster_s = S > 1, used when there are fewer tiles than SMs): tile = blockIdx / S, the S CTAs of
  //     one cluster take consecutive K slices of it and reduce through distributed shared memory -- no global round trip.
  const int KS = args.KS;
  const int CS = args.cluster_s;
  const bool clustered = CS > 1;
  const int c_rank = clustered ? (int)cluster_ctarank() : 0;
  const Partition part(clustered ? KS : args.total_units, clustered ? CS : (int)gridDim.x);
  const int u_base = clustered ? (int)(blockIdx.x / CS) * KS : 0;
  const int u_begin = u_base + part.start(clustered ? c_rank : (int)blockIdx.x);
  const int u_end = u_base + part.start((clustered ? c_rank : (int)blockIdx.x) + 1);
#ifdef B200_GEMM_ABLATE
  // ablation 8: every role finds an empty unit range -> the kernel is launch + prologue + TMEM alloc / free + teardown only
  const int n_units = EXP_IS(8) ? 0 : u_end - u_begin;
#else
  const int n_units = u_end - u_begin;
#endif
  const int gps = args.groups_per_stage;
  const int silu = FUSE ? 0 : args.silu_inter;   // the all-reduce instantiation carries none of the SiLU code

  // everyone but the producers needs the TMEM base address: named barrier 2 after the allocator warp's tcgen05.alloc
  auto get_tmem_base = [&]() -> uint32_t {
    tc_fence_before();
    asm volatile("bar.sync 2, %0;" ::"n"(Cfg::THREADS - kNumProducers * 32) : "memory");
    tc_fence_after();
    return *reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_slot - smem_base));
  };

  if (warp < kNumProducers) {
    setmaxnreg_dec<Cfg::REG_LOW>();
    // ================= TMA producers: warp 0 streams packed weights, warp 1 activations + group scales ================
    // (the whole warp runs the loops so that coordinates / descriptors stay warp-uniform; one elected lane issues)
    // (tile, k-stage) of the first unit by division ONCE, then stepped: runtime integer divisions cost ~50 dependent
    // instructions each, and three of them per unit in these single-warp loops cost 6-9 % of the kernel time
    // (production before / after: profiles/r2_gemm_ablations.txt)
    int ks, nt, mc;
    {
      const int tile0 = u_begin / KS;
      ks = u_begin - tile0 * KS;
      nt = tile0 % args.n_tiles;
      mc = tile0 / args.n_tiles;
    }
    auto advance = [&]() {
      if (++ks == KS) {
        ks = 0;
        if (++nt == args.n_tiles) { nt = 0; ++mc; }
      }
    };
    if (warp == 0) {
      // weights do not depend on the previous kernel of the stream: no PDL wait on this path at all
      const uint64_t pol_w = policy_evict_first();   // weights are streamed exactly once
      const uint64_t pol_s = policy_evict_first();
      int s = 0, q = 0;
      uint32_t ph = 1, qph = 1;                      // parity to wait for on the "empty" barriers: ((it / ring) & 1) ^ 1
      for (int it = 0; it < n_units; ++it) {
        if (it >= NW) mbar_wait(wempty_bar(s), ph);
        TRACE(9, it, 0);
        // the tile's two 64-column halves are separate boxes ([half][slab][64 columns] in shared memory): adjacent
        // columns normally, the gate / up pair (n, n + inter) when the SiLU epilogue is on
        const int col0 = silu ? nt * 64 : nt * kTileN;
        const int col1 = silu ? silu + nt * 64 : nt * kTileN + 64;
        if (EXP_IS(6)) {   // ablation: no weight traffic at all
          if (elect_one()) mbar_arrive(wfull_bar(s));
        } else
        if (elect_one()) {
          mbar_arrive_expect_tx(wfull_bar(s), Cfg::W_STAGE_BYTES);
          tma_2d_g2s_hint(w_base + s * Cfg::W_STAGE_BYTES, &tmap_w, col0 * (BITS / 4), ks * (kStageK / 16), wfull_bar(s), pol_w);
          tma_2d_g2s_hint(w_base + s * Cfg::W_STAGE_BYTES + Cfg::W_STAGE_BYTES / 2, &tmap_w, col1 * (BITS / 4),
                          ks * (kStageK / 16), wfull_bar(s), pol_w);
        }
        __syncwarp();
        if (GM == kGroup128) {
          // the stage's two group-scale rows (static data like the weights) -> the epilogue's scale ring
          if (it >= NSC) mbar_wait(scempty_bar(q), qph);
          if (elect_one()) {
            mbar_arrive_expect_tx(scfull_bar(q), Cfg::SCQ_STAGE_BYTES);
            tma_2d_g2s_hint(scq_base + q * Cfg::SCQ_STAGE_BYTES, &tmap_s, col0, ks * 2, scfull_bar(q), pol_s);
            tma_2d_g2s_hint(scq_base + q * Cfg::SCQ_STAGE_BYTES + Cfg::SCQ_STAGE_BYTES / 2, &tmap_s, col1, ks * 2, scfull_bar(q), pol_s);
          }
          __syncwarp();
          if (++q == NSC) { q = 0; qph ^= 1; }
        }
        if (++s == NW) { s = 0; ph ^= 1; }
        advance();
      }
    } else {
      const uint64_t pol_a = policy_evict_last();    // activations / scales are re-read by every CTA
      const uint32_t tx = Cfg::A_STAGE_BYTES + (GM == kGroupSmall ? (uint32_t)gps * 256u : 0u);
      pdl_wait();
      int s = 0;
      uint32_t ph = 1;
      for (int it = 0; it < n_units; ++it) {
        if (it >= NA) mbar_wait(actempty_bar(s), ph);
        TRACE(10, it, 0);
        if (EXP_IS(5)) {   // ablation: no activation traffic at all
          if (elect_one()) mbar_arrive(actfull_bar(s));
        } else
        if (elect_one()) {
          mbar_arrive_expect_tx(actfull_bar(s), tx);
          tma_3d_g2s_hint(act_base + s * Cfg::A_STAGE_BYTES, &tmap_a, 0, mc * MPAD, ks * (kStageK / 64), actfull_bar(s), pol_a);
          if (GM == kGroupSmall) {
            tma_2d_g2s_hint(sc_base + s * Cfg::S_STAGE_BYTES, &tmap_s, silu ? nt * 64 : nt * kTileN, ks * gps, actfull_bar(s), pol_a);
            tma_2d_g2s_hint(sc_base + s * Cfg::S_STAGE_BYTES + gps * 128, &tmap_s, silu ? silu + nt * 64 : nt * kTileN + 64,
                            ks * gps, actfull_bar(s), pol_a);
          }
        }
        __syncwarp();
        if (++s == NA) { s = 0; ph ^= 1; }
        advance();
      }
    }
  } else if (warp < kIssuerWarp0) {
    setmaxnreg_dec<Cfg::REG_LOW>();
    if (warp == kAllocWarp) tmem_alloc<Cfg::TMEM_COLS>(tmem_slot);
    (void)get_tmem_base();
    if (warp == kAllocWarp + 1 && !clustered && n_units > 0 && (u_begin % KS) != 0) {
      // stream-K contributor: this CTA's first segment starts inside a tile, so its partial goes to the tile's owner.
      // The epilogue warps store it and arrive on named barrier 3; this warp publishes it.
      pdl_wait();
      asm volatile("bar.sync 3, %0;" ::"n"(Cfg::NEPI * 32 + 32) : "memory");
      if (lane == 0) {
        __threadfence();
        atomicAdd(&args.counters[u_begin / KS], 1);
      }
    }
  } else if (warp < kDequantWarp0) {
    setmaxnreg_dec<Cfg::REG_LOW>();
    const uint32_t tmem_base = get_tmem_base();
    if (warp < kIssuerWarp0 + NOWN) {
    // ================= MMA issuers: own-group o turns half stages hu % NOWN == o, each into a FRESH accumulator ====
    // (one half stage = 128 k = one quantisation group at g128: the epilogue warps scale and sum the accumulators
    // in registers, so the issuer knows nothing about tiles)
    const int o = warp - kIssuerWarp0, h = o & 1;
    constexpr uint32_t idesc = make_idesc(1 /*F32*/, Pack2<WT>::FMT, Pack2<T>::FMT, 128, MPAD);
    const uint32_t a_tmem0 = tmem_base + 64 * o;
    long long p_mark = PROF_ON(args.prof) ? clock64() : 0, p_wacc = 0, p_wfull = 0, p_wafull = 0, p_issue = 0;
    int jc = 0;
    for (int hu = o; hu < 2 * n_units; hu += NOWN, ++jc) {
      const int ab = jc % NBUF;
      mbar_wait(accempty_bar(o, ab), ((jc / NBUF) & 1) ^ 1);
      PROF_ACC(p_wacc);
      TRACE(o, jc, 0);
      const uint32_t d_tmem = tmem_base + Cfg::ACC_COL0 + (o * NBUF + ab) * MPAD;
      const int it = hu >> 1, s = it % NA;
      mbar_wait(actfull_bar(s), (it / NA) & 1);
      PROF_ACC(p_wfull);
      TRACE(o, jc, 1);
      const uint32_t b_smem = act_base + s * Cfg::A_STAGE_BYTES;
#pragma unroll
      for (int sub = 0; sub < 2; ++sub) {
        mbar_wait(afull_bar(o, sub), jc & 1);
        PROF_ACC(p_wafull);
        TRACE(o, jc, 2 + 2 * sub);
        tc_fence_after();
        if (elect_one()) {
#pragma unroll
          for (int rr = 0; rr < 4; ++rr) {
            const int R = h * kSlabs + sub * 4 + rr;   // k16 slab within the stage
            if (!EXP_IS(2)) mma_f16_ts(d_tmem, a_tmem0 + sub * 32 + rr * 8,
                       make_kmajor_sw128_desc(b_smem + (R >> 2) * Cfg::A_BOX_BYTES + (R & 3) * 32), idesc,
                       (sub == 0 && rr == 0) ? 0u : 1u);
          }
          tc_commit(aempty_bar(o, sub));
          if (sub == 1) {
            tc_commit(actempty_bar(s));
            tc_commit(accfull_bar(o, ab));
          }
        }
        __syncwarp();
        PROF_ACC(p_issue);
        TRACE(o, jc, 3 + 2 * sub);
      }
    }
    if (PROF_ON(args.prof) && o == 0 && lane == 0) { long long* q = args.prof + blockIdx.x * 16; q[2] = p_wacc; q[3] = p_wfull; q[4] = p_wafull; q[5] = p_issue; }
    }
  } else if (warp < Cfg::EPI_WARP0) {
    const uint32_t tmem_base = get_tmem_base();
    // ================= dequant warps: smem (packed) -> registers -> TMEM =================
    const int qd = warp & 3;                       // TMEM lane quarter: rows 32 qd .. 32 qd + 31 of the tile
    const int o = (warp - kDequantWarp0) >> 2;     // own-group: half stages hu % NOWN == o
    const int h = o & 1;                           // ... which are always the same half of their stage
    const int t64 = qd >> 1, jpair = qd & 1;
    const int c = lane >> 2;
    constexpr int WPT = BITS == 4 ? 2 : 4;         // packed words per thread per slab
    const int gph = gps >> 1;                      // scale rows per half stage (kGroupSmall: 2 or 4)
    const int lspg = gph == 4 ? 1 : 2;             // log2(slabs per group) for kGroupSmall
    const uint32_t a_col = tmem_base + ((uint32_t)(32 * qd) << 16) + 64 * o;
    // weight stage in shared memory: [64-column half t64][16 slabs][32 lanes x 16 B (4-bit) / 32 B (8-bit)]
    const uint8_t* wsm = smem_gen + (w_base - smem_base) + t64 * (Cfg::W_STAGE_BYTES / 2) + h * (Cfg::W_HALF_BYTES / 2) +
                         lane * (4 * BITS) + jpair * (2 * BITS);
    // small-group scale rows: [half t64][gps rows][64 columns]
    const uint8_t* ssm = smem_gen + (sc_base - smem_base) + t64 * gps * 128 + h * gph * 128 + (8 * c + 4 * jpair) * 2;
    long long p_mark = PROF_ON(args.prof) ? clock64() : 0, p_wfull = 0, p_lds = 0, p_waempty = 0, p_deq = 0, p_st = 0;
    int jc = 0;
    for (int hu = o; hu < 2 * n_units; hu += NOWN, ++jc) {
      const int it = hu >> 1;
      const int s = it % NW, sa = it % NA;
      mbar_wait(wfull_bar(s), (it / NW) & 1);
      PROF_ACC(p_wfull);
      if (qd == 0) TRACE(4 + o, jc, 0);
      // packed words of this thread for all 8 slabs: 64-chunk t64, fragment lane, n16-tiles (2 jpair, 2 jpair + 1)
      uint32_t w[kSlabs][WPT];
#pragma unroll
      for (int r = 0; r < kSlabs; ++r) {
        if (BITS == 4) {
          const uint2 v = *reinterpret_cast<const uint2*>(wsm + s * Cfg::W_STAGE_BYTES + r * (Cfg::W_ROW_BYTES / 2));
          w[r][0] = v.x; w[r][1] = v.y;
        } else {
          const uint4 v = *reinterpret_cast<const uint4*>(wsm + s * Cfg::W_STAGE_BYTES + r * (Cfg::W_ROW_BYTES / 2));
          w[r][0] = v.x; w[r][1] = v.y; w[r][2] = v.z; w[r][3] = v.w;
        }
      }
      // kGroupSmall only (groups of 32 / 64 are finer than a half stage, so the scale has to go onto the weights, in
      // the activation dtype like the reference :270-278).  scale_perm (gptq_marlin.py:30-33): position 8 c + b' of a
      // 64-chunk holds column c + 8 b'; this thread needs b' = 4 jpair .. 4 jpair + 3 -> one 8-byte read
      uint32_t s2[4] = {0, 0, 0, 0};  // broadcast scales for (jj, b): index 2 jj + b
      auto load_scales = [&](int row) {
        const uint2 raw = *reinterpret_cast<const uint2*>(ssm + sa * Cfg::S_STAGE_BYTES + row * 128);
        s2[0] = bcast16(raw.x);
        s2[1] = bcast16_hi(raw.x);
        s2[2] = bcast16(raw.y);
        s2[3] = bcast16_hi(raw.y);
      };
      // the packed words are in registers: hand the weight stage back to the producer right away (the arrive has
      // release semantics, so the shared-memory reads above are performed before it)
      __syncwarp();
      if (lane == 0) mbar_arrive(wempty_bar(s));
      if (GM == kGroupSmall) mbar_wait(actfull_bar(sa), (it / NA) & 1);   // the staged scale rows ride with the activations
      PROF_ACC(p_lds);
      if (qd == 0) TRACE(4 + o, jc, 1);
      // one n16-tile (jj) of one 64-k sub-slot: 4 slabs x 4 registers in tcgen05.st order
      auto deq = [&](const int sub, const int jj, uint32_t* ov) {
#pragma unroll
        for (int rr = 0; rr < 4; ++rr) {
          const int r = sub * 4 + rr;
          if (GM == kGroupSmall && (r & ((1 << lspg) - 1)) == 0) load_scales(r >> lspg);
          uint32_t* o4 = &ov[4 * rr];
          if (EXP_IS(1)) { o4[0] = o4[1] = o4[2] = o4[3] = w[r][jj]; }
          else if (BITS == 4) dequant4<WT>(w[r][jj], o4);
          else dequant8<WT>(w[r][2 * jj], w[r][2 * jj + 1], o4);
          if (GM == kGroupSmall) {
            o4[0] = Pack2<T>::mul(o4[0], s2[2 * jj]);
            o4[1] = Pack2<T>::mul(o4[1], s2[2 * jj + 1]);
            o4[2] = Pack2<T>::mul(o4[2], s2[2 * jj]);
            o4[3] = Pack2<T>::mul(o4[3], s2[2 * jj + 1]);
          }
        }
      };
      auto slot_wait = [&](const int sub) {
        mbar_wait(aempty_bar(o, sub), (jc & 1) ^ 1);
        PROF_ACC(p_waempty);
        if (qd == 0) TRACE(4 + o, jc, 2 + 2 * sub);
        tc_fence_after();
      };
      auto slot_publish = [&](const int sub) {
        tc_wait_st();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(afull_bar(o, sub));
        PROF_ACC(p_st);
        if (qd == 0) TRACE(4 + o, jc, 3 + 2 * sub);
      };
      auto st = [&](const int sub, const int jj, const uint32_t* ov) {
        if (!EXP_IS(4)) tmem_st_16x128b_x8(a_col + ((uint32_t)(16 * jj) << 16) + sub * 32, ov);
      };
      if constexpr (Cfg::DEFER_ST) {
        // the first n16-tile of sub-slot 1 is expanded while sub-slot 0's TMEM stores are still in flight
        uint32_t ova[16], ovb[16], ovc[16];
        deq(0, 0, ova);
        slot_wait(0);
        st(0, 0, ova);
        deq(0, 1, ovb);
        st(0, 1, ovb);
        deq(1, 0, ovc);
        slot_publish(0);
        slot_wait(1);
        st(1, 0, ovc);
        deq(1, 1, ova);
        st(1, 1, ova);
        slot_publish(1);
      } else {
#pragma unroll
        for (int sub = 0; sub < 2; ++sub) {
          uint32_t ova[16], ovb[16];
          deq(sub, 0, ova);
          slot_wait(sub);
          st(sub, 0, ova);
          deq(sub, 1, ovb);
          st(sub, 1, ovb);
          slot_publish(sub);
        }
      }
      if (GM == kGroupSmall) {
        __syncwarp();
        if (lane == 0) mbar_arrive(actempty_bar(sa));   // scale rows consumed
      }
    }
    if (PROF_ON(args.prof) && warp == kDequantWarp0 && lane == 0) {
      long long* q = args.prof + blockIdx.x * 16; q[6] = p_wfull; q[7] = p_lds; q[8] = p_waempty; q[9] = p_deq; q[10] = p_st;
    }
  } else {
    // ================= epilogue warps: per half stage TMEM -> registers, scale, accumulate; per tile segment -> global ===
    // thread = one output channel (TMEM lane), acc[m] = its MPAD outputs in fp32.  At g128 a half stage is exactly one
    // quantisation group, so the group scale is applied HERE, once per (group, channel, m), in fp32 -- the dequant warps
    // never touch scales.  Channelwise scales are applied once at the end; small groups were scaled by the dequant warps.
    setmaxnreg_inc<Cfg::REG_EPI>();
    const uint32_t tmem_base = get_tmem_base();
    const int qd = warp & 3;
    const int row = 32 * qd + lane;  // output channel within the tile == TMEM lane
    const int et = threadIdx.x - Cfg::EPI_WARP0 * 32;  // 0 .. 32 NEPI - 1
    constexpr int ACCW = Cfg::ACCW;                    // this thread's share of the MPAD columns (m) of its channel:
    const int jb = (et >> 7) * ACCW;                   // columns jb .. jb + ACCW - 1 (two warpgroups split them at MPAD >= 64)
    constexpr int CH = ACCW < 32 ? ACCW : 32;          // accumulator columns per TMEM read batch
    // position of this channel's scale in a 128-column slice of a scale row (scale_perm, 64-chunks: pos 8 c + b' <-> col c + 8 b')
    // (the ring slot holds [64-column half][2 rows][64 positions])
    const int sc_pos = 128 * (row >> 6) + 8 * (row & 7) + ((row & 63) >> 3);
    const uint32_t acc_lane = tmem_base + ((uint32_t)(32 * qd) << 16) + Cfg::ACC_COL0;
    pdl_wait();   // C, the partial slots and the tile counters may still be in use by the previous kernel
    constexpr bool fuse = FUSE;   // separate instantiation: the plain kernel carries none of the exchange code
    uint32_t ar_epoch = 0;
    if constexpr (FUSE) {
      uint8_t* self = nullptr;
#pragma unroll
      for (int q = 0; q < kArMaxWorld; ++q)
        if (q == args.ar.rank) self = args.ar.region[q];
      ar_epoch = *reinterpret_cast<volatile uint32_t*>(self + args.ar.ctr_off) + 1;
    }
    const size_t ar_npad = (size_t)args.n_tiles * kTileN;
    // two-shot exchange: 32 consecutive channels (one epilogue warp) belong to one rank; -1 = one-shot
    auto ar_owner = [&](int n) { return args.ar.two_shot ? (n >> 5) % args.ar.world : -1; };
    // stores this thread's ACCW final outputs (or, fused, exchanges them with the peers first): rows jb .. jb + ACCW - 1
    auto store_rows = [&](float (&acc)[ACCW], float ch_scale, T* cdst, int n, bool n_ok, int m_valid) {
      if (silu) {
        // gate_up layer: round gate and up to T (what the unfused GEMM would have stored), pair them through shared memory,
        // then silu_and_mul with the reference's roundings (activation_kernels.cu:14-40: T(x / (1 + e^-x)) * y).  Every
        // epilogue thread of the CTA gets here together (CTA-uniform branch).
        T* xch = reinterpret_cast<T*>(smem_gen + (xch_base - smem_base));
        // the activation math is split evenly: the gate thread of a channel finishes the first half of its columns, the up
        // thread the second half; each hands the partner the half it does NOT finish ([column][64 channels] in xch).  (Only
        // the gate threads doing all of it stalled the accumulator hand-off for ~3 us per tile at M = 64.)
        constexpr int HW = ACCW / 2;
        const bool is_up = row >= 64;
        const int r64 = row & 63;
        const int j_send = is_up ? 0 : HW, j_keep = is_up ? HW : 0;
#pragma unroll
        for (int j = 0; j < HW; ++j) {
          const float send = is_up ? acc[j] : acc[HW + j];
          xch[(jb + j_send + j) * 64 + r64] = from_f32<T>(send * ch_scale);
        }
        asm volatile("bar.sync 1, %0;" ::"n"(Cfg::NEPI * 32) : "memory");
#pragma unroll
        for (int j = 0; j < HW; ++j) {
          const int col = jb + j_keep + j;
          if (col < m_valid) {
            const float own = to_f32<T>(from_f32<T>((is_up ? acc[HW + j] : acc[j]) * ch_scale));
            const float other = to_f32<T>(xch[col * 64 + r64]);
            const float x = is_up ? other : own, y = is_up ? own : other;
            cdst[(size_t)col * silu] = from_f32<T>(to_f32<T>(from_f32<T>(x / (1.0f + expf(-x)))) * y);
          }
        }
        asm volatile("bar.sync 1, %0;" ::"n"(Cfg::NEPI * 32) : "memory");
        return;
      }
      if (!n_ok) return;
      if constexpr (FUSE) {
#pragma unroll
        for (int g = 0; g < ACCW / 4; ++g) {
          const float v[4] = {acc[4 * g] * ch_scale, acc[4 * g + 1] * ch_scale, acc[4 * g + 2] * ch_scale, acc[4 * g + 3] * ch_scale};
          ar_push4<T>(args.ar, ar_epoch, (size_t)(jb / 4 + g) * ar_npad + n, v, ar_owner(n));
        }
#pragma unroll
        for (int g = 0; g < ACCW / 4; ++g) {
          float v[4] = {acc[4 * g] * ch_scale, acc[4 * g + 1] * ch_scale, acc[4 * g + 2] * ch_scale, acc[4 * g + 3] * ch_scale};
          ar_sum4<T>(args.ar, ar_epoch, (size_t)(jb / 4 + g) * ar_npad + n, v, ar_owner(n));
#pragma unroll
          for (int i = 0; i < 4; ++i)
            if (jb + 4 * g + i < m_valid) cdst[(size_t)(jb + 4 * g + i) * args.N] = from_f32<T>(v[i]);
        }
        return;
      }
#pragma unroll
      for (int j = 0; j < ACCW; ++j)
        if (jb + j < m_valid) cdst[(size_t)(jb + j) * args.N] = from_f32<T>(acc[j] * ch_scale);
    };
    long long p_mark = PROF_ON(args.prof) ? clock64() : 0, p_wacc = 0, p_epi = 0, p_red = 0;
#ifdef B200_GEMM_ABLATE
    for (int u = u_begin; u < u_begin + n_units;) {
#else
    for (int u = u_begin; u < u_end;) {
#endif
      const int tile = u / KS, ks0 = u - tile * KS;
      const int len = min(KS - ks0, u_end - u);
      const int it0 = u - u_begin, it1 = it0 + len;
      const int nt = tile % args.n_tiles, mc = tile / args.n_tiles;
      const int n = silu ? (row < 64 ? nt * 64 + row : silu + nt * 64 + (row - 64)) : nt * kTileN + row;
      const int m0 = mc * MPAD;
      const int m_valid = min(MPAD, args.M - m0);
      const bool n_ok = n < args.N;
      const bool full_k = (ks0 == 0 && len == KS);
      const bool owner = (ks0 == 0 && !full_k && !clustered);   // this CTA reduces the tile; the others only contribute
      T* cdst = silu ? reinterpret_cast<T*>(args.C) + (size_t)m0 * silu + nt * 64 + (row & 63)
                     : reinterpret_cast<T*>(args.C) + (size_t)m0 * args.N + n;
      float acc[ACCW];
#pragma unroll
      for (int j = 0; j < ACCW; ++j) acc[j] = 0.f;
      float sc_h[2] = {1.f, 1.f};
      for (int hu = 2 * it0; hu < 2 * it1; ++hu) {
        const int o = hu & (NOWN - 1), jc = hu / NOWN, ab = jc % NBUF;
        if (GM == kGroup128 && (hu & 1) == 0) {
          // both group scales of this stage from the scale ring, then release the slot
          const int it = hu >> 1, q = it % NSC;
          mbar_wait(scfull_bar(q), (it / NSC) & 1);
          const uint8_t* sp = smem_gen + (scq_base - smem_base) + q * Cfg::SCQ_STAGE_BYTES + 2 * sc_pos;
          sc_h[0] = to_f32<T>(*reinterpret_cast<const T*>(sp));
          sc_h[1] = to_f32<T>(*reinterpret_cast<const T*>(sp + 128));
          __syncwarp();
          if (lane == 0) mbar_arrive(scempty_bar(q));
        }
        const float sc = (MPAD >= 128) ? ((hu & 1) ? sc_h[1] : sc_h[0]) : sc_h[hu & 1];   // (static indices keep sc_h in registers next to the asm below)
        if (et < 32) TRACE(8, hu, 0);
        mbar_wait(accfull_bar(o, ab), (jc / NBUF) & 1);
        PROF_ACC(p_wacc);
        if (et < 32) TRACE(8, hu, 1);
        tc_fence_after();
        const uint32_t src = acc_lane + (o * NBUF + ab) * MPAD + jb;
#pragma unroll
        for (int c0 = 0; c0 < ACCW; c0 += CH) {
          uint32_t t[CH];
#pragma unroll
          for (int c1 = 0; c1 < CH; c1 += 16) if (!EXP_IS(3)) tmem_ld_32x32b_x16(src + c0 + c1, *reinterpret_cast<uint32_t(*)[16]>(&t[c1]));
          tc_wait_ld();
          if (c0 + CH == ACCW) {
            if (et < 32) TRACE(8, hu, 2);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(accempty_bar(o, ab));
          }
          if constexpr (MPAD >= 128) {
            // prefill-sized M: packed fp32 FMA (FFMA2) halves the arithmetic of the 64-column epilogue threads (2k-token
            // Llama-3-8B prefill 38.0 -> 35.7 ms).  NOT used below 128: measured 5-8 % SLOWER on the decode shapes, whose
            // timing is sensitive to the instruction schedule of this loop (profiles/r2_gemm_ablations.txt)
#pragma unroll
            for (int j = 0; j < CH; j += 2) {
              float2 av = make_float2(acc[c0 + j], acc[c0 + j + 1]);
              const float2 tv = make_float2(__uint_as_float(t[j]), __uint_as_float(t[j + 1]));
              const float2 sv = make_float2(sc, sc);
              asm("fma.rn.f32x2 %0, %1, %2, %0;"
                  : "+l"(*reinterpret_cast<unsigned long long*>(&av))
                  : "l"(*reinterpret_cast<const unsigned long long*>(&sv)), "l"(*reinterpret_cast<const unsigned long long*>(&tv)));
              acc[c0 + j] = av.x;
              acc[c0 + j + 1] = av.y;
            }
          } else {
#pragma unroll
            for (int j = 0; j < CH; ++j) acc[c0 + j] = fmaf(sc, __uint_as_float(t[j]), acc[c0 + j]);
          }
        }
        PROF_ACC(p_epi);
        if (et < 32) TRACE(8, hu, 3);
      }
      float ch_scale = 1.f;
      if (GM == kChannelwise && n_ok)
        ch_scale = to_f32<T>(reinterpret_cast<const T*>(args.scales)[(n & ~31) + scale_single_pos(n & 31)]);

      if (full_k) {
        store_rows(acc, ch_scale, cdst, n, n_ok, m_valid);
      } else if (owner) {
        // contributors are the CTAs after this one whose unit ranges intersect [tile*KS, (tile+1)*KS); for each of them
        // the tile is the FIRST segment of its range, so its partial has long been in its slot when this CTA (whose
        // LAST segment this is) gets here
        const int n_other = part.owner(tile * KS + KS - 1) - (int)blockIdx.x;
        if (et == 0) {
          while (ld_acquire_gpu(&args.counters[tile]) < n_other) {}
          args.counters[tile] = 0;   // self-reset: exactly n_other increments ever arrive
        }
        asm volatile("bar.sync 1, %0;" ::"n"(Cfg::NEPI * 32) : "memory");
        __threadfence();
        for (int b = 1; b <= n_other; ++b) {
          const float* src = args.scratch + (size_t)(blockIdx.x + b) * (kTileN * MPAD) + row;
#pragma unroll
          for (int j = 0; j < ACCW; ++j)
            if (jb + j < m_valid) acc[j] += __ldcg(src + (jb + j) * kTileN);
        }
        PROF_ACC(p_red);
        store_rows(acc, ch_scale, cdst, n, n_ok, m_valid);
      } else if (clustered) {
        // partial of this CTA's K slice -> own shared memory (the activation ring is dead: every MMA of this CTA has
        // completed, or the accumulators above could not have been read), laid out [m][128 channels]
        float* red = reinterpret_cast<float*>(smem_gen + (act_base - smem_base));
#pragma unroll
        for (int j = 0; j < ACCW; ++j)
          if (jb + j < m_valid) red[(jb + j) * kTileN + row] = acc[j];
        // cluster barrier 1: every CTA's partial is in its shared memory; rank r then finishes rows m = r, r + S, ...
        cluster_arrive();
        cluster_wait();
        const uint32_t red_u32 = act_base + (uint32_t)row * 4u;
        // all remote loads of a group of rows are issued before the first add: DSMEM latency is paid once per group
        auto row_group = [&](int mb, float (&out4)[4]) {   // rows mb, mb + CS, mb + 2 CS, mb + 3 CS of this channel, summed over the cluster
          float pv[4][8];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int m = mb + i * CS;
#pragma unroll
            for (int c = 0; c < 8; ++c)
              pv[i][c] = (c < CS && m < m_valid) ? ld_dsmem_f32(mapa(red_u32 + (uint32_t)m * (kTileN * 4u), (uint32_t)c)) : 0.f;
          }
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            float a2 = 0.f;
#pragma unroll
            for (int c = 0; c < 8; ++c) a2 += pv[i][c];
            out4[i] = a2 * ch_scale;
          }
        };
        if constexpr (FUSE) {
          // push every row group first, then collect: one NVLink traversal for the whole tile, not one per group
          constexpr int MAXIT = (MPAD / (8 * Cfg::NEG)) > 0 ? MPAD / (8 * Cfg::NEG) : 1;
          const int G = (MPAD + 4 * CS - 1) / (4 * CS);
          float keep[MAXIT][4];
#pragma unroll
          for (int it = 0; it < MAXIT; ++it) {
            const int k = (et >> 7) + it * Cfg::NEG;
            const int mb = c_rank + k * 4 * CS;
            if (mb < m_valid) {
              row_group(mb, keep[it]);
              if (n_ok) ar_push4<T>(args.ar, ar_epoch, ((size_t)c_rank * G + k) * ar_npad + n, keep[it], ar_owner(n));
            }
          }
#pragma unroll
          for (int it = 0; it < MAXIT; ++it) {
            const int k = (et >> 7) + it * Cfg::NEG;
            const int mb = c_rank + k * 4 * CS;
            if (mb < m_valid && n_ok) {
              ar_sum4<T>(args.ar, ar_epoch, ((size_t)c_rank * G + k) * ar_npad + n, keep[it], ar_owner(n));
#pragma unroll
              for (int i = 0; i < 4; ++i)
                if (mb + i * CS < m_valid) cdst[(size_t)(mb + i * CS) * args.N] = from_f32<T>(keep[it][i]);
            }
          }
        } else {
          for (int mb = c_rank + (et >> 7) * 4 * CS; mb < m_valid; mb += 4 * CS * Cfg::NEG) {
            float r4[4];
            row_group(mb, r4);
#pragma unroll
            for (int i = 0; i < 4; ++i)
              if (n_ok && mb + i * CS < m_valid) cdst[(size_t)(mb + i * CS) * args.N] = from_f32<T>(r4[i]);
          }
        }
      } else {
        float* part_dst = args.scratch + (size_t)blockIdx.x * (kTileN * MPAD);
#pragma unroll
        for (int j = 0; j < ACCW; ++j)
          if (jb + j < m_valid) part_dst[(jb + j) * kTileN + row] = acc[j];
        // hand the "partial is complete" signal to the idle warp 3 (named barrier 3: arrive here, sync there): the
        // gpu-scope fence costs microseconds under load and these warps have accumulators to drain
        asm volatile("bar.arrive 3, %0;" ::"n"(Cfg::NEPI * 32 + 32) : "memory");
      }
      u += len;
      PROF_ACC(p_red);
    }
    if (PROF_ON(args.prof) && et == 0) { long long* q = args.prof + blockIdx.x * 16; q[13] = p_wacc; q[14] = p_epi; q[15] = p_red; }
  }

  if (clustered) {
    // the epilogue warps passed barrier 1 above; barrier 2 keeps every CTA's shared memory alive until all ranks have
    // finished reading it
    if (warp < Cfg::EPI_WARP0) {
      cluster_arrive();
      cluster_wait();
    }
    cluster_arrive();
    cluster_wait();
  }
  tc_fence_before();
  __syncthreads();
  if (FUSE && threadIdx.x == 0) {
    // every CTA is past its exchanges: the last one to get here closes the call (see ar_ticket)
    uint8_t* self = nullptr;
#pragma unroll
    for (int q = 0; q < kArMaxWorld; ++q)
      if (q == args.ar.rank) self = args.ar.region[q];
    volatile uint32_t* ctr = reinterpret_cast<volatile uint32_t*>(self + args.ar.ctr_off);
    ar_ticket(ctr, ctr[0] + 1);
  }
  if (threadIdx.x == 0) { TRACE(10, 63, 7); }
  if (PROF_ON(args.prof) && threadIdx.x == 0) { args.prof[blockIdx.x * 16 + 11] = clock64() - k_t0; args.prof[blockIdx.x * 16 + 12] = n_units; }
  if (warp == kAllocWarp) tmem_dealloc<Cfg::TMEM_COLS>(*reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_slot - smem_base)));
}

// ---- host side: one launch ---------------------------------------------------------------------------------------
template <typename T, int MPAD, int BITS, int GM, bool FUSE>
int launch_gemm(const GemmArgs& a_in, const void* A, int64_t lda, cudaStream_t st) {
  using Cfg = GemmCfg<MPAD, BITS>;
  GemmArgs a = a_in;
  a.m_chunks = (a.M + MPAD - 1) / MPAD;
  const int tiles = a.n_tiles * a.m_chunks;
  a.total_units = tiles * a.KS;
  int grid = num_sms();
  if (grid > a.total_units) grid = a.total_units;
  // fewer tiles than SMs: split each tile's K range over a cluster (<= 8 CTAs) instead of stream-K with global fixups
  a.cluster_s = 0;
  static const int no_cluster = getenv("B200_GEMM_NO_CLUSTER") ? atoi(getenv("B200_GEMM_NO_CLUSTER")) : 0;
  auto kern = marlin_gemm_kernel<T, MPAD, BITS, GM, FUSE>;
  int dev = 0;
  cudaGetDevice(&dev);
  static bool attr_set[64] = {};
  if (!attr_set[dev & 63]) {
    // setmaxnreg.inc in the epilogue warps can only be satisfied if the launch really reserves REG_LAUNCH per thread
    cudaFuncAttributes fa;
    if (cudaFuncGetAttributes(&fa, kern) != cudaSuccess || fa.numRegs != Cfg::REG_LAUNCH)
      return set_error(B200_ECUDA, "marlin gemm: kernel built with %d registers per thread, the register re-balance needs %d",
                       fa.numRegs, Cfg::REG_LAUNCH);
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
    attr_set[dev & 63] = true;
  }
  if (!no_cluster && a.silu_inter == 0 && a.m_chunks == 1 && tiles * 2 <= num_sms()) {
    // largest cluster size whose clusters are all co-resident (GPC boundaries limit this: 45 clusters of 3 on B200)
    static int max_active[64][9] = {};
    int cs = num_sms() / tiles;
    if (cs > 8) cs = 8;
    if (cs > a.KS) cs = a.KS;
    for (; cs > 1; --cs) {
      int& ma = max_active[dev & 63][cs];
      if (ma == 0) {
        cudaLaunchConfig_t q = {};
        q.gridDim = dim3(cs * 64);
        q.blockDim = dim3(Cfg::THREADS);
        q.dynamicSmemBytes = Cfg::SMEM_BYTES;
        cudaLaunchAttribute qa[1];
        qa[0].id = cudaLaunchAttributeClusterDimension;
        qa[0].val.clusterDim.x = (unsigned)cs;
        qa[0].val.clusterDim.y = 1;
        qa[0].val.clusterDim.z = 1;
        q.attrs = qa;
        q.numAttrs = 1;
        int n = 0;
        if (cudaOccupancyMaxActiveClusters(&n, kern, &q) != cudaSuccess) { cudaGetLastError(); n = -1; }
        ma = n > 0 ? n : -1;
      }
      if (ma >= tiles) break;
    }
    if (cs > 1) {
      a.cluster_s = cs;
      grid = tiles * cs;
    }
  }
  GemmScratch sc;
  if (int e = gemm_ensure_scratch(dev, (size_t)grid * kTileN * MPAD, tiles, st, &sc)) return e;
  a.scratch = sc.partials;
  a.counters = sc.counters;
  CUtensorMap tmap_a, tmap_w, tmap_s;
  const CUtensorMapDataType dt = std::is_same<T, __half>::value ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  // activations [M, K] (row pitch lda) as {64, M, K/64}: one box = 4 chunks of 64 (k) x MPAD rows, 128-byte swizzle,
  // rows beyond M and chunks beyond K read as zero
  bool ok = make_tmap_act3d(&tmap_a, A, dt, (uint64_t)a.M, (uint64_t)a.K, (uint64_t)lda * 2, MPAD, kStageK / 64);
  // packed weights [K/16, N*16*BITS/8 bytes] viewed as 8-byte elements: box = 16 slabs x one 64-column half tile
  const uint64_t wrow_bytes = (uint64_t)a.N * 16 * BITS / 8;
  ok = ok && make_tmap_2d(&tmap_w, a.B, CU_TENSOR_MAP_DATA_TYPE_UINT64, 8, (uint64_t)a.K / 16, wrow_bytes / 8, wrow_bytes,
                          kStageK / 16, Cfg::W_ROW_BYTES / 16, CU_TENSOR_MAP_SWIZZLE_NONE);
  // group scales [num_groups, N]: box = (scale rows of one stage) x 64 columns
  if (GM != kChannelwise)
    ok = ok && make_tmap_2d(&tmap_s, a.scales, dt, 2, (uint64_t)a.num_groups, (uint64_t)a.N, (uint64_t)a.N * 2,
                            (uint32_t)a.groups_per_stage, kTileN / 2, CU_TENSOR_MAP_SWIZZLE_NONE);
  else
    tmap_s = tmap_a;
  if (!ok)
    return set_error(B200_ECUDA, "marlin gemm: cuTensorMapEncodeTiled failed (A=%p M=%d K=%d N=%d lda=%lld)", A, a.M, a.K,
                     a.N, (long long)lda);
#ifdef B200_GEMM_PROF
  static const bool prof_on = getenv("B200_GEMM_PROF") != nullptr;
#else
  static const bool prof_on = false;
#endif
  a.prof = nullptr;
  a.trace = nullptr;
  a.trace_cta = 0;
  a.exp = 0;
  if (getenv("B200_GEMM_EXP")) a.exp = atoi(getenv("B200_GEMM_EXP"));
#ifdef B200_GEMM_PROF
  if (prof_on && getenv("B200_GEMM_TRACE")) {
    a.trace_cta = atoi(getenv("B200_GEMM_TRACE"));
    cudaMallocManaged(&a.trace, sizeof(long long) * 11 * 64 * 8);
    cudaMemset(a.trace, 0, sizeof(long long) * 11 * 64 * 8);
  }
#endif
  if (prof_on) { cudaMallocManaged(&a.prof, sizeof(long long) * 16 * grid); cudaMemset(a.prof, 0, sizeof(long long) * 16 * grid); }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(Cfg::THREADS);
  cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if (pdl_enabled()) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  if (a.cluster_s > 1) {
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = (unsigned)a.cluster_s;
    attr[na].val.clusterDim.y = 1;
    attr[na].val.clusterDim.z = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, tmap_a, tmap_w, tmap_s, a);
  if (e != cudaSuccess) return set_error(B200_ECUDA, "marlin gemm launch: %s", cudaGetErrorString(e));
  if (prof_on) {
    cudaDeviceSynchronize();
    static const char* names[16] = {"-", "-", "mma.wait_accempty", "mma.wait_full", "mma.wait_afull", "mma.issue",
                                    "deq.wait_full", "deq.lds", "deq.wait_aempty", "deq.compute", "deq.st", "cta.total", "cta.units",
                                    "epi.wait_accfull", "epi.work", "epi.reduce_prefetch"};
    fprintf(stderr, "[gemm prof] M=%d N=%d K=%d MPAD=%d grid=%d stages=%d:", a.M, a.N, a.K, MPAD, grid, Cfg::NW);
    for (int i = 2; i < 16; ++i) {
      double sum = 0, mx = 0;
      for (int b = 0; b < grid; ++b) { sum += (double)a.prof[b * 16 + i]; if ((double)a.prof[b * 16 + i] > mx) mx = (double)a.prof[b * 16 + i]; }
      fprintf(stderr, " %s=%.0f(max %.0f)", names[i], sum / grid, mx);
    }
    fprintf(stderr, "\n");
    if (a.trace) {
      const long long t0 = a.trace[(9 * 64 + 63) * 8 + 7];
      for (int r = 0; r < 11; ++r)
        for (int i = 0; i < 64; ++i)
          for (int e = 0; e < 8; ++e)
            if (a.trace[(r * 64 + i) * 8 + e]) fprintf(stderr, "[gemm trace] %d %d %d %lld\n", r, i, e, a.trace[(r * 64 + i) * 8 + e] - t0);
      cudaFree(a.trace);
    }
    cudaFree(a.prof);
  }
  return B200_OK;
}

template <typename T, int BITS, int GM>
int dispatch_mpad(const GemmArgs& a, const void* A, int64_t lda, cudaStream_t st) {
  if (a.ar.world > 1) {   // row-parallel layer with the all-reduce fused into the epilogue (b200_ar_fuse_next_gemm)
    if (a.M <= 16) return launch_gemm<T, 16, BITS, GM, true>(a, A, lda, st);
    if (a.M <= 32) return launch_gemm<T, 32, BITS, GM, true>(a, A, lda, st);
    if (a.M <= 64) return launch_gemm<T, 64, BITS, GM, true>(a, A, lda, st);
    return launch_gemm<T, 128, BITS, GM, true>(a, A, lda, st);
  }
  if (a.M <= 16) return launch_gemm<T, 16, BITS, GM, false>(a, A, lda, st);
  if (a.M <= 32) return launch_gemm<T, 32, BITS, GM, false>(a, A, lda, st);
  if (a.M <= 64) return launch_gemm<T, 64, BITS, GM, false>(a, A, lda, st);
  return launch_gemm<T, 128, BITS, GM, false>(a, A, lda, st);
}

// one translation unit per (T, BITS): marlin_gemm_inst_*.cu
template <typename T, int BITS>
int dispatch_gemm(const GemmArgs& a, const void* A, int64_t lda, cudaStream_t st) {
  if (a.groups_per_stage == 0) return dispatch_mpad<T, BITS, kChannelwise>(a, A, lda, st);
  if (a.groups_per_stage == 2) return dispatch_mpad<T, BITS, kGroup128>(a, A, lda, st);
  return dispatch_mpad<T, BITS, kGroupSmall>(a, A, lda, st);
}
int dispatch_gemm_f16_4(const GemmArgs& a, const void* A, int64_t lda, cudaStream_t st);
int dispatch_gemm_f16_8(const GemmArgs& a, const void* A, int64_t lda, cudaStream_t st);
int dispatch_gemm_bf16_4(const GemmArgs& a, const void* A, int64_t lda, cudaStream_t st);
int dispatch_gemm_bf16_8(const GemmArgs& a, const void* A, int64_t lda, cudaStream_t st);

}  // namespace b200
