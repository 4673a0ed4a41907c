// W8A8 scaled GEMM for sm_100a:  out[M,N] = a_s (.) (b_s (.) (A[M,K] . B[K,N])) (+ bias)     (FP8 e4m3 or INT8 operands)
//
// Replaces cutlass_scaled_mm of the reference (csrc/quantization/cutlass_w8a8/scaled_mm_entry.cu:48-100, epilogue
// semantics scaled_mm_c3x.cu:100-183: Compute0 = b_s * acc, Compute1 = a_s * that (+ bias), one rounding to the output
// type; int8 accumulates in s32, fp8 in f32).  The reference's mainloop is CUTLASS 3.5 (not in its tree, sm_90a at
// best); this is an independent tcgen05 kernel:
//
//   * swap-AB: B is column-major [K,N] == a K-major [N,K] weight, so a 128-row slab of output channels is the M=128
//     side of tcgen05.mma (kind::f8f6f4 / kind::i8, K = 32 per instruction) and the token batch (padded to
//     MT in {16,32,64,128,256}) is the N side.  Decode batches fill the tensor core; both operands are K-major in
//     shared memory with the 128-byte swizzle, loaded by tensor-map TMA (zero fill beyond M, N and K).
//   * the accumulator [128 channels x MT tokens] lives in TMEM, double buffered, so the epilogue of one tile overlaps
//     the main loop of the next.
//   * stream-K: the (tile, 128-deep k block) space is cut into one contiguous range per CTA (grid = #SMs).  A tile
//     whose K range is shared writes 32-bit partials to per-CTA slots; the LAST CTA to arrive (atomic ticket) sums the
//     slots in CTA order -- deterministic, nobody spins -- and applies the epilogue.  INT8 partials are exact s32.
//   * programmatic dependent launch: barrier setup, TMEM allocation and the first ring of weight tiles are issued
//     before griddepcontrol.wait.
//
// Warps: w0 TMA producer | w1 MMA issuer | w2 TMEM allocator | w4-7 epilogue (TMEM lane quarter = warp % 4).
#include "common.cuh"
#include "ptx.cuh"
#include "streamk.cuh"
#include "tma_host.cuh"
#include <stdlib.h>

namespace b200 {
using namespace ptx;

namespace smm {

constexpr int kTileN = 128;   // output channels per tile (UMMA M)
constexpr int kBlockK = 128;  // k elements (= bytes) per pipeline stage: one 128-byte swizzle row
constexpr int kThreads = 256;
constexpr int kEpiWarp0 = 4;

template <int MT>
struct Cfg {
  static constexpr int W_BYTES = kTileN * kBlockK;  // 16 KB
  static constexpr int A_BYTES = MT * kBlockK;
  static constexpr int STAGE_BYTES = W_BYTES + A_BYTES;
  static constexpr int NS_RAW = (200 * 1024) / STAGE_BYTES;
  static constexpr int NS = NS_RAW > 10 ? 10 : NS_RAW;
  static constexpr int NBARS = 2 * NS + 4;
  static constexpr int SMEM_BYTES = 1024 + NS * STAGE_BYTES + NBARS * 8 + 32;
  static constexpr int TMEM_COLS = 2 * MT < 32 ? 32 : 2 * MT;
  static_assert(NS >= 3, "pipeline depth");
  static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");
};

struct Args {
  void* out;
  const float* a_scales;
  const float* b_scales;
  const void* bias;
  float* scratch;  // [grid][2 slots][MT * 128] 32-bit partials
  int* counters;   // [tiles], zero between launches
  int M, N, K;
  int64_t ldc;
  int a_scales_numel, b_scales_numel;
  int n_tiles, m_chunks, KS, total_units;
  int ns;          // pipeline stages actually used (<= Cfg::NS): small enough for two CTAs per SM when MT <= 128
};

}  // namespace smm

template <bool IS_FP8, typename OutT, int MT>
__global__ void __launch_bounds__(smm::kThreads, 1)
scaled_mm_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_w,
                 const smm::Args args) {
  using namespace smm;
  using C = Cfg<MT>;
  const int NS = args.ns;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t w_base = smem_base;
  const uint32_t a_base = w_base + NS * C::W_BYTES;
  const uint32_t bar_base = a_base + NS * C::A_BYTES;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (NS + s); };
  auto accfull_bar = [&](int b) { return bar_base + 8u * (2 * NS + b); };
  auto accempty_bar = [&](int b) { return bar_base + 8u * (2 * NS + 2 + b); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * NS + 4);
  const uint32_t flag_slot = tmem_slot + 8;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    prefetch_tmap(&tmap_w);
    prefetch_tmap(&tmap_a);
    for (int s = 0; s < NS; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(accfull_bar(b), 1);
      mbar_init(accempty_bar(b), 4);
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<C::TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_launch_dependents();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_slot - smem_base));

  const Partition part(args.total_units, gridDim.x);
  const int u_begin = part.start(blockIdx.x), u_end = part.start(blockIdx.x + 1);
  const int n_units = u_end - u_begin;
  const int KS = args.KS;

  if (warp == 0) {
    // ================= TMA producer: one weight box [128 n x 128 k] + one activation box [MT m x 128 k] per stage =======
    const uint64_t pol_w = policy_evict_first();
    const uint64_t pol_a = policy_evict_last();
#ifdef B200_SMM_DIV_COORDS   // A/B build only: the producer loop as it was (runtime divisions per unit)
    auto coords = [&](int it, int& nt, int& mc, int& ks) {
      const int u = u_begin + it;
      const int tile = u / KS;
      ks = u - tile * KS;
      nt = tile / args.m_chunks;
      mc = tile - nt * args.m_chunks;
    };
    auto issue_w = [&](int it) {
      int nt, mc, ks;
      coords(it, nt, mc, ks);
      const int s = it % NS;
      if (elect_one()) {
        mbar_arrive_expect_tx(full_bar(s), C::STAGE_BYTES);
        tma_2d_g2s_hint(w_base + s * C::W_BYTES, &tmap_w, ks * kBlockK, nt * kTileN, full_bar(s),
                        args.m_chunks > 1 ? pol_a : pol_w);
      }
      __syncwarp();
    };
    auto issue_a = [&](int it) {
      int nt, mc, ks;
      coords(it, nt, mc, ks);
      const int s = it % NS;
      if (elect_one()) tma_2d_g2s_hint(a_base + s * C::A_BYTES, &tmap_a, ks * kBlockK, mc * MT, full_bar(s), pol_a);
      __syncwarp();
    };
    // weights do not depend on the previous kernel of the stream: first ring pass before the PDL wait
    const int first = n_units < NS ? n_units : NS;
    for (int it = 0; it < first; ++it) issue_w(it);
    pdl_wait();
    for (int it = 0; it < first; ++it) issue_a(it);
    for (int it = NS; it < n_units; ++it) {
      mbar_wait(empty_bar(it % NS), ((it / NS) & 1) ^ 1);
      issue_w(it);
      issue_a(it);
    }
#else
    // (tile, k-block) cursor: divided ONCE, then stepped -- runtime integer divisions per unit in this single-warp loop cost
    // 6-9 % of the Marlin kernel's time before they were removed there (profiles/r2_gemm_ablations.txt); same pattern here
    struct Cursor { int ks, nt, mc; };
    auto cursor_at = [&](int u) {
      Cursor c;
      const int tile = u / KS;
      c.ks = u - tile * KS;
      c.nt = tile / args.m_chunks;
      c.mc = tile - c.nt * args.m_chunks;
      return c;
    };
    auto advance = [&](Cursor& c) {
      if (++c.ks == KS) {
        c.ks = 0;
        if (++c.mc == args.m_chunks) { c.mc = 0; ++c.nt; }
      }
    };
    auto issue_w = [&](int it, const Cursor& c) {
      const int s = it % NS;
      if (elect_one()) {
        mbar_arrive_expect_tx(full_bar(s), C::STAGE_BYTES);
        tma_2d_g2s_hint(w_base + s * C::W_BYTES, &tmap_w, c.ks * kBlockK, c.nt * kTileN, full_bar(s),
                        args.m_chunks > 1 ? pol_a : pol_w);
      }
      __syncwarp();
    };
    auto issue_a = [&](int it, const Cursor& c) {
      const int s = it % NS;
      if (elect_one()) tma_2d_g2s_hint(a_base + s * C::A_BYTES, &tmap_a, c.ks * kBlockK, c.mc * MT, full_bar(s), pol_a);
      __syncwarp();
    };
    // weights do not depend on the previous kernel of the stream: first ring pass before the PDL wait
    const int first = n_units < NS ? n_units : NS;
    Cursor cw = cursor_at(u_begin), ca = cw;
    for (int it = 0; it < first; ++it) { issue_w(it, cw); advance(cw); }
    pdl_wait();
    for (int it = 0; it < first; ++it) { issue_a(it, ca); advance(ca); }
    for (int it = NS; it < n_units; ++it) {
      mbar_wait(empty_bar(it % NS), ((it / NS) & 1) ^ 1);
      issue_w(it, cw);
      issue_a(it, cw);
      advance(cw);
    }
#endif
  } else if (warp == 1) {
    // ================= MMA issuer =================
    constexpr uint32_t idesc = IS_FP8 ? make_idesc(1 /*F32*/, 0 /*E4M3*/, 0, kTileN, MT)
                                      : make_idesc(2 /*S32*/, 1 /*signed 8*/, 1, kTileN, MT);
    int seg = 0;
    for (int u = u_begin; u < u_end; ++seg) {
      const int ks0 = u % KS;
      const int len = min(KS - ks0, u_end - u);
      const int buf = seg & 1;
      mbar_wait(accempty_bar(buf), ((seg >> 1) & 1) ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + buf * MT;
      for (int i = 0; i < len; ++i) {
        const int it = u - u_begin + i, s = it % NS;
        mbar_wait(full_bar(s), (it / NS) & 1);
        tc_fence_after();
        if (elect_one()) {
#pragma unroll
          for (int kk = 0; kk < kBlockK / 32; ++kk) {
            const uint64_t wd = make_kmajor_sw128_desc(w_base + s * C::W_BYTES + kk * 32);
            const uint64_t ad = make_kmajor_sw128_desc(a_base + s * C::A_BYTES + kk * 32);
            if (IS_FP8) mma_f8f6f4_ss(d_tmem, wd, ad, idesc, (i | kk) ? 1u : 0u);
            else mma_i8_ss(d_tmem, wd, ad, idesc, (i | kk) ? 1u : 0u);
          }
          tc_commit(empty_bar(s));
          if (i == len - 1) tc_commit(accfull_bar(buf));
        }
        __syncwarp();
      }
      u += len;
    }
  } else if (warp >= kEpiWarp0) {
    // ================= epilogue =================
    const int qd = warp & 3;
    const int row = 32 * qd + lane;                      // output channel within the tile == TMEM lane
    const int et = threadIdx.x - kEpiWarp0 * 32;
    volatile int* flag = reinterpret_cast<volatile int*>(smem_gen + (flag_slot - smem_base));
    pdl_wait();  // out / partial slots / counters may still be in use by the previous kernel
    int seg = 0;
    for (int u = u_begin; u < u_end; ++seg) {
      const int tile = u / KS, ks0 = u - tile * KS;
      const int len = min(KS - ks0, u_end - u);
      const int nt = tile / args.m_chunks, mc = tile - nt * args.m_chunks;
      const int n = nt * kTileN + row;
      const int m0 = mc * MT;
      const int m_valid = min(MT, args.M - m0);
      const bool n_ok = n < args.N;
      const bool full_k = (ks0 == 0 && len == KS);
      const int buf = seg & 1;
      const uint32_t acc = tmem_base + ((uint32_t)(32 * qd) << 16) + buf * MT;
      const float bs = n_ok ? args.b_scales[args.b_scales_numel > 1 ? n : 0] : 0.f;
      const float bi = (args.bias && n_ok) ? to_f32<OutT>(reinterpret_cast<const OutT*>(args.bias)[n]) : 0.f;
      OutT* cdst = reinterpret_cast<OutT*>(args.out) + (size_t)m0 * args.ldc + n;
      auto finish = [&](const uint32_t* t, int j0) {
        float as[16];   // per-token scales first: the loads must not be serialised behind the stores below
#pragma unroll
        for (int j = 0; j < 16; ++j)
          as[j] = (j0 + j < m_valid) ? __ldg(args.a_scales + (args.a_scales_numel > 1 ? m0 + j0 + j : 0)) : 0.f;
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          if (j0 + j < m_valid && n_ok) {
            const float a = IS_FP8 ? __uint_as_float(t[j]) : (float)(int)t[j];
            cdst[(size_t)(j0 + j) * args.ldc] = from_f32<OutT>(fmaf(as[j], bs * a, bi));
          }
        }
      };
      mbar_wait(accfull_bar(buf), (seg >> 1) & 1);
      tc_fence_after();
      // a segment at the start of this CTA's range uses slot 0, one that only ends the range slot 1
      uint32_t* my_slot = reinterpret_cast<uint32_t*>(args.scratch) +
                          ((size_t)blockIdx.x * 2 + (u == u_begin ? 0 : 1)) * (kTileN * MT);
      for (int j0 = 0; j0 < MT; j0 += 16) {
        if (j0 < m_valid) {
          uint32_t t[16];
          tmem_ld_32x32b_x16(acc + j0, t);
          tc_wait_ld();
          if (full_k) {
            finish(t, j0);
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (j0 + j < m_valid) __stcg(my_slot + (j0 + j) * kTileN + row, t[j]);
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(accempty_bar(buf));   // TMEM buffer is free for the segment after next
      if (!full_k) {
        // ticket: the last contributor of the tile reduces all slots in CTA order (deterministic) and writes the output
        const int c_first = part.owner(tile * KS), c_last = part.owner(tile * KS + KS - 1);
        __threadfence();
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (et == 0) {
          const int old = atomicAdd(&args.counters[tile], 1);
          const int last = (old == c_last - c_first);
          if (last) args.counters[tile] = 0;
          *flag = last;
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
        const bool is_last = *flag != 0;
        asm volatile("bar.sync 1, 128;" ::: "memory");   // flag is rewritten by the next segment
        if (is_last) {
          __threadfence();
          for (int j0 = 0; j0 < MT; j0 += 16) {
            if (j0 >= m_valid) break;
            uint32_t t[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) t[j] = 0;
            for (int c = c_first; c <= c_last; ++c) {
              const int c_start = part.start(c);
              const int slot = (c_start >= tile * KS) ? 0 : 1;
              const uint32_t* src = reinterpret_cast<const uint32_t*>(args.scratch) +
                                    ((size_t)c * 2 + slot) * (kTileN * MT) + j0 * kTileN + row;
              uint32_t v[16];
#pragma unroll
              for (int j = 0; j < 16; ++j) v[j] = (j0 + j < m_valid) ? __ldcg(src + j * kTileN) : 0u;
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                if (IS_FP8) t[j] = __float_as_uint(__uint_as_float(t[j]) + __uint_as_float(v[j]));
                else t[j] = (uint32_t)((int)t[j] + (int)v[j]);
              }
            }
            finish(t, j0);
          }
        }
      }
      u += len;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<C::TMEM_COLS>(tmem_base);
}

// ---- host side -------------------------------------------------------------------------------------------------------
template <bool IS_FP8, typename OutT, int MT>
static int launch_scaled_mm(smm::Args a, const void* A, const void* B, int64_t lda, int64_t ldb, cudaStream_t st) {
  using namespace smm;
  using C = Cfg<MT>;
  a.m_chunks = (a.M + MT - 1) / MT;
  const int tiles = a.n_tiles * a.m_chunks;
  a.total_units = tiles * a.KS;
  int grid = num_sms();
  if (grid > a.total_units) grid = a.total_units;
  // every CTA whose range ends inside a tile publishes MT x 128 32-bit partials and the last arriver of the tile re-reads all
  // of them: with few tiles and a large MT the fix-up outweighs the streaming (qkv shard [8192 x 1280], M = 64: 37.6 us on 148
  // CTAs, 26 us with 8 k-blocks per CTA; o shard [1024 x 8192], M = 16: 8.0 -> 4.8 us).  Keep at least `min_units` k-blocks
  // (16 KB of weights each) per CTA.
  static const int env_minu = getenv("B200_SMM_MIN_UNITS") ? atoi(getenv("B200_SMM_MIN_UNITS")) : 0;
  const int min_units = env_minu > 0 ? env_minu : 8;   // measured on the 70B TP8 shards (profiles/r2_smm_min_units.txt): 8 beats 1 / 4 / 16 / 32
  if (grid > a.total_units / min_units) grid = a.total_units / min_units;
  if (grid < 1) grid = 1;
  int dev = 0;
  cudaGetDevice(&dev);
  GemmScratch sc;
  if (int e = gemm_ensure_scratch(dev, (size_t)grid * 2 * kTileN * MT, tiles, st, &sc)) return e;
  a.scratch = sc.partials;
  a.counters = sc.counters;
  CUtensorMap tmap_a, tmap_w;
  bool ok = make_tmap_2d(&tmap_a, A, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, (uint64_t)a.M, (uint64_t)a.K, (uint64_t)lda, MT,
                         kBlockK);
  ok = ok && make_tmap_2d(&tmap_w, B, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, (uint64_t)a.N, (uint64_t)a.K, (uint64_t)ldb,
                          kTileN, kBlockK);
  if (!ok) return set_error(B200_ECUDA, "scaled_mm: cuTensorMapEncodeTiled failed (M=%d N=%d K=%d lda=%lld ldb=%lld)", a.M,
                            a.N, a.K, (long long)lda, (long long)ldb);
  auto kern = scaled_mm_kernel<IS_FP8, OutT, MT>;
  static bool attr_set[64] = {};
  if (!attr_set[dev & 63]) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES);
    attr_set[dev & 63] = true;
  }
  // decode-sized tiles keep the footprint under half an SM (shared memory and TMEM) so that the CTAs of the NEXT kernel
  // of the stream (programmatic dependent launch) are already resident and prefetching their weights while this
  // kernel drains; the 128 x 256 prefill tile takes the whole SM
  int ns = C::NS;
  if (MT <= 128) {
    const int half = (110 * 1024 - 1024 - 256) / C::STAGE_BYTES;
    if (ns > half) ns = half;
  }
  static const int ns_env = getenv("B200_SMM_NS") ? atoi(getenv("B200_SMM_NS")) : 0;
  if (ns_env >= 2 && ns_env <= C::NS) ns = ns_env;
  a.ns = ns;
  const size_t smem = 1024 + (size_t)ns * C::STAGE_BYTES + (2 * ns + 4) * 8 + 32;
  cudaError_t e = launch_pdl(kern, dim3(grid), dim3(kThreads), smem, st, tmap_a, tmap_w, a);
  if (e != cudaSuccess) return set_error(B200_ECUDA, "scaled_mm launch: %s", cudaGetErrorString(e));
  return B200_OK;
}

template <bool IS_FP8, typename OutT>
static int dispatch_mt(const smm::Args& a, const void* A, const void* B, int64_t lda, int64_t ldb, cudaStream_t st) {
  if (a.M <= 16) return launch_scaled_mm<IS_FP8, OutT, 16>(a, A, B, lda, ldb, st);
  if (a.M <= 32) return launch_scaled_mm<IS_FP8, OutT, 32>(a, A, B, lda, ldb, st);
  if (a.M <= 64) return launch_scaled_mm<IS_FP8, OutT, 64>(a, A, B, lda, ldb, st);
  if (a.M <= 128) return launch_scaled_mm<IS_FP8, OutT, 128>(a, A, B, lda, ldb, st);
  return launch_scaled_mm<IS_FP8, OutT, 256>(a, A, B, lda, ldb, st);
}

}  // namespace b200

extern "C" int b200_scaled_mm(void* out, const void* a, const void* b, const float* a_scales, const float* b_scales,
                              const void* bias, int64_t size_m, int64_t size_n, int64_t size_k, int64_t lda,
                              int64_t ldb, int64_t ldc, int a_scales_numel, int b_scales_numel, int q8_dtype,
                              int out_dtype, void* stream) {
  using namespace b200;
  B200_CHECK(q8_dtype == B200_Q8_INT8 || q8_dtype == B200_Q8_FP8_E4M3, "scaled_mm: operands must be int8 or float8_e4m3fn");
  B200_CHECK(out_dtype == B200_F16 || out_dtype == B200_BF16, "scaled_mm: output must be bfloat16 or float16");
  B200_CHECK(a_scales_numel == 1 || a_scales_numel == size_m, "scaled_mm: a_scales.numel() must be 1 or M");
  B200_CHECK(b_scales_numel == 1 || b_scales_numel == size_n, "scaled_mm: b_scales.numel() must be 1 or N");
  // reference: c.stride(0) % 16 == 0 && b.stride(1) % 16 == 0 (scaled_mm_entry.cu:68-71); TMA needs the same of a
  B200_CHECK(ldc % 16 == 0 && ldb % 16 == 0, "scaled_mm: c.stride(0) and b.stride(1) must be multiples of 16");
  B200_CHECK(lda % 16 == 0 && (reinterpret_cast<uintptr_t>(a) & 15) == 0 && (reinterpret_cast<uintptr_t>(b) & 15) == 0,
             "scaled_mm: a and b must be 16-byte aligned with 16-byte multiple leading strides");
  if (size_m == 0 || size_n == 0) return B200_OK;
  B200_CHECK(size_k > 0, "scaled_mm: K must be positive");
  smm::Args ga;
  ga.out = out;
  ga.a_scales = a_scales;
  ga.b_scales = b_scales;
  ga.bias = bias;
  ga.scratch = nullptr;
  ga.counters = nullptr;
  ga.M = (int)size_m; ga.N = (int)size_n; ga.K = (int)size_k;
  ga.ldc = ldc;
  ga.a_scales_numel = a_scales_numel;
  ga.b_scales_numel = b_scales_numel;
  ga.n_tiles = (int)((size_n + smm::kTileN - 1) / smm::kTileN);
  ga.m_chunks = 1;
  ga.KS = (int)((size_k + smm::kBlockK - 1) / smm::kBlockK);
  ga.total_units = 0;
  cudaStream_t st = (cudaStream_t)stream;
  int rc;
  if (q8_dtype == B200_Q8_FP8_E4M3)
    rc = out_dtype == B200_F16 ? dispatch_mt<true, __half>(ga, a, b, lda, ldb, st)
                               : dispatch_mt<true, __nv_bfloat16>(ga, a, b, lda, ldb, st);
  else
    rc = out_dtype == B200_F16 ? dispatch_mt<false, __half>(ga, a, b, lda, ldb, st)
                               : dispatch_mt<false, __nv_bfloat16>(ga, a, b, lda, ldb, st);
  if (rc) return rc;
  B200_LAUNCH_CHECK("scaled_mm");
  return B200_OK;
}

// reference: scaled_mm_entry.cu:32-46 (capability >= 89 with a recent enough toolkit); this build is sm_100a only
extern "C" int b200_scaled_mm_supports_fp8(int cap) { return cap >= 89 ? 1 : 0; }
