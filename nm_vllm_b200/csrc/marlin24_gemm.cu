// Sparse-Marlin (2:4) W4A16 / W8A16 GEMM for sm_100a:  C[M,N] = A[M,K] . W24,  W stored compressed (2 of every 4 along K)
//
// Replaces gptq_marlin_24_gemm of the reference (csrc/quantization/marlin/sparse/marlin_24_cuda_kernel.cu:1017-1125 op,
// :111-856 kernel).  The STORED layouts are the reference's: b_q_weight [K/32, N*16/pf] int32 in the Marlin-24 tile order
// (marlin_24_perms.py:11-36), b_meta int16 [K/32, 2N] = the CUTLASS sm80 reordered 2-bit indices reinterpreted
// (marlin_utils.py:161-162, format_24.py:21-46), scales [K/128 or 1, N] permuted in 64-chunks (marlin_24_perms.py:37-43).
// What those layouts mean for one warp (derived in DESIGN.md "Sparse-Marlin operand map"):
//   * one (k32 block kb, 64-column tile t) of weights is 32 lanes x 4 int32 (4-bit) / 8 int32 (8-bit); int32 jj of lane
//     l = 4 col + q dequantises DIRECTLY into the sparse-A registers {a0,a1,a2,a3} of mma.sp m16n8k32 for the 16 output
//     channels n = 64 t + 8 col' + jj (+4 for rows 8..15), compressed k = {2q,2q+1,2q+8,2q+9};
//   * the metadata register of lane l for the instruction pair jj in {2h, 2h+1} is the 32-bit word
//     b_meta32[kb][64 t + 32 h + l], used with sparsity selector jj % 2;
//   * the two group scales of (jj, rows 0-7 / 8-15) are the adjacent halves 8 col + 2 jj, +1 of the permuted scale row.
// So the compressed operand goes straight from global memory into the sparse tensor-core instruction: no shared-memory
// staging of weights, no shuffles.
//
// Work decomposition (HBM-bound, M <= 64 per pass): many small CTAs instead of a deep async pipeline.  CTA = (64-column
// tile, K split s of S); its 8 warps take the k32 blocks of the split round-robin, each streaming its 512 B weight +
// 256 B metadata rows with a 4-deep register prefetch (ld.global.nc, L1 bypass); 4-6 CTAs are resident per SM.  Warp
// partials are summed in shared memory; K splits are summed through fp32 slots in CTA order by the last split to arrive
// (atomic ticket): deterministic, nobody spins.  NOTE this kernel uses the warp-level sparse MMA (HMMA.SP); feeding
// tcgen05.mma.sp from TMEM is the planned next step.
#include "marlin_gemm_kernel.cuh"

namespace b200 {
namespace m24 {

constexpr int kWarps = 4;
constexpr int kThreads = kWarps * 32;

struct Args {
  const __half* A;
  const uint32_t* Bq;
  const uint32_t* meta;
  const __half* scales;
  __half* C;
  float* scratch;
  int* counters;
  int M, N, K;
  int64_t lda;
  int num_groups;   // 1 = channelwise
  int splits;       // K splits per tile
  int kb_per_split; // k32 blocks per split (last split may be shorter)
  int n_tiles;
};

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async4(uint32_t dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ uint4 ldg_stream(const void* p) {
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ uint32_t ldg_stream32(const void* p) {
  uint32_t v;
  asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}

template <int SEL>
__device__ __forceinline__ void mma_sp(float* c, const uint32_t* a, const uint32_t* b, uint32_t e) {
  asm volatile(
      "mma.sp::ordered_metadata.sync.aligned.m16n8k32.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, "
      "{%8,%9,%10,%11}, {%0,%1,%2,%3}, %12, %13;"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]), "r"(b[2]), "r"(b[3]), "r"(e), "n"(SEL));
}

__device__ __forceinline__ void cp_async16_zfill(uint32_t dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr) : "memory");
}

template <int BITS, int MB, int NH>
struct Stage {
  uint32_t w[BITS == 4 ? 4 : 8];
  uint32_t e[2];
  uint32_t s[4];
  uint32_t b[MB][NH][4];   // activation fragments (B operand of mma.sp) of the k32 block
};

// bytes of one warp's ring slot (32 private lane parts + 8 x 80 bytes per 8-row activation half block) and the ring depth
// that keeps a CTA at <= 54 KB of shared memory (4 CTAs per SM: a 448-tile GEMM must not spill into a second wave)
constexpr int slot_bytes(int bits, int mb, int nh) { return 32 * (bits == 4 ? 48 : 80) + mb * nh * 640; }
constexpr int prefetch_depth(int bits, int mb, int nh) {
  int d = (54 * 1024) / (kWarps * slot_bytes(bits, mb, nh));
  return d > 8 ? 8 : d;
}

}  // namespace m24

// NH = 8-row activation halves per 16-row block that carry real rows (1 when M <= 8: half the tensor-core work)
template <int BITS, int MB, bool GROUPED, int NH>
__global__ void __launch_bounds__(m24::kThreads, 4)
marlin24_gemm_kernel(const m24::Args args) {
  using namespace m24;
  constexpr int WPL = BITS == 4 ? 4 : 8;   // packed int32 per lane per (k32 block, 64-column tile)
  constexpr int MC = 16 * MB;              // activation rows per pass
  constexpr int NACC = 32 * MB;            // fp32 accumulators per thread
  extern __shared__ __align__(16) uint8_t smem24[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int col = lane >> 2, q = lane & 3;
  const int t = blockIdx.x / args.splits, sp = blockIdx.x - t * args.splits;
  const int m0 = blockIdx.y * MC;
  const int KB = args.K / 32;
  const int kb0 = sp * args.kb_per_split;
  const int kb1 = min(KB, kb0 + args.kb_per_split);
  const int wrow = args.N * (BITS == 4 ? 2 : 4);
  // (walking the k32 blocks in a per-CTA rotated order, so that the SMs do not all ask L2 for the same activation lines at
  // the same moment, was measured: no gain, and it breaks the DRAM page locality of neighbouring tiles.)
  const int n_it = (kb1 - kb0 + kWarps - 1) / kWarps;
  auto kb_of = [&](int i) { return i < n_it ? kb0 + i * kWarps + warp : kb1; };   // k32 block of this warp's i-th iteration
  const uint32_t* wp = args.Bq + (size_t)t * (32 * WPL) + lane * WPL;
  const uint32_t* mp = args.meta + (size_t)t * 64 + lane;
  const __half* sp_ptr = args.scales + (size_t)t * 64 + col * 8;

  // Each lane streams ITS OWN 16 (32) bytes of weights, 8 bytes of metadata and 16 bytes of scales of a k32 block into a
  // private part of a kPrefetch-deep shared-memory ring with cp.async and reads the same bytes back later (no cross-lane
  // hazard).  The activation fragments ride the same ring: each 8-row half block is 8 x 64 bytes, copied by the warp as 32
  // 16-byte cp.async (rows beyond M zero-filled) into rows 80 bytes apart and read back with ONE ldmatrix.x4 whose matrix i
  // is the 16-byte column i of the 8 rows -- exactly word 4 i + q of row col, the B fragment of mma.sp.  History
  // (profiles/r2_m24_ncu.txt): with the activations as LDGs every warp iteration stalled ~2000 clk on the first HMMA.SP
  // (long scoreboard) however far ahead they were issued; as 4-byte cp.async per fragment word the MIO queue became the
  // limiter (12 LDGSTS per iteration).
  constexpr int LANE_BYTES = BITS == 4 ? 48 : 80;   // odd multiple of 16: a quarter warp's 16-byte reads hit distinct banks
  constexpr int ACT_ROW = 80;                       // same for the ldmatrix rows
  constexpr int ACT_OFF = 32 * LANE_BYTES;
  constexpr int SLOT_BYTES = m24::slot_bytes(BITS, MB, NH);
  constexpr int kPrefetch = m24::prefetch_depth(BITS, MB, NH);
  const uint32_t ring_off = (uint32_t)warp * (kPrefetch * SLOT_BYTES);
  const uint32_t lane_off = ring_off + (uint32_t)lane * LANE_BYTES;
  const uint32_t ring = ptx::smem_u32(smem24) + lane_off;
  const uint32_t act_dst = ptx::smem_u32(smem24) + ring_off + ACT_OFF + (lane >> 2) * ACT_ROW + (lane & 3) * 16;
  const uint32_t act_src = ptx::smem_u32(smem24) + ring_off + ACT_OFF + (lane & 7) * ACT_ROW + (lane >> 3) * 16;
  const __half* arow[MB][NH];
  uint32_t arow_bytes[MB][NH];   // 16 = real row, 0 = beyond M (cp.async zero-fills)
#pragma unroll
  for (int mb = 0; mb < MB; ++mb)
#pragma unroll
    for (int h = 0; h < NH; ++h) {
      const int m = m0 + 16 * mb + 8 * h + col;
      arow_bytes[mb][h] = m < args.M ? 16u : 0u;
      arow[mb][h] = args.A + (size_t)(m < args.M ? m : 0) * args.lda + q * 8;
    }
  auto issue_weights = [&](int slot, int kb) {
    if (kb < kb1) {
      const uint32_t dst = ring + slot * SLOT_BYTES;
      const uint32_t* w = wp + (size_t)kb * wrow;
      cp_async16(dst, w);
      if (BITS == 8) cp_async16(dst + 16, w + 4);
      if (GROUPED) cp_async16(dst + WPL * 4, sp_ptr + (size_t)(kb >> 2) * args.N);   // group_size 128 = 4 k32 blocks
      cp_async4(dst + WPL * 4 + 16, mp + (size_t)kb * args.N);
      cp_async4(dst + WPL * 4 + 20, mp + (size_t)kb * args.N + 32);
    }
  };
  auto issue_acts = [&](int slot, int kb) {
    if (kb < kb1) {
#pragma unroll
      for (int mb = 0; mb < MB; ++mb)
#pragma unroll
        for (int h = 0; h < NH; ++h)
          cp_async16_zfill(act_dst + slot * SLOT_BYTES + (mb * NH + h) * (8 * ACT_ROW), arow[mb][h] + kb * 32, arow_bytes[mb][h]);
    }
  };
  auto read_stage = [&](Stage<BITS, MB, NH>& st_, int slot) {
    const uint8_t* src = smem24 + lane_off + slot * SLOT_BYTES;
    const uint4 v0 = *reinterpret_cast<const uint4*>(src);
    st_.w[0] = v0.x; st_.w[1] = v0.y; st_.w[2] = v0.z; st_.w[3] = v0.w;
    if (BITS == 8) {
      const uint4 v1 = *reinterpret_cast<const uint4*>(src + 16);
      st_.w[4 % WPL] = v1.x; st_.w[5 % WPL] = v1.y; st_.w[6 % WPL] = v1.z; st_.w[7 % WPL] = v1.w;
    }
    const uint2 e = *reinterpret_cast<const uint2*>(src + WPL * 4 + 16);
    st_.e[0] = e.x; st_.e[1] = e.y;
    if (GROUPED) {
      const uint4 sv = *reinterpret_cast<const uint4*>(src + WPL * 4);
      st_.s[0] = sv.x; st_.s[1] = sv.y; st_.s[2] = sv.z; st_.s[3] = sv.w;
    }
#pragma unroll
    for (int mb = 0; mb < MB; ++mb)
#pragma unroll
      for (int h = 0; h < NH; ++h) ldmatrix_x4(st_.b[mb][h], act_src + slot * SLOT_BYTES + (mb * NH + h) * (8 * ACT_ROW));
  };
  // weights / metadata / scales do not depend on the previous kernel: start streaming them before the PDL wait; the
  // activations follow as kPrefetch more groups.  wait_group<kPrefetch-1> in iteration i then covers both halves of stage i
  // (groups retire in order: W0..W{P-1}, A0..A{P-1}, then one combined group per iteration).
#pragma unroll
  for (int u = 0; u < kPrefetch; ++u) { issue_weights(u, kb_of(u)); cp_async_commit(); }
  pdl_wait();
  pdl_launch_dependents();
#pragma unroll
  for (int u = 0; u < kPrefetch; ++u) { issue_acts(u, kb_of(u)); cp_async_commit(); }

  float acc[NACC];
#pragma unroll
  for (int i = 0; i < NACC; ++i) acc[i] = 0.f;

  auto compute = [&](const Stage<BITS, MB, NH>& s) {
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) {
      uint32_t a[4];
      if (BITS == 4) dequant4<__half>(s.w[jj], a);
      else dequant8<__half>(s.w[(2 * jj) % WPL], s.w[(2 * jj + 1) % WPL], a);
      if (GROUPED) {
        const uint32_t s0 = bcast16(s.s[jj]), s1 = bcast16_hi(s.s[jj]);   // rows 0-7 (n), rows 8-15 (n + 4)
        a[0] = Pack2<__half>::mul(a[0], s0);
        a[2] = Pack2<__half>::mul(a[2], s0);
        a[1] = Pack2<__half>::mul(a[1], s1);
        a[3] = Pack2<__half>::mul(a[3], s1);
      }
      const uint32_t e = s.e[jj >> 1];
#pragma unroll
      for (int mb = 0; mb < MB; ++mb)
#pragma unroll
        for (int h = 0; h < NH; ++h) {
          float* c = &acc[((mb * 4 + jj) * 2 + h) * 4];
          if (jj & 1) mma_sp<1>(c, a, s.b[mb][h], e);
          else mma_sp<0>(c, a, s.b[mb][h], e);
        }
    }
  };

  {
    int slot = 0;
    for (int i = 0; i < n_it; ++i) {
      const bool live = kb_of(i) < kb1;      // the last round is ragged when the split is not a multiple of kWarps
      cp_async_wait<kPrefetch - 1>();        // the oldest stage of this lane has landed ...
      __syncwarp();                          // ... and of every other lane (the activation rows are shared by the warp)
      Stage<BITS, MB, NH> st_;
      read_stage(st_, slot);
      __syncwarp();                          // everybody has read the slot before anybody refills it
      const int kn = kb_of(i + kPrefetch);   // refill the slot just read (same lane: no hazard)
      issue_weights(slot, kn);
      issue_acts(slot, kn);
      cp_async_commit();                     // one group per iteration, issued or not, so that the wait always counts stages
      if (live) compute(st_);
      if (++slot == kPrefetch) slot = 0;
    }
    cp_async_wait<0>();
  }
  __syncthreads();   // the ring is reused as the reduction buffer below

  // ---- sum the 8 warps' partials, un-permute into a [MC][64] tile ------------------------------------------------------
  float* red = reinterpret_cast<float*>(smem24);
  float* tile = red + kWarps * NACC * 32;
#pragma unroll
  for (int i = 0; i < NACC; ++i) red[(warp * NACC + i) * 32 + lane] = acc[i];
  __syncthreads();
  for (int idx = threadIdx.x; idx < NACC * 32; idx += kThreads) {
    const int a = idx >> 5, l = idx & 31;
    float v = 0.f;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) v += red[(w * NACC + a) * 32 + l];
    const int ci = a & 3, h = (a >> 2) & 1, jj = (a >> 3) & 3, mb = a >> 5;
    const int n_local = 8 * (l >> 2) + jj + 4 * (ci >> 1);
    const int m_local = 16 * mb + 8 * h + 2 * (l & 3) + (ci & 1);
    tile[m_local * 64 + n_local] = v;
  }
  __syncthreads();

  const int m_valid = min(MC, args.M - m0);
  __half* cdst = args.C + (size_t)m0 * args.N + (size_t)t * 64;
  auto finish = [&](int idx, float v) {
    const int m = idx >> 6, n = idx & 63;
    if (!GROUPED) v *= __half2float(args.scales[(size_t)t * 64 + n]);
    cdst[(size_t)m * args.N + n] = __float2half_rn(v);
  };
  if (args.splits == 1) {
    for (int idx = threadIdx.x; idx < m_valid * 64; idx += kThreads) finish(idx, tile[idx]);
    return;
  }
  // ---- K splits: fp32 slots + ticket; the last split to arrive sums the slots in split order ------------------------------
  __shared__ int s_last;
  const int tile_id = blockIdx.y * args.n_tiles + t;
  float* slots = args.scratch + (size_t)tile_id * args.splits * (MC * 64);
  for (int idx = threadIdx.x; idx < m_valid * 64; idx += kThreads) __stcg(slots + (size_t)sp * (MC * 64) + idx, tile[idx]);
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const int old = atomicAdd(&args.counters[tile_id], 1);
    s_last = (old == args.splits - 1);
    if (s_last) args.counters[tile_id] = 0;
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  for (int idx = threadIdx.x; idx < m_valid * 64; idx += kThreads) {
    float v = 0.f;
    for (int s = 0; s < args.splits; ++s) v += __ldcg(slots + (size_t)s * (MC * 64) + idx);
    finish(idx, v);
  }
}

// ---- host side -------------------------------------------------------------------------------------------------------
template <int BITS, int MB, bool GROUPED, int NH>
static int launch_m24(m24::Args a, cudaStream_t st) {
  using namespace m24;
  constexpr int MC = 16 * MB;
  const int KB = a.K / 32;
  const int m_chunks = (a.M + MC - 1) / MC;
  a.n_tiles = a.N / 64;
  // ~4 CTAs per SM; a split is at least 8 k32 blocks per warp (amortises the CTA prologue / reduction) and a whole
  // number of scale groups
  const int target = num_sms() * 4;
  int splits = target / (a.n_tiles * m_chunks);
  int min_kb = 8 * kWarps;
  if (GROUPED) { const int gb = (a.K / a.num_groups) / 32; if (min_kb % gb) min_kb = ((min_kb + gb - 1) / gb) * gb; }
  if (splits > KB / min_kb) splits = KB / min_kb;
  if (splits < 1) splits = 1;
  int kbps = (KB + splits - 1) / splits;
  if (GROUPED) { const int gb = (a.K / a.num_groups) / 32; kbps = ((kbps + gb - 1) / gb) * gb; }
  splits = (KB + kbps - 1) / kbps;
  a.splits = splits;
  a.kb_per_split = kbps;
  int dev = 0;
  cudaGetDevice(&dev);
  GemmScratch sc;
  if (int e = gemm_ensure_scratch(dev, splits > 1 ? (size_t)a.n_tiles * m_chunks * splits * MC * 64 : 1, a.n_tiles * m_chunks, st, &sc)) return e;
  a.scratch = sc.partials;
  a.counters = sc.counters;
  const size_t smem_red = (size_t)(kWarps * 32 * MB * 32 + MC * 64) * 4;
  const size_t smem_ring = (size_t)kWarps * prefetch_depth(BITS, MB, NH) * slot_bytes(BITS, MB, NH);
  const size_t smem = smem_red > smem_ring ? smem_red : smem_ring;
  auto kern = marlin24_gemm_kernel<BITS, MB, GROUPED, NH>;
  static bool attr_set[64] = {};
  if (!attr_set[dev & 63]) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    attr_set[dev & 63] = true;
  }
  cudaError_t e = launch_pdl(kern, dim3(a.n_tiles * splits, m_chunks), dim3(kThreads), smem, st, a);
  if (e != cudaSuccess) return set_error(B200_ECUDA, "gptq_marlin_24_gemm launch: %s", cudaGetErrorString(e));
  return B200_OK;
}

template <int BITS, bool GROUPED>
static int dispatch_m24(const m24::Args& a, cudaStream_t st) {
  if (a.M <= 8) return launch_m24<BITS, 1, GROUPED, 1>(a, st);
  if (a.M <= 16) return launch_m24<BITS, 1, GROUPED, 2>(a, st);
  return launch_m24<BITS, 2, GROUPED, 2>(a, st);   // M > 32: passes of 32 rows (grid.y), weights re-read through L2
}

}  // namespace b200

extern "C" int b200_gptq_marlin_24_gemm(void* c, const void* a, const int32_t* b_q_weight, const int16_t* b_meta,
                                        const void* b_scales, int32_t* workspace, int64_t workspace_numel, int num_bits,
                                        int64_t size_m, int64_t size_n, int64_t size_k, int64_t lda, int num_groups,
                                        int dtype, void* stream) {
  using namespace b200;
  (void)workspace;
  // preconditions of the reference op (marlin_24_cuda_kernel.cu:1026-1115)
  B200_CHECK(num_bits == 4 || num_bits == 8, "num_bits must be 4 or 8. Got = %d", num_bits);
  B200_CHECK(dtype == B200_F16, "gptq_marlin_24_gemm only supports float16 activations (gptq_marlin_24.py:145-147)");
  B200_CHECK(size_k % 128 == 0, "size_k = %lld is not divisible by tile_k_size = 128", (long long)size_k);
  B200_CHECK(size_n % 128 == 0, "size_n = %lld is not divisible by tile_n_size = 128", (long long)size_n);
  B200_CHECK(num_groups >= 1 && size_k % num_groups == 0, "size_k = %lld is not divisible by b_scales.size(0) = %d",
             (long long)size_k, num_groups);
  const int gs = (int)(size_k / num_groups);
  if (num_groups > 1) B200_CHECK(gs == 128, "Unexpected group_size = %d (Sparse-Marlin supports -1 and 128)", gs);
  B200_CHECK(workspace_numel >= (size_n / 128) * 64, "workspace.numel = %lld is below min_workspace_size = %lld",
             (long long)workspace_numel, (long long)((size_n / 128) * 64));
  B200_CHECK((reinterpret_cast<uintptr_t>(a) & 15) == 0 && (lda * 2) % 16 == 0, "A must be 16-byte aligned with a 16-byte multiple row pitch");
  if (size_m == 0) return B200_OK;
  m24::Args ga;
  ga.A = reinterpret_cast<const __half*>(a);
  ga.Bq = reinterpret_cast<const uint32_t*>(b_q_weight);
  ga.meta = reinterpret_cast<const uint32_t*>(b_meta);
  ga.scales = reinterpret_cast<const __half*>(b_scales);
  ga.C = reinterpret_cast<__half*>(c);
  ga.scratch = nullptr;
  ga.counters = nullptr;
  ga.M = (int)size_m; ga.N = (int)size_n; ga.K = (int)size_k;
  ga.lda = lda;
  ga.num_groups = num_groups;
  ga.splits = 1; ga.kb_per_split = 0; ga.n_tiles = 0;
  cudaStream_t st = (cudaStream_t)stream;
  int rc;
  if (num_groups > 1) rc = num_bits == 4 ? dispatch_m24<4, true>(ga, st) : dispatch_m24<8, true>(ga, st);
  else rc = num_bits == 4 ? dispatch_m24<4, false>(ga, st) : dispatch_m24<8, false>(ga, st);
  if (rc) return rc;
  B200_LAUNCH_CHECK("gptq_marlin_24_gemm");
  return B200_OK;
}
