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
// Work decomposition (HBM-bound, M <= 64 per pass).  CTA = (four adjacent 64-column tiles, K split s of S, 64-row pass):
// warp w streams the weights / metadata / scales of tile w % 4 lane-privately (cp.async into a multi-stage shared-memory
// ring, read back by the same lane); the activation rows of the k32 block are copied ONCE per CTA (16-byte cp.async, rows
// 80 bytes apart) and every warp takes its B fragments from them with one ldmatrix.x4 per 8-row half block.  One
// __syncthreads per k32 block (classic multistage pipeline).  With more than 32 rows a second group of four warps works on
// rows 32..63 of the same tiles from the same staged weights, so the weights cross L2 once per 64 rows.  No cross-warp
// reduction: every warp owns its (rows, tile) outputs; K splits are summed through fp32 slots in split order by the last
// split to arrive (atomic ticket): deterministic, nobody spins.
// History (profiles/r2_m24_ncu.txt): the first version gave every warp its own k32 blocks and loaded its activation
// fragments with LDGs -- 105 MB (M = 16) to 470 MB (M = 64) of L2->SM activation traffic against 46 MB of weights, and every
// iteration stalled ~2000 clk on the first HMMA.SP.  NOTE this kernel uses the warp-level sparse MMA (HMMA.SP); feeding
// tcgen05.mma.sp (dequantised tile in shared memory, metadata in TMEM) is not done.
#include "marlin_gemm_kernel.cuh"

namespace b200 {
namespace m24 {


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
  int n_tiles;      // 64-column tiles
  int n_ctiles;     // CTA columns (four tiles each)
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

// shared-memory stage of one k32 block for a CTA: [4 tiles x 32 lanes x weights][4 x 32 x 8 B metadata][4 x 8 cols x 16 B
// scales][MC activation rows x 80 B]
constexpr int kActRow = 80;   // odd multiple of 16 bytes: the 8 rows of an ldmatrix phase hit distinct banks
constexpr int w_bytes(int bits) { return 32 * (bits == 4 ? 16 : 32); }
constexpr int e_off(int bits) { return 4 * w_bytes(bits); }
constexpr int s_off(int bits) { return e_off(bits) + 4 * 256; }
constexpr int a_off(int bits) { return s_off(bits) + 4 * 128; }
constexpr int stage_bytes(int bits, int mc) { return a_off(bits) + mc * kActRow; }
// ring depth: <= 54 KB per 128-thread CTA (4 CTAs per SM), <= 100 KB per 256-thread CTA (2 per SM)
constexpr int ring_depth(int bits, int mc, int mh) {
  int d = ((mh == 1 ? 54 : 100) * 1024) / stage_bytes(bits, mc);
  return d > 8 ? 8 : d;
}

}  // namespace m24

// MB = 16-row blocks per warp, NH = 8-row halves of a block that carry real rows (1 when M <= 8: half the tensor-core
// work), MH = groups of four warps (each group takes 16 MB rows of the same four tiles)
template <int BITS, int MB, bool GROUPED, int NH, int MH>
__global__ void __launch_bounds__(128 * MH, MH == 1 ? 4 : 2)
marlin24_gemm_kernel(const m24::Args args) {
  using namespace m24;
  constexpr int WPL = BITS == 4 ? 4 : 8;   // packed int32 per lane per (k32 block, 64-column tile)
  constexpr int MCW = 16 * MB;             // activation rows per warp
  constexpr int MC = MCW * MH;             // activation rows per CTA pass
  constexpr int NACC = 32 * MB;            // fp32 accumulators per thread
  constexpr int THREADS = 128 * MH;
  constexpr int WB = w_bytes(BITS), E_OFF = e_off(BITS), S_OFF = s_off(BITS), A_OFF = a_off(BITS);
  constexpr int STAGE = stage_bytes(BITS, MC);
  constexpr int P = ring_depth(BITS, MC, MH);
  static_assert(P >= 3, "ring too shallow");
  extern __shared__ __align__(16) uint8_t smem24[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tw = warp & 3, mh = warp >> 2;
  const int col = lane >> 2, q = lane & 3;
  const int ct = blockIdx.x / args.splits, sp = blockIdx.x - ct * args.splits;
  const int t = ct * 4 + tw;
  const bool tile_ok = t < args.n_tiles;   // N is a multiple of 128, not of 256: the last CTA may own two tiles
  const int m0 = blockIdx.y * MC;
  const int KB = args.K / 32;
  const int kb0 = sp * args.kb_per_split;
  const int kb1 = min(KB, kb0 + args.kb_per_split);
  const int n_it = kb1 - kb0;
  const int wrow = args.N * (BITS == 4 ? 2 : 4);

  const uint32_t smem_base = ptx::smem_u32(smem24);
  // weights, metadata, scales: group 0 of warps copies them (lane-private slots; the other group reads the same slots)
  const bool w_issuer = tile_ok && mh == 0;
  const uint32_t* wp = args.Bq + (size_t)t * (32 * WPL) + lane * WPL;
  const uint32_t* mp = args.meta + (size_t)t * 64 + lane;
  const __half* sp_ptr = args.scales + (size_t)t * 64 + col * 8;
  const uint32_t w_dst = smem_base + tw * WB + lane * (WPL * 4);
  const uint32_t e_dst = smem_base + E_OFF + tw * 256 + lane * 8;
  const uint32_t s_dst = smem_base + S_OFF + tw * 128 + col * 16;
  // activations: thread c copies 16-byte chunk c & 3 of row c >> 2 of the pass (rows beyond M are zero-filled)
  constexpr int A_ROWS = NH == 1 ? 8 : MC;
  const int a_row = threadIdx.x >> 2;
  const bool a_issuer = a_row < A_ROWS;
  const uint32_t a_bytes = (a_issuer && m0 + a_row < args.M) ? 16u : 0u;
  const __half* a_src = args.A + (size_t)(a_bytes ? m0 + a_row : 0) * args.lda + (threadIdx.x & 3) * 8;
  const uint32_t a_dst = smem_base + A_OFF + a_row * kActRow + (threadIdx.x & 3) * 16;
  const uint32_t a_frag = smem_base + A_OFF + (mh * MCW + (lane & 7)) * kActRow + (lane >> 3) * 16;

  auto issue_weights = [&](int slot, int kb) {
    if (w_issuer) {
      const uint32_t so = slot * STAGE;
      const uint32_t* w = wp + (size_t)kb * wrow;
      cp_async16(w_dst + so, w);
      if (BITS == 8) cp_async16(w_dst + so + 16, w + 4);
      cp_async4(e_dst + so, mp + (size_t)kb * args.N);
      cp_async4(e_dst + so + 4, mp + (size_t)kb * args.N + 32);
      if (GROUPED && q == 0) cp_async16(s_dst + so, sp_ptr + (size_t)(kb >> 2) * args.N);   // group_size 128 = 4 k32 blocks
    }
  };
  auto issue_acts = [&](int slot, int kb) {
    if (a_issuer) cp_async16_zfill(a_dst + slot * STAGE, a_src + kb * 32, a_bytes);
  };

  // weights / metadata / scales do not depend on the previous kernel: start streaming them before the PDL wait; the
  // activations follow as P-1 more groups.  wait_group<P-2> in iteration i then covers both halves of stage i (groups
  // retire in order: W0..W{P-2}, A0..A{P-2}, then one combined group per iteration).
#pragma unroll
  for (int u = 0; u < P - 1; ++u) { if (u < n_it) issue_weights(u, kb0 + u); cp_async_commit(); }
  pdl_wait();
  pdl_launch_dependents();
#pragma unroll
  for (int u = 0; u < P - 1; ++u) { if (u < n_it) issue_acts(u, kb0 + u); cp_async_commit(); }

  float acc[NACC];
#pragma unroll
  for (int i = 0; i < NACC; ++i) acc[i] = 0.f;

  int slot = 0, slot_fill = P - 1;
  for (int i = 0; i < n_it; ++i) {
    cp_async_wait<P - 2>();   // this thread's copies of stage i have landed ...
    __syncthreads();          // ... and everybody else's; everybody is also done reading the slot refilled below
    if (i + P - 1 < n_it) { issue_weights(slot_fill, kb0 + i + P - 1); issue_acts(slot_fill, kb0 + i + P - 1); }
    cp_async_commit();        // one group per iteration, issued or not, so that the wait always counts stages
    if (tile_ok) {
      const uint8_t* st = smem24 + slot * STAGE;
      uint32_t w[WPL], s4[4] = {0, 0, 0, 0};
      const uint4 v0 = *reinterpret_cast<const uint4*>(st + tw * WB + lane * (WPL * 4));
      w[0] = v0.x; w[1] = v0.y; w[2] = v0.z; w[3] = v0.w;
      if (BITS == 8) {
        const uint4 v1 = *reinterpret_cast<const uint4*>(st + tw * WB + lane * (WPL * 4) + 16);
        w[4 % WPL] = v1.x; w[5 % WPL] = v1.y; w[6 % WPL] = v1.z; w[7 % WPL] = v1.w;
      }
      const uint2 e2 = *reinterpret_cast<const uint2*>(st + E_OFF + tw * 256 + lane * 8);
      if (GROUPED) {
        const uint4 sv = *reinterpret_cast<const uint4*>(st + S_OFF + tw * 128 + col * 16);
        s4[0] = sv.x; s4[1] = sv.y; s4[2] = sv.z; s4[3] = sv.w;
      }
      uint32_t b[MB][NH][4];
#pragma unroll
      for (int mb = 0; mb < MB; ++mb)
#pragma unroll
        for (int h = 0; h < NH; ++h) ldmatrix_x4(b[mb][h], a_frag + slot * STAGE + (16 * mb + 8 * h) * kActRow);
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) {
        uint32_t a[4];
        if (BITS == 4) dequant4<__half>(w[jj], a);
        else dequant8<__half>(w[(2 * jj) % WPL], w[(2 * jj + 1) % WPL], a);
        if (GROUPED) {
          const uint32_t s0 = bcast16(s4[jj]), s1 = bcast16_hi(s4[jj]);   // rows 0-7 (n), rows 8-15 (n + 4)
          a[0] = Pack2<__half>::mul(a[0], s0);
          a[2] = Pack2<__half>::mul(a[2], s0);
          a[1] = Pack2<__half>::mul(a[1], s1);
          a[3] = Pack2<__half>::mul(a[3], s1);
        }
        const uint32_t e = jj < 2 ? e2.x : e2.y;
#pragma unroll
        for (int mb = 0; mb < MB; ++mb)
#pragma unroll
          for (int h = 0; h < NH; ++h) {
            float* c = &acc[((mb * 4 + jj) * 2 + h) * 4];
            if (jj & 1) mma_sp<1>(c, a, b[mb][h], e);
            else mma_sp<0>(c, a, b[mb][h], e);
          }
      }
    }
    if (++slot == P) slot = 0;
    if (++slot_fill == P) slot_fill = 0;
  }
  cp_async_wait<0>();
  __syncthreads();   // the ring is reused as the un-permute buffer below

  // ---- every warp un-permutes its own [MCW][64] outputs through shared memory ---------------------------------------------
  constexpr int TP = 65;   // row pitch in floats
  float* tile = reinterpret_cast<float*>(smem24) + warp * (MCW * TP);
#pragma unroll
  for (int a = 0; a < NACC; ++a) {
    const int ci = a & 3, h = (a >> 2) & 1, jj = (a >> 3) & 3, mb = a >> 5;
    if (NH == 1 && h == 1) continue;
    const int n_local = 8 * col + jj + 4 * (ci >> 1);
    const int m_local = 16 * mb + 8 * h + 2 * q + (ci & 1);
    tile[m_local * TP + n_local] = acc[a];
  }
  __syncwarp();

  const int mw0 = m0 + mh * MCW;                       // first row of this warp
  const int m_valid = max(0, min(NH == 1 ? 8 : MCW, args.M - mw0));
  auto finish = [&](int r, int n, float v) {
    if (!GROUPED) v *= __half2float(args.scales[(size_t)t * 64 + n]);
    args.C[(size_t)(mw0 + r) * args.N + (size_t)t * 64 + n] = __float2half_rn(v);
  };
  if (args.splits == 1) {
    if (tile_ok)
      for (int idx = lane; idx < m_valid * 64; idx += 32) finish(idx >> 6, idx & 63, tile[(idx >> 6) * TP + (idx & 63)]);
    return;
  }
  // ---- K splits: fp32 slots + ticket; the last split to arrive sums the slots in split order ------------------------------
  __shared__ int s_last;
  const int cta_id = blockIdx.y * args.n_ctiles + ct;
  float* cta_slots = args.scratch + (size_t)cta_id * args.splits * (MC * 256);
  if (tile_ok) {
    float* mine = cta_slots + (size_t)sp * (MC * 256) + (size_t)(mh * MCW) * 256 + tw * 64;
    for (int idx = lane; idx < m_valid * 64; idx += 32) __stcg(mine + (idx >> 6) * 256 + (idx & 63), tile[(idx >> 6) * TP + (idx & 63)]);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();   // release: cumulative over the CTA's slot writes ordered before it by the barrier
    const int old = atomicAdd(&args.counters[cta_id], 1);
    s_last = (old == args.splits - 1);
    if (s_last) args.counters[cta_id] = 0;
    __threadfence();   // acquire
  }
  __syncthreads();
  if (!s_last) return;
  // the whole CTA sums the [rows][256] slot planes, four columns per thread, four splits in flight per accumulator (the
  // ncu capture of a per-warp scalar loop showed this tail as the top stall of the kernel)
  const int rows = min(NH == 1 ? 8 : MC, args.M - m0);
  const int n_cols = min(256, args.N - ct * 256);
  for (int idx = threadIdx.x; idx < rows * 64; idx += THREADS) {
    const int r = idx >> 6, n = (idx & 63) * 4;
    if (n >= n_cols) continue;
    const float* src = cta_slots + r * 256 + n;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    int s2 = 0;
    for (; s2 + 4 <= args.splits; s2 += 4) {
      const float4 x0 = __ldcg(reinterpret_cast<const float4*>(src + (size_t)(s2 + 0) * (MC * 256)));
      const float4 x1 = __ldcg(reinterpret_cast<const float4*>(src + (size_t)(s2 + 1) * (MC * 256)));
      const float4 x2 = __ldcg(reinterpret_cast<const float4*>(src + (size_t)(s2 + 2) * (MC * 256)));
      const float4 x3 = __ldcg(reinterpret_cast<const float4*>(src + (size_t)(s2 + 3) * (MC * 256)));
      v.x += x0.x; v.y += x0.y; v.z += x0.z; v.w += x0.w;
      v.x += x1.x; v.y += x1.y; v.z += x1.z; v.w += x1.w;
      v.x += x2.x; v.y += x2.y; v.z += x2.z; v.w += x2.w;
      v.x += x3.x; v.y += x3.y; v.z += x3.z; v.w += x3.w;
    }
    for (; s2 < args.splits; ++s2) {
      const float4 x0 = __ldcg(reinterpret_cast<const float4*>(src + (size_t)s2 * (MC * 256)));
      v.x += x0.x; v.y += x0.y; v.z += x0.z; v.w += x0.w;
    }
    const size_t nn = (size_t)ct * 256 + n;
    if (!GROUPED) {
      const __half2 c0 = *reinterpret_cast<const __half2*>(args.scales + nn), c1 = *reinterpret_cast<const __half2*>(args.scales + nn + 2);
      v.x *= __low2float(c0); v.y *= __high2float(c0); v.z *= __low2float(c1); v.w *= __high2float(c1);
    }
    const __half2 o0 = __floats2half2_rn(v.x, v.y), o1 = __floats2half2_rn(v.z, v.w);
    uint2 o;
    o.x = *reinterpret_cast<const uint32_t*>(&o0);
    o.y = *reinterpret_cast<const uint32_t*>(&o1);
    *reinterpret_cast<uint2*>(args.C + (size_t)(m0 + r) * args.N + nn) = o;
  }
}

// ---- host side -------------------------------------------------------------------------------------------------------
template <int BITS, int MB, bool GROUPED, int NH, int MH>
static int launch_m24(m24::Args a, cudaStream_t st) {
  using namespace m24;
  constexpr int MC = 16 * MB * MH;
  const int KB = a.K / 32;
  const int m_chunks = (a.M + MC - 1) / MC;
  a.n_tiles = a.N / 64;
  a.n_ctiles = (a.n_tiles + 3) / 4;
  // one wave of resident CTAs (4 per SM at 128 threads, 2 at 256); a split is at least 16 k32 blocks (amortises the prologue
  // and the slot traffic) and a whole number of scale groups
  const int target = num_sms() * (MH == 1 ? 4 : 2);
  int splits = target / (a.n_ctiles * m_chunks);
  const int min_kb = 16;
  if (splits > KB / min_kb) splits = KB / min_kb;
  if (splits < 1) splits = 1;
  int kbps = (KB + splits - 1) / splits;
  if (GROUPED) kbps = ((kbps + 3) / 4) * 4;   // group_size 128 = 4 k32 blocks
  splits = (KB + kbps - 1) / kbps;
  a.splits = splits;
  a.kb_per_split = kbps;
  int dev = 0;
  cudaGetDevice(&dev);
  GemmScratch sc;
  if (int e = gemm_ensure_scratch(dev, splits > 1 ? (size_t)a.n_ctiles * m_chunks * splits * MC * 256 : 1, a.n_ctiles * m_chunks, st, &sc)) return e;
  a.scratch = sc.partials;
  a.counters = sc.counters;
  const size_t smem_tile = (size_t)(4 * MH) * (16 * MB) * 65 * 4;
  const size_t smem_ring = (size_t)ring_depth(BITS, MC, MH) * stage_bytes(BITS, MC);
  const size_t smem = smem_tile > smem_ring ? smem_tile : smem_ring;
  auto kern = marlin24_gemm_kernel<BITS, MB, GROUPED, NH, MH>;
  static bool attr_set[64] = {};
  if (!attr_set[dev & 63]) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    attr_set[dev & 63] = true;
  }
  cudaError_t e = launch_pdl(kern, dim3(a.n_ctiles * splits, m_chunks), dim3(128 * MH), smem, st, a);
  if (e != cudaSuccess) return set_error(B200_ECUDA, "gptq_marlin_24_gemm launch: %s", cudaGetErrorString(e));
  return B200_OK;
}

template <int BITS, bool GROUPED>
static int dispatch_m24(const m24::Args& a, cudaStream_t st) {
  if (a.M <= 8) return launch_m24<BITS, 1, GROUPED, 1, 1>(a, st);
  if (a.M <= 16) return launch_m24<BITS, 1, GROUPED, 2, 1>(a, st);
  if (a.M <= 32) return launch_m24<BITS, 2, GROUPED, 2, 1>(a, st);
  return launch_m24<BITS, 2, GROUPED, 2, 2>(a, st);   // M > 64: passes of 64 rows (grid.y), weights re-read through L2
}

// ---- allocation-free variant (b200_gptq_marlin_24_gemm_static) ------------------------------------------------------------
// The K-split slots and tile counters live in file-scope device arrays instead of the grow-on-demand scratch of
// gemm_ensure_scratch(): nothing is allocated, ever.  Every GPU that loads the library holds the 20 MiB + 16 KiB until exit.
// All calls share the arrays: calls that overlap on the GPU (two streams, or two threads on one device) race.
constexpr size_t kStaticPartialFloats = B200_M24_STATIC_MAX_PARTIAL_FLOATS;   // 5 Mi floats = 20 MiB
constexpr int kStaticCounters = B200_M24_STATIC_MAX_COUNTERS;                 // 4096 ints
static __device__ __align__(256) float m24_static_partials[kStaticPartialFloats];
static __device__ __align__(256) int m24_static_counters[kStaticCounters];

// In host code the two names above are host-side stand-ins, not device addresses: the addresses a kernel can use come from
// cudaGetSymbolAddress, once per device, with that device current.
static int m24_static_scratch(int dev, float** partials, int** counters) {
  static float* p_cache[64];
  static int* c_cache[64];
  static bool have[64];
  if (dev < 0 || dev >= 64) return set_error(B200_EINVAL, "gptq_marlin_24_gemm_static: device index %d out of range", dev);
  if (!have[dev]) {
    float* p = nullptr;
    int* c = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&p, m24_static_partials);
    if (e != cudaSuccess) return set_error(B200_ECUDA, "gptq_marlin_24_gemm_static: cudaGetSymbolAddress(partials): %s", cudaGetErrorString(e));
    e = cudaGetSymbolAddress((void**)&c, m24_static_counters);
    if (e != cudaSuccess) return set_error(B200_ECUDA, "gptq_marlin_24_gemm_static: cudaGetSymbolAddress(counters): %s", cudaGetErrorString(e));
    p_cache[dev] = p;
    c_cache[dev] = c;
    have[dev] = true;
  }
  *partials = p_cache[dev];
  *counters = c_cache[dev];
  return B200_OK;
}

// same work decomposition and kernel as launch_m24; only the origin of the scratch differs
template <int BITS, int MB, bool GROUPED, int NH, int MH>
static int launch_m24_static(m24::Args a, cudaStream_t st) {
  using namespace m24;
  constexpr int MC = 16 * MB * MH;
  const int KB = a.K / 32;
  const int m_chunks = (a.M + MC - 1) / MC;
  a.n_tiles = a.N / 64;
  a.n_ctiles = (a.n_tiles + 3) / 4;
  const int target = num_sms() * (MH == 1 ? 4 : 2);
  int splits = target / (a.n_ctiles * m_chunks);
  const int min_kb = 16;
  if (splits > KB / min_kb) splits = KB / min_kb;
  if (splits < 1) splits = 1;
  int kbps = (KB + splits - 1) / splits;
  if (GROUPED) kbps = ((kbps + 3) / 4) * 4;
  splits = (KB + kbps - 1) / kbps;
  a.splits = splits;
  a.kb_per_split = kbps;
  int dev = 0;
  cudaGetDevice(&dev);
  if (int e = m24_static_scratch(dev, &a.scratch, &a.counters)) return e;
  if (splits > 1) {
    const size_t need = (size_t)a.n_ctiles * m_chunks * splits * MC * 256;
    const size_t need_counters = (size_t)a.n_ctiles * m_chunks;
    if (need > kStaticPartialFloats || need_counters > (size_t)kStaticCounters)
      return set_error(B200_EINVAL,
                       "gptq_marlin_24_gemm_static: M=%d N=%d K=%d needs %zu fp32 K-split slots and %zu tile counters; the static "
                       "scratch holds %zu and %d (use b200_gptq_marlin_24_gemm)",
                       a.M, a.N, a.K, need, need_counters, kStaticPartialFloats, kStaticCounters);
    // the counters must start at zero; the kernel leaves them zero, but an aborted earlier call may not have.  The slots
    // are written before they are read, so they need no clearing.
    cudaError_t e = cudaMemsetAsync(a.counters, 0, need_counters * sizeof(int), st);
    if (e != cudaSuccess) return set_error(B200_ECUDA, "gptq_marlin_24_gemm_static: clearing the tile counters: %s", cudaGetErrorString(e));
  }
  const size_t smem_tile = (size_t)(4 * MH) * (16 * MB) * 65 * 4;
  const size_t smem_ring = (size_t)ring_depth(BITS, MC, MH) * stage_bytes(BITS, MC);
  const size_t smem = smem_tile > smem_ring ? smem_tile : smem_ring;
  auto kern = marlin24_gemm_kernel<BITS, MB, GROUPED, NH, MH>;
  static bool attr_set[64] = {};
  if (!attr_set[dev & 63]) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    attr_set[dev & 63] = true;
  }
  cudaError_t e = launch_pdl(kern, dim3(a.n_ctiles * splits, m_chunks), dim3(128 * MH), smem, st, a);
  if (e != cudaSuccess) return set_error(B200_ECUDA, "gptq_marlin_24_gemm_static launch: %s", cudaGetErrorString(e));
  return B200_OK;
}

template <int BITS, bool GROUPED>
static int dispatch_m24_static(const m24::Args& a, cudaStream_t st) {
  if (a.M <= 8) return launch_m24_static<BITS, 1, GROUPED, 1, 1>(a, st);
  if (a.M <= 16) return launch_m24_static<BITS, 1, GROUPED, 2, 1>(a, st);
  if (a.M <= 32) return launch_m24_static<BITS, 2, GROUPED, 2, 1>(a, st);
  return launch_m24_static<BITS, 2, GROUPED, 2, 2>(a, st);
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
  ga.splits = 1; ga.kb_per_split = 0; ga.n_tiles = 0; ga.n_ctiles = 0;
  cudaStream_t st = (cudaStream_t)stream;
  int rc;
  if (num_groups > 1) rc = num_bits == 4 ? dispatch_m24<4, true>(ga, st) : dispatch_m24<8, true>(ga, st);
  else rc = num_bits == 4 ? dispatch_m24<4, false>(ga, st) : dispatch_m24<8, false>(ga, st);
  if (rc) return rc;
  B200_LAUNCH_CHECK("gptq_marlin_24_gemm");
  return B200_OK;
}

// Allocation-free variant of b200_gptq_marlin_24_gemm (no counterpart in the reference; see include/b200_ops.h).  Same
// arguments, preconditions, kernel and results; the K-split scratch is the file-scope device array above.  Calls no
// cudaMalloc / cudaFree / allocating library routine.
extern "C" int b200_gptq_marlin_24_gemm_static(void* c, const void* a, const int32_t* b_q_weight, const int16_t* b_meta,
                                               const void* b_scales, int32_t* workspace, int64_t workspace_numel,
                                               int num_bits, int64_t size_m, int64_t size_n, int64_t size_k, int64_t lda,
                                               int num_groups, int dtype, void* stream) {
  using namespace b200;
  (void)workspace;
  B200_CHECK(num_bits == 4 || num_bits == 8, "num_bits must be 4 or 8. Got = %d", num_bits);
  B200_CHECK(dtype == B200_F16, "gptq_marlin_24_gemm_static only supports float16 activations (gptq_marlin_24.py:145-147)");
  B200_CHECK(size_m >= 0 && size_m <= INT32_MAX && size_n > 0 && size_n <= INT32_MAX && size_k > 0 && size_k <= INT32_MAX,
             "sizes out of range: m=%lld n=%lld k=%lld", (long long)size_m, (long long)size_n, (long long)size_k);
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
  ga.splits = 1; ga.kb_per_split = 0; ga.n_tiles = 0; ga.n_ctiles = 0;
  cudaStream_t st = (cudaStream_t)stream;
  int rc;
  if (num_groups > 1) rc = num_bits == 4 ? dispatch_m24_static<4, true>(ga, st) : dispatch_m24_static<8, true>(ga, st);
  else rc = num_bits == 4 ? dispatch_m24_static<4, false>(ga, st) : dispatch_m24_static<8, false>(ga, st);
  if (rc) return rc;
  B200_LAUNCH_CHECK("gptq_marlin_24_gemm_static");
  return B200_OK;
}
