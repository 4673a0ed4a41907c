// Prefill attention over the paged KV cache ("prefix" / chunked prefill): every query token of a sequence attends to
// the sequence's CONTEXT (tokens already in the block-table cache, no mask) and to the new tokens of the same call up
// to and including itself (causal), with optional sliding window or ALiBi bias.
//
// Replaces the reference's Triton kernels `_fwd_kernel` / `_fwd_kernel_alibi` behind `context_attention_fwd`
// (vllm/attention/ops/prefix_prefill.py:14-435, :437-672, launcher :674-799; caller PagedAttention.forward_prefix,
// vllm/attention/ops/paged_attn.py:185-216).  Same arguments and masking rules: scale = 1/sqrt(head_size) (:696), keys of
// the context beyond b_ctx_len and new keys beyond the query length are excluded, the sliding window replaces logits
// with -10000 after scaling (:120-133), ALiBi adds slope * (key_pos - query_pos) (:551-557), probabilities are rounded
// to the value dtype before P.V (:169).  Cache layouts are the decode kernels': K [NB, Hkv, D/8, BS, 8], V [NB, Hkv, D, BS].
//
// Shape of the kernel: one CTA = 64 query rows of one (sequence, head), 4 warps x 16 rows, flash-attention style online
// softmax in registers.  Both cache layouts are mma.m16n8k16 fragment friendly: a K "chunk" row (8 consecutive d of one
// token, 16 bytes) is what the 4 lanes of a quad need for one k-step, and consecutive tokens of a chunk are contiguous,
// so every warp-wide 32-bit load is one 128-byte line; V rows hold consecutive tokens of one d, the k-pairs of the P.V
// B operand.  New keys / values come from the dense k / v tensors of the call.
#include "common.cuh"
#include <stdlib.h>

namespace b200 {

struct PrefillArgs {
  void* o;
  const void* q;
  const void* k;
  const void* v;
  const void* kc;
  const void* vc;
  const int* block_tables;
  const int* start_loc;
  const int* seq_lens;
  const int* ctx_lens;
  const float* alibi;
  int64_t bt_stride;
  int64_t q_s0, q_s1, k_s0, k_s1, v_s0, v_s1, o_s0, o_s1;   // element strides: token, head
  int64_t kc_bs, kc_hs, vc_bs, vc_hs;                     // cache strides: block, kv head
  int num_heads, num_kv_heads, head_size, block_size;
  int sliding_window;
  float scale;
};

template <typename T> struct MmaT;
template <> struct MmaT<__half> {
  static __device__ __forceinline__ void mma(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
  static __device__ __forceinline__ uint32_t pack(float x, float y) {
    __half2 h = __floats2half2_rn(x, y);
    return *reinterpret_cast<uint32_t*>(&h);
  }
};
template <> struct MmaT<__nv_bfloat16> {
  static __device__ __forceinline__ void mma(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
  static __device__ __forceinline__ uint32_t pack(float x, float y) {
    __nv_bfloat162 h = __floats2bfloat162_rn(x, y);
    return *reinterpret_cast<uint32_t*>(&h);
  }
};

__device__ __forceinline__ uint32_t ld32(const void* p) { return *reinterpret_cast<const uint32_t*>(p); }
__device__ __forceinline__ uint32_t ld16(const void* p) { return *reinterpret_cast<const unsigned short*>(p); }

constexpr int kPrefillRows = 64;   // query rows per CTA

template <typename T, int DPAD>
__global__ void __launch_bounds__(128) prefill_paged_attn_kernel(const PrefillArgs a) {
  pdl_wait();
  pdl_launch_dependents();
  const int seq = blockIdx.z, head = blockIdx.y;
  const int ctx = a.ctx_lens[seq];
  const int qlen = a.seq_lens[seq] - ctx;
  const int m0 = blockIdx.x * kPrefillRows;
  if (m0 >= qlen) return;
  const int start = a.start_loc[seq];
  const int kvh = head / (a.num_heads / a.num_kv_heads);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, q4 = lane & 3;
  const int D = a.head_size, BS = a.block_size, SW = a.sliding_window;
  const int row0 = m0 + 16 * warp + g;                 // this thread's rows: row0 and row0 + 8 (query index within the call)
  if (m0 + 16 * warp >= qlen) return;                  // whole warp beyond the query: nothing to do (no CTA-wide sync below)
  constexpr int KSTEPS = DPAD / 16, NDT = DPAD / 8;
  const T* Q = reinterpret_cast<const T*>(a.q);
  const T* Kn = reinterpret_cast<const T*>(a.k);
  const T* Vn = reinterpret_cast<const T*>(a.v);
  const T* KC = reinterpret_cast<const T*>(a.kc);
  const T* VC = reinterpret_cast<const T*>(a.vc);
  const int* bt = a.block_tables + (int64_t)seq * a.bt_stride;
  const float slope = a.alibi ? a.alibi[head] : 0.f;

  // Q fragments (A operand of Q.K^T): rows row0 / row0 + 8, d = 16 kk + 2 q4 (+8)
  uint32_t qa[KSTEPS][4];
#pragma unroll
  for (int kk = 0; kk < KSTEPS; ++kk) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int r = row0 + 8 * (i & 1), d = 16 * kk + 2 * q4 + 8 * (i >> 1);
      qa[kk][i] = (r < qlen && d < D) ? ld32(Q + (int64_t)(start + r) * a.q_s0 + (int64_t)head * a.q_s1 + d) : 0u;
    }
  }
  float acc[NDT][4];
#pragma unroll
  for (int i = 0; i < NDT; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.f;
  float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};

  // one 16-key tile: s[nt][0..1] = (row0, keys 8 nt + 2 q4, +1), s[nt][2..3] = (row0 + 8, same keys); logits already
  // scaled / masked (-inf = excluded).  vb(nd, half) returns the P.V B fragment register for d-tile nd, keys 8 half + 2 q4, +1
  auto softmax_pv = [&](float (&s)[2][4], auto&& vb) {
    float mx[2], alpha[2], rs[2] = {0.f, 0.f};
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      float v = fmaxf(fmaxf(s[0][2 * h], s[0][2 * h + 1]), fmaxf(s[1][2 * h], s[1][2 * h + 1]));
      v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 1));
      v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 2));
      mx[h] = fmaxf(m_run[h], v);
      const float msafe = mx[h] == -INFINITY ? 0.f : mx[h];   // a row with nothing to attend to yet
      alpha[h] = __expf(m_run[h] - msafe);
#pragma unroll
      for (int nt = 0; nt < 2; ++nt) {
        s[nt][2 * h] = __expf(s[nt][2 * h] - msafe);
        s[nt][2 * h + 1] = __expf(s[nt][2 * h + 1] - msafe);
        rs[h] += s[nt][2 * h] + s[nt][2 * h + 1];
      }
      rs[h] += __shfl_xor_sync(0xffffffffu, rs[h], 1);
      rs[h] += __shfl_xor_sync(0xffffffffu, rs[h], 2);
      l_run[h] = l_run[h] * alpha[h] + rs[h];
      m_run[h] = mx[h];
    }
    uint32_t pa[4];   // P as the A operand (16 rows x 16 keys), rounded to the value dtype like the reference
    pa[0] = MmaT<T>::pack(s[0][0], s[0][1]);
    pa[1] = MmaT<T>::pack(s[0][2], s[0][3]);
    pa[2] = MmaT<T>::pack(s[1][0], s[1][1]);
    pa[3] = MmaT<T>::pack(s[1][2], s[1][3]);
#pragma unroll
    for (int nd = 0; nd < NDT; ++nd) {
      acc[nd][0] *= alpha[0]; acc[nd][1] *= alpha[0];
      acc[nd][2] *= alpha[1]; acc[nd][3] *= alpha[1];
      MmaT<T>::mma(acc[nd], pa, vb(nd, 0), vb(nd, 1));
    }
  };
  // logit of (query row r, key position kp in the sequence) after scaling, with the window / ALiBi rules of the reference
  auto finish = [&](float raw, int r, int kpos, bool valid) -> float {
    if (!valid) return -INFINITY;
    float x = raw * a.scale;
    const int qpos = ctx + r;
    if (a.alibi) x += slope * (float)(kpos - qpos);
    if (SW > 0 && qpos - kpos >= SW) x = -10000.f;
    return x;
  };

  // ---- context: keys 0 .. ctx-1 from the paged cache, no causal mask ---------------------------------------------------
  for (int t0 = 0; t0 < ctx; t0 += 16) {
    float s[2][4];
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
      s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
      const int tok = t0 + 8 * nt + g;           // the key this thread's B fragment column holds
      if (t0 + 8 * nt < ctx) {                   // warp-uniform: the 8 keys of this n-tile are not all beyond the context
        const bool ok = tok < ctx;
        const T* kb = KC + (ok ? (int64_t)bt[tok / BS] * a.kc_bs + (int64_t)kvh * a.kc_hs + (tok % BS) * 8 : 0);
#pragma unroll
        for (int kk = 0; kk < KSTEPS; ++kk) {
          const uint32_t b0 = (ok && 16 * kk + 2 * q4 < D) ? ld32(kb + (int64_t)(2 * kk) * BS * 8 + 2 * q4) : 0u;
          const uint32_t b1 = (ok && 16 * kk + 8 + 2 * q4 < D) ? ld32(kb + (int64_t)(2 * kk + 1) * BS * 8 + 2 * q4) : 0u;
          MmaT<T>::mma(s[nt], qa[kk], b0, b1);
        }
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int kp = t0 + 8 * nt + 2 * q4 + (i & 1);
        s[nt][i] = finish(s[nt][i], row0 + 8 * (i >> 1), kp, kp < ctx);
      }
    }
    // V pages of the two 8-key halves (block sizes 8 / 16 / 32)
    const T* vbase[2];
    int voff[2];
    bool v0ok[2], v1ok[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int tok = t0 + 8 * h + 2 * q4;
      v0ok[h] = tok < ctx;
      v1ok[h] = tok + 1 < ctx;
      vbase[h] = VC + (v0ok[h] ? (int64_t)bt[tok / BS] * a.vc_bs + (int64_t)kvh * a.vc_hs : 0);
      voff[h] = tok % BS;
    }
    softmax_pv(s, [&](int nd, int h) -> uint32_t {
      const int d = 8 * nd + g;
      if (!v0ok[h] || d >= D) return 0u;
      uint32_t w = ld32(vbase[h] + (int64_t)d * BS + voff[h]);
      return v1ok[h] ? w : (w & 0xffffu);      // never let cache garbage beyond the context meet a zero probability
    });
  }

  // ---- the call's own tokens: keys 0 .. min(qlen, last row of this warp + 1) - 1, causal ---------------------------------
  const int n_end = min(qlen, m0 + 16 * warp + 16);
  for (int t0 = 0; t0 < n_end; t0 += 16) {
    float s[2][4];
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
      s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
      const int tok = t0 + 8 * nt + g;
      if (t0 + 8 * nt < n_end) {
        const bool ok = tok < qlen;
        const T* kb = Kn + (ok ? (int64_t)(start + tok) * a.k_s0 + (int64_t)kvh * a.k_s1 : 0);
#pragma unroll
        for (int kk = 0; kk < KSTEPS; ++kk) {
          const uint32_t b0 = (ok && 16 * kk + 2 * q4 < D) ? ld32(kb + 16 * kk + 2 * q4) : 0u;
          const uint32_t b1 = (ok && 16 * kk + 8 + 2 * q4 < D) ? ld32(kb + 16 * kk + 8 + 2 * q4) : 0u;
          MmaT<T>::mma(s[nt], qa[kk], b0, b1);
        }
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int n = t0 + 8 * nt + 2 * q4 + (i & 1), r = row0 + 8 * (i >> 1);
        s[nt][i] = finish(s[nt][i], r, ctx + n, n < qlen && n <= r);
      }
    }
    const T* vb0[2];
    bool v0ok[2], v1ok[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int tok = t0 + 8 * h + 2 * q4;
      v0ok[h] = tok < qlen;
      v1ok[h] = tok + 1 < qlen;
      vb0[h] = Vn + (v0ok[h] ? (int64_t)(start + tok) * a.v_s0 + (int64_t)kvh * a.v_s1 : 0);
    }
    softmax_pv(s, [&](int nd, int h) -> uint32_t {
      const int d = 8 * nd + g;
      if (!v0ok[h] || d >= D) return 0u;
      uint32_t w = ld16(vb0[h] + d);
      if (v1ok[h]) w |= ld16(vb0[h] + a.v_s0 + d) << 16;
      return w;
    });
  }

  // ---- normalise and store -----------------------------------------------------------------------------------------------
  T* O = reinterpret_cast<T*>(a.o);
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int r = row0 + 8 * h;
    if (r >= qlen) continue;
    const float inv = l_run[h] > 0.f ? 1.f / l_run[h] : 0.f;
    T* orow = O + (int64_t)(start + r) * a.o_s0 + (int64_t)head * a.o_s1;
#pragma unroll
    for (int nd = 0; nd < NDT; ++nd) {
      const int d = 8 * nd + 2 * q4;
      if (d < D) *reinterpret_cast<uint32_t*>(orow + d) = MmaT<T>::pack(acc[nd][2 * h] * inv, acc[nd][2 * h + 1] * inv);
    }
  }
}


// ---- tiled variant: K / V tiles of 64 keys staged in shared memory once per CTA ------------------------------------------
// The direct-from-global kernel above re-loads every K / V fragment in each of the 4 warps (and the dense new-token V needs
// two 16-bit loads per register): 17 TFLOP/s on a 2k Llama-3-8B prompt.  Here the CTA copies a 64-key tile with 16-byte
// cp.async (zero-filled beyond the context / query / head size), XOR-swizzled by row so that ldmatrix is conflict free:
//   K  [64 keys][DPAD]   from the dense k rows or from the cache chunks (a chunk row IS an ldmatrix row)
//   V  [64 keys][DPAD]   dense new tokens, read with ldmatrix.trans;   V^T [DPAD][64 keys] cache rows, read without
// and every warp pulls its B fragments with ldmatrix.x4 (16 per 16 keys instead of 64-190 global loads).
template <int DPAD>
struct PrefillTile {
  static constexpr int BN = 64;
  static constexpr int CPR = DPAD / 8;                     // 16-byte chunks of a key row that carry data
  static constexpr int ROWC = (CPR + 7) / 8 * 8;           // row pitch in chunks (multiple of 8 for the swizzle)
  static constexpr int K_BYTES = BN * ROWC * 16;
  static constexpr int VT_BYTES = DPAD * 128;              // V^T: DPAD rows of 64 keys
  static constexpr int V_BYTES = K_BYTES > VT_BYTES ? K_BYTES : VT_BYTES;
};

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, int src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
  asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
}
__device__ __forceinline__ void ldsm4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldsm4_t(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}

template <typename T, int DPAD>
__global__ void __launch_bounds__(128, 3) prefill_paged_attn_tiled_kernel(const PrefillArgs a) {   // 3 CTAs / SM: <= 168 registers, 3 x 64 KB
  using TL = PrefillTile<DPAD>;
  extern __shared__ __align__(128) uint8_t smem[];       // two buffers of (K tile, V tile)
  constexpr uint32_t kBufBytes = TL::K_BYTES + TL::V_BYTES;
  pdl_wait();
  pdl_launch_dependents();
  const int seq = blockIdx.z, head = blockIdx.y;
  const int ctx = a.ctx_lens[seq];
  const int qlen = a.seq_lens[seq] - ctx;
  // heaviest query blocks first: block x attends to ctx + 64 (x + 1) keys, so the last blocks of a long prompt run ~32x longer
  // than the first; launched in index order they would make up the tail of the grid (2.3 waves of 3 CTAs / SM on a 2k prompt)
  const int m0 = (int)(gridDim.x - 1 - blockIdx.x) * kPrefillRows;
  if (m0 >= qlen) return;                                // CTA-uniform
  const int start = a.start_loc[seq];
  const int kvh = head / (a.num_heads / a.num_kv_heads);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, q4 = lane & 3;
  const int D = a.head_size, BS = a.block_size, SW = a.sliding_window;
  const int row0 = m0 + 16 * warp + g;
  const bool warp_active = m0 + 16 * warp < qlen;        // idle warps still copy tiles and meet the barriers
  constexpr int KSTEPS = DPAD / 16, NDT = DPAD / 8, CPR = TL::CPR, ROWC = TL::ROWC;
  const T* Q = reinterpret_cast<const T*>(a.q);
  const T* Kn = reinterpret_cast<const T*>(a.k);
  const T* Vn = reinterpret_cast<const T*>(a.v);
  const T* KC = reinterpret_cast<const T*>(a.kc);
  const T* VC = reinterpret_cast<const T*>(a.vc);
  const int* bt = a.block_tables + (int64_t)seq * a.bt_stride;
  const float slope = a.alibi ? a.alibi[head] : 0.f;
  const uint32_t Ks = (uint32_t)__cvta_generic_to_shared(smem), Vs = Ks + TL::K_BYTES;   // buffer 0; buffer 1 at + kBufBytes

  uint32_t qa[KSTEPS][4];
#pragma unroll
  for (int kk = 0; kk < KSTEPS; ++kk) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int r = row0 + 8 * (i & 1), d = 16 * kk + 2 * q4 + 8 * (i >> 1);
      qa[kk][i] = (r < qlen && d < D) ? ld32(Q + (int64_t)(start + r) * a.q_s0 + (int64_t)head * a.q_s1 + d) : 0u;
    }
  }
  float acc[NDT][4];
#pragma unroll
  for (int i = 0; i < NDT; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.f;
  float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};

  auto kv_off = [&](int t, int c) -> uint32_t { return (uint32_t)(t * ROWC + ((c & ~7) | ((c ^ t) & 7))) * 16u; };
  auto vt_off = [&](int d, int j) -> uint32_t { return (uint32_t)(d * 8 + ((j ^ d) & 7)) * 16u; };

  // S = Q.K^T over the staged 64 keys, masks, online softmax, O += P.V.  NEW: keys are the call's own tokens (causal, V row-major);
  // otherwise cached context (V^T)
  auto compute = [&](const int t0, const bool is_new, const uint32_t Kb, const uint32_t Vb) {
    float s[8][4];
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
    const int mid = lane >> 3, rr = lane & 7;
#pragma unroll
    for (int kk = 0; kk < KSTEPS; ++kk) {
#pragma unroll
      for (int np = 0; np < 4; ++np) {
        uint32_t r[4];
        ldsm4(r, Kb + kv_off(16 * np + 8 * (mid >> 1) + rr, 2 * kk + (mid & 1)));
        MmaT<T>::mma(s[2 * np], qa[kk], r[0], r[1]);
        MmaT<T>::mma(s[2 * np + 1], qa[kk], r[2], r[3]);
      }
    }
    float mx[2] = {-INFINITY, -INFINITY};
    // most tiles need no mask at all: every key valid and (own tokens) strictly below this warp's diagonal, no window, no ALiBi
    const bool plain = SW == 0 && a.alibi == nullptr &&
                       (is_new ? (t0 + TL::BN <= qlen && t0 + TL::BN - 1 <= m0 + 16 * warp) : (t0 + TL::BN <= ctx));
    if (plain) {
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          s[nt][i] *= a.scale;
          mx[i >> 1] = fmaxf(mx[i >> 1], s[nt][i]);
        }
      }
    } else
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int kidx = t0 + 8 * nt + 2 * q4 + (i & 1), r = row0 + 8 * (i >> 1);
        const bool valid = is_new ? (kidx < qlen && kidx <= r) : (kidx < ctx);
        float x = -INFINITY;
        if (valid) {
          const int kpos = is_new ? ctx + kidx : kidx, qpos = ctx + r;
          x = s[nt][i] * a.scale;
          if (a.alibi) x += slope * (float)(kpos - qpos);
          if (SW > 0 && qpos - kpos >= SW) x = -10000.f;
        }
        s[nt][i] = x;
        mx[i >> 1] = fmaxf(mx[i >> 1], x);
      }
    }
    float alpha[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      float v = mx[h];
      v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 1));
      v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 2));
      const float mnew = fmaxf(m_run[h], v);
      const float msafe = mnew == -INFINITY ? 0.f : mnew;
      alpha[h] = __expf(m_run[h] - msafe);
      float rs = 0.f;
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
        s[nt][2 * h] = __expf(s[nt][2 * h] - msafe);
        s[nt][2 * h + 1] = __expf(s[nt][2 * h + 1] - msafe);
        rs += s[nt][2 * h] + s[nt][2 * h + 1];
      }
      rs += __shfl_xor_sync(0xffffffffu, rs, 1);
      rs += __shfl_xor_sync(0xffffffffu, rs, 2);
      l_run[h] = l_run[h] * alpha[h] + rs;
      m_run[h] = mnew;
    }
#pragma unroll
    for (int nd = 0; nd < NDT; ++nd) {
      acc[nd][0] *= alpha[0]; acc[nd][1] *= alpha[0];
      acc[nd][2] *= alpha[1]; acc[nd][3] *= alpha[1];
    }
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {
      uint32_t pa[4];
      pa[0] = MmaT<T>::pack(s[2 * ks][0], s[2 * ks][1]);
      pa[1] = MmaT<T>::pack(s[2 * ks][2], s[2 * ks][3]);
      pa[2] = MmaT<T>::pack(s[2 * ks + 1][0], s[2 * ks + 1][1]);
      pa[3] = MmaT<T>::pack(s[2 * ks + 1][2], s[2 * ks + 1][3]);
#pragma unroll
      for (int dp = 0; dp < NDT / 2; ++dp) {
        uint32_t r[4];
        if (is_new) ldsm4_t(r, Vb + kv_off(16 * ks + 8 * (mid & 1) + rr, 2 * dp + (mid >> 1)));
        else ldsm4(r, Vb + vt_off(8 * (2 * dp + (mid >> 1)) + rr, 2 * ks + (mid & 1)));
        MmaT<T>::mma(acc[2 * dp], pa, r[0], r[1]);
        MmaT<T>::mma(acc[2 * dp + 1], pa, r[2], r[3]);
      }
    }
  };

  // ---- one list of 64-key tiles: the cached context first, then the call's own tokens; two buffers, the copy of tile
  // i + 1 is in flight while tile i is computed ------------------------------------------------------------------------------
  const int n_end = min(qlen, m0 + kPrefillRows);
  const int n_ctx_tiles = (ctx + TL::BN - 1) / TL::BN, n_tiles = n_ctx_tiles + (n_end + TL::BN - 1) / TL::BN;
  auto load_tile = [&](const int ti) {
    const uint32_t kb = Ks + (uint32_t)(ti & 1) * kBufBytes, vb = Vs + (uint32_t)(ti & 1) * kBufBytes;
    if (ti < n_ctx_tiles) {
      const int t0 = ti * TL::BN;
      for (int id = tid; id < TL::BN * CPR; id += 128) {
        const int t = id / CPR, c = id - t * CPR, tok = t0 + t;
        const bool ok = tok < ctx && 8 * c < D;
        const T* src = ok ? KC + (int64_t)bt[tok / BS] * a.kc_bs + (int64_t)kvh * a.kc_hs + ((int64_t)c * BS + tok % BS) * 8 : KC;
        cp_async16(kb + kv_off(t, c), src, ok ? 16 : 0);
      }
      for (int id = tid; id < DPAD * 8; id += 128) {
        const int d = id >> 3, j = id & 7, tok0 = t0 + 8 * j;
        int nv = ctx - tok0;                               // valid keys of this 8-key chunk; the rest is zero-filled so that
        nv = (d < D) ? (nv > 8 ? 8 : (nv < 0 ? 0 : nv)) : 0;   // stale cache contents never meet a zero probability
        const T* src = nv > 0 ? VC + (int64_t)bt[tok0 / BS] * a.vc_bs + (int64_t)kvh * a.vc_hs + (int64_t)d * BS + tok0 % BS : VC;
        cp_async16(vb + vt_off(d, j), src, nv * 2);
      }
    } else {
      const int t0 = (ti - n_ctx_tiles) * TL::BN;
      for (int id = tid; id < TL::BN * CPR; id += 128) {
        const int t = id / CPR, c = id - t * CPR, tok = t0 + t;
        const bool ok = tok < qlen && 8 * c < D;
        const int64_t off = ok ? (int64_t)(start + tok) * a.k_s0 + (int64_t)kvh * a.k_s1 + 8 * c : 0;
        const int64_t voff = ok ? (int64_t)(start + tok) * a.v_s0 + (int64_t)kvh * a.v_s1 + 8 * c : 0;
        cp_async16(kb + kv_off(t, c), Kn + off, ok ? 16 : 0);
        cp_async16(vb + kv_off(t, c), Vn + voff, ok ? 16 : 0);
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  if (n_tiles > 0) load_tile(0);
  for (int ti = 0; ti < n_tiles; ++ti) {
    if (ti + 1 < n_tiles) {
      load_tile(ti + 1);                                   // its buffer was released by the barrier that ended tile ti - 1
      asm volatile("cp.async.wait_group 1;" ::: "memory");
    } else {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncthreads();                                       // tile ti has landed for every thread
    const uint32_t kb = Ks + (uint32_t)(ti & 1) * kBufBytes, vb = Vs + (uint32_t)(ti & 1) * kBufBytes;
    if (ti < n_ctx_tiles) {
      if (warp_active) compute(ti * TL::BN, false, kb, vb);
    } else {
      const int t0 = (ti - n_ctx_tiles) * TL::BN;
      if (warp_active && t0 <= m0 + 16 * warp + 15) compute(t0, true, kb, vb);   // tiles entirely above the warp's rows are all masked
    }
    __syncthreads();                                       // everyone is done with buffer ti & 1 before tile ti + 2 overwrites it
  }

  if (!warp_active) return;
  T* O = reinterpret_cast<T*>(a.o);
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int r = row0 + 8 * h;
    if (r >= qlen) continue;
    const float inv = l_run[h] > 0.f ? 1.f / l_run[h] : 0.f;
    T* orow = O + (int64_t)(start + r) * a.o_s0 + (int64_t)head * a.o_s1;
#pragma unroll
    for (int nd = 0; nd < NDT; ++nd) {
      const int d = 8 * nd + 2 * q4;
      if (d < D) *reinterpret_cast<uint32_t*>(orow + d) = MmaT<T>::pack(acc[nd][2 * h] * inv, acc[nd][2 * h + 1] * inv);
    }
  }
}

template <typename T>
static int launch_prefill(const PrefillArgs& a, int batch, int max_input_len, cudaStream_t st) {
  dim3 grid((unsigned)((max_input_len + kPrefillRows - 1) / kPrefillRows), (unsigned)a.num_heads, (unsigned)batch);
  const int D = a.head_size;
  cudaError_t e;
  // the tiled kernel copies 16-byte chunks: k / v / cache strides and bases must keep them aligned (always true for the
  // tensors vLLM passes); anything else takes the direct-from-global kernel
  static const bool no_tiled = getenv("B200_PREFILL_DIRECT") != nullptr;
  const bool aligned = ((a.k_s0 | a.k_s1 | a.v_s0 | a.v_s1 | a.kc_bs | a.kc_hs | a.vc_bs | a.vc_hs) % 8 == 0) &&
                       (((uintptr_t)a.k | (uintptr_t)a.v | (uintptr_t)a.kc | (uintptr_t)a.vc) % 16 == 0);
  if (aligned && !no_tiled) {
#define PREFILL_GO(DP)                                                                                                   \
  do {                                                                                                                   \
    constexpr int smem = 2 * (PrefillTile<DP>::K_BYTES + PrefillTile<DP>::V_BYTES);                                      \
    static bool attr_done = false;   /* > 48 KB of dynamic shared memory needs the opt-in, once per instantiation */     \
    if (!attr_done) {                                                                                                    \
      cudaFuncSetAttribute(prefill_paged_attn_tiled_kernel<T, DP>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);   \
      attr_done = true;                                                                                                  \
    }                                                                                                                    \
    e = launch_pdl(prefill_paged_attn_tiled_kernel<T, DP>, grid, dim3(128), smem, st, a);                                \
  } while (0)
    if (D <= 32) PREFILL_GO(32);
    else if (D <= 64) PREFILL_GO(64);
    else if (D <= 96) PREFILL_GO(96);
    else PREFILL_GO(128);
#undef PREFILL_GO
    if (e != cudaSuccess) return set_error(B200_ECUDA, "context_attention_fwd launch: %s", cudaGetErrorString(e));
    return B200_OK;
  }
  if (D <= 32) e = launch_pdl(prefill_paged_attn_kernel<T, 32>, grid, dim3(128), 0, st, a);
  else if (D <= 64) e = launch_pdl(prefill_paged_attn_kernel<T, 64>, grid, dim3(128), 0, st, a);
  else if (D <= 96) e = launch_pdl(prefill_paged_attn_kernel<T, 96>, grid, dim3(128), 0, st, a);
  else e = launch_pdl(prefill_paged_attn_kernel<T, 128>, grid, dim3(128), 0, st, a);
  if (e != cudaSuccess) return set_error(B200_ECUDA, "context_attention_fwd launch: %s", cudaGetErrorString(e));
  return B200_OK;
}

}  // namespace b200

extern "C" int b200_context_attention_fwd(void* o, const void* q, const void* k, const void* v, const void* k_cache,
                                          const void* v_cache, const int32_t* block_tables, int64_t bt_stride,
                                          const int32_t* b_start_loc, const int32_t* b_seq_len, const int32_t* b_ctx_len,
                                          int batch, int max_input_len, int num_heads, int num_kv_heads, int head_size,
                                          int block_size, int x, const int64_t* qkvo_strides, int64_t kc_block_stride,
                                          int64_t kc_head_stride, int64_t vc_block_stride, int64_t vc_head_stride,
                                          const float* alibi_slopes, int sliding_window, int dtype, void* stream) {
  using namespace b200;
  B200_CHECK(dtype == B200_F16 || dtype == B200_BF16, "context_attention_fwd: only float16 / bfloat16 (the cache holds the query dtype)");
  B200_CHECK(head_size % 8 == 0 && head_size >= 8, "context_attention_fwd: head size %d is not a multiple of 8", head_size);
  if (head_size > 128) return set_error(B200_ENOTSUP, "context_attention_fwd: head size %d > 128 is not supported", head_size);
  B200_CHECK(x == 8, "context_attention_fwd: key cache must be [num_blocks, kv_heads, head_size/8, block_size, 8], got x = %d", x);
  B200_CHECK(block_size == 8 || block_size == 16 || block_size == 32, "Unsupported block size: %d", block_size);
  B200_CHECK(num_kv_heads > 0 && num_heads % num_kv_heads == 0, "context_attention_fwd: num_heads %d not a multiple of num_kv_heads %d",
             num_heads, num_kv_heads);
  for (int i = 0; i < 8; ++i)
    B200_CHECK(qkvo_strides[i] % 2 == 0, "context_attention_fwd: q / k / v / o strides must be even (4-byte vector accesses)");
  B200_CHECK(((uintptr_t)o | (uintptr_t)q | (uintptr_t)k | (uintptr_t)v | (uintptr_t)k_cache | (uintptr_t)v_cache) % 4 == 0,
             "context_attention_fwd: tensors must be 4-byte aligned");
  if (batch == 0 || max_input_len <= 0) return B200_OK;
  PrefillArgs a;
  a.o = o; a.q = q; a.k = k; a.v = v; a.kc = k_cache; a.vc = v_cache;
  a.block_tables = block_tables; a.start_loc = b_start_loc; a.seq_lens = b_seq_len; a.ctx_lens = b_ctx_len;
  a.alibi = alibi_slopes;
  a.bt_stride = bt_stride;
  a.q_s0 = qkvo_strides[0]; a.q_s1 = qkvo_strides[1]; a.k_s0 = qkvo_strides[2]; a.k_s1 = qkvo_strides[3];
  a.v_s0 = qkvo_strides[4]; a.v_s1 = qkvo_strides[5]; a.o_s0 = qkvo_strides[6]; a.o_s1 = qkvo_strides[7];
  a.kc_bs = kc_block_stride; a.kc_hs = kc_head_stride; a.vc_bs = vc_block_stride; a.vc_hs = vc_head_stride;
  a.num_heads = num_heads; a.num_kv_heads = num_kv_heads; a.head_size = head_size; a.block_size = block_size;
  a.sliding_window = sliding_window > 0 ? sliding_window : 0;
  a.scale = 1.0f / sqrtf((float)head_size);
  cudaStream_t st = (cudaStream_t)stream;
  int rc = dtype == B200_F16 ? launch_prefill<__half>(a, batch, max_input_len, st) : launch_prefill<__nv_bfloat16>(a, batch, max_input_len, st);
  if (rc) return rc;
  B200_LAUNCH_CHECK("context_attention_fwd");
  return B200_OK;
}
