// GQA-aware paged-attention decode for sm_100a (fast path of paged_attention_v1 / v2).
//
// Reference semantics: csrc/attention/attention_kernels.cu:90-496 (one CTA per QUERY head, so a kv head is re-read
// H/Hkv times) and :567-669 (v2 merge).  Here one CTA serves all G = H/Hkv query heads of one (sequence, kv head,
// partition): every KV byte is read from HBM once.  Covered: head_size 128, block_size 16, fp16/bf16 queries, kv cache
// "auto", G <= 8, optional alibi.  Everything else returns 0 ("not covered") and the generic kernel runs.
//
// Each warp walks the pages p = w, w + W, ... of its token range with an online softmax (running max / sum per head):
//   S^T[16 tokens x 8 heads] = K_page[16 x 128] . Q^T[128 x 8]   8 x mma.m16n8k16 (A = K fragments, B = Q fragments)
//   O^T[128 d x 8 heads]    += V_page^T[128 x 16] . P^T[16 x 8]  8 x mma.m16n8k16 (A = V fragments, B = P^T)
// The vLLM cache layouts are fragment friendly: K [D/8][16 tok][8] puts the 4-byte (d, d+1) pair lane (g,q) needs for
// token g at offset tok*16 + 4q of an 8-d chunk -> every K load instruction is one fully used 128-byte line; V [D][16 tok]
// puts lane (g,q)'s token pair (2q,2q+1) of row d = g at d*32 + 4q -> two loads cover 256 contiguous bytes.  So K and V
// go from global memory straight into tensor-core registers: no shared-memory staging, no conversion, and all 64
// loads of a page are issued before the first MMA (8 KB in flight per warp).  P^T (the B operand of the second MMA) is the
// transpose of the accumulator layout of the first: two movmatrix instructions.  p is rounded to the query dtype before
// P.V and the denominators carry +1e-6 like the reference (:342, :398-400, :652).
// Warps are merged through shared memory (v1: final output; v2: per-partition tmp_out / exp_sums / max_logits in the
// caller's buffers, merged by the same reduce kernel as the generic path).
#include "common.cuh"
#include <stdlib.h>
#include <type_traits>


namespace b200 {
namespace amma {

// warps per CTA is a template parameter (2, 3 or 4): at ~148 registers per thread an SM holds 3 CTAs of 4 warps, 4 of 3 or 6 of
// 2, so the launcher picks the widest CTA whose grid still fits in ONE wave (64 sequences x 8 kv heads = 512 CTAs do not fit
// the 444 slots of 4-warp CTAs: the 68 left-over CTAs cost a second, latency-bound wave)
constexpr int D = 128, BS = 16, GMAX = 8;
constexpr int kPartition = 512;

template <typename T> struct Mma;
template <> struct Mma<__half> {
  static __device__ __forceinline__ void run(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
  static __device__ __forceinline__ uint32_t pack(float x, float y) {
    __half2 h = __floats2half2_rn(x, y);
    return *reinterpret_cast<uint32_t*>(&h);
  }
};
template <> struct Mma<__nv_bfloat16> {
  static __device__ __forceinline__ void run(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
  static __device__ __forceinline__ uint32_t pack(float x, float y) {
    __nv_bfloat162 h = __floats2bfloat162_rn(x, y);
    return *reinterpret_cast<uint32_t*>(&h);
  }
};

__device__ __forceinline__ uint32_t movmatrix_trans(uint32_t x) {
  uint32_t y;
  asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(y) : "r"(x));
  return y;
}
// NC: read-only path (the caches are not written by this kernel).  The fused kernel below writes the new token's K / V
// itself and must read them back coherently.
template <bool NC>
__device__ __forceinline__ uint32_t ldg32(const void* p) {
  uint32_t v;
  if (NC) asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p));
  else asm volatile("ld.global.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// rotary embedding (neox) + cache write of the step's new token, folded into the attention kernel's prologue
// (reference: csrc/pos_encoding_kernels.cu:10-93 and csrc/cache_kernels.cu:153-204; rounding points as in glue.cu:
// T(T(x*cos) - T(y*sin))).  All pointers may be null when the kernel is not fused.
struct RopeCacheArgs {
  const int64_t* positions;     // [S]
  const void* cos_sin;          // [max_pos, 128]: cos [0,64) | sin [64,128)
  const void* key;              // [S, Hkv, 128], row stride key_stride
  const void* value;
  const int64_t* slot_mapping;  // [S], -1 = padding
  int64_t key_stride, value_stride;
  void* k_cache_w;
  void* v_cache_w;
};

template <typename T>
__device__ __forceinline__ T rope_x(T x, T y, T c, T s_) {   // x' = T(T(x c) - T(y s))
  return from_f32<T>(to_f32<T>(from_f32<T>(to_f32<T>(x) * to_f32<T>(c))) - to_f32<T>(from_f32<T>(to_f32<T>(y) * to_f32<T>(s_))));
}
template <typename T>
__device__ __forceinline__ T rope_y(T x, T y, T c, T s_) {   // y' = T(T(y c) + T(x s))
  return from_f32<T>(to_f32<T>(from_f32<T>(to_f32<T>(y) * to_f32<T>(c))) + to_f32<T>(from_f32<T>(to_f32<T>(x) * to_f32<T>(s_))));
}
// packed pairs: (x2, y2) = elements (d, d+1) and (d+64, d+65); (c2, s2) = cos / sin of (d, d+1)
template <typename T>
__device__ __forceinline__ void rope_pair(uint32_t& x2, uint32_t& y2, uint32_t c2, uint32_t s2) {
  const T* x = reinterpret_cast<const T*>(&x2);
  const T* y = reinterpret_cast<const T*>(&y2);
  const T* c = reinterpret_cast<const T*>(&c2);
  const T* s_ = reinterpret_cast<const T*>(&s2);
  T xo[2], yo[2];
#pragma unroll
  for (int i = 0; i < 2; ++i) { xo[i] = rope_x<T>(x[i], y[i], c[i], s_[i]); yo[i] = rope_y<T>(x[i], y[i], c[i], s_[i]); }
  x2 = *reinterpret_cast<const uint32_t*>(xo);
  y2 = *reinterpret_cast<const uint32_t*>(yo);
}

// four fp8 bytes (little endian v0..v3) -> {half2(v0,v1), half2(v2,v3)}: exact (every e4m3 / e5m2 value is an fp16 value).
// With an fp8 cache the MMAs run in fp16 whatever the query dtype: K and V need no further conversion, kv_scale is folded
// into the softmax scale (K) and the final normalisation (V), Q and P are converted once per page.  The reference rounds
// k * kv_scale to the query dtype before the dot product (quant_utils.cuh:294-360); the difference is one rounding.
template <int KV>
__device__ __forceinline__ void fp8x4_to_h2(uint32_t v, uint32_t& lo, uint32_t& hi) {
  const __half2_raw h01 = __nv_cvt_fp8x2_to_halfraw2((__nv_fp8x2_storage_t)(v & 0xffffu), Fp8Kind<KV>::v);
  const __half2_raw h23 = __nv_cvt_fp8x2_to_halfraw2((__nv_fp8x2_storage_t)(v >> 16), Fp8Kind<KV>::v);
  lo = *reinterpret_cast<const uint32_t*>(&h01);
  hi = *reinterpret_cast<const uint32_t*>(&h23);
}
// packed pair of T -> packed pair of MT
template <typename T, typename MT>
__device__ __forceinline__ uint32_t repack(uint32_t v) {
  if constexpr (std::is_same<T, MT>::value) return v;
  else {
    const T* t = reinterpret_cast<const T*>(&v);
    return Mma<MT>::pack(to_f32<T>(t[0]), to_f32<T>(t[1]));
  }
}

}  // namespace amma

template <typename T, int KV, bool FUSED, int kWarps>
__global__ void __launch_bounds__(kWarps * 32, kWarps == 4 ? 3 : (kWarps == 3 ? 4 : 6))
paged_attn_gqa_kernel(float* __restrict__ exp_sums, float* __restrict__ max_logits, T* __restrict__ out,
                      const T* __restrict__ q, const void* __restrict__ k_cache_v, const void* __restrict__ v_cache_v,
                      int num_heads, int num_kv_heads, float scale, const int* __restrict__ block_tables,
                      const int* __restrict__ seq_lens, int max_blocks_per_seq, const float* __restrict__ alibi_slopes,
                      int64_t q_stride, int64_t kv_block_stride, int64_t kv_head_stride, int partitioned, float kv_scale,
                      const amma::RopeCacheArgs rc) {
  using namespace amma;
  constexpr int kThreads = kWarps * 32;
  constexpr bool FP8 = KV != B200_KV_AUTO;
  constexpr bool NC = !FUSED;
  using MT = typename std::conditional<FP8, __half, T>::type;   // MMA operand type
  const float s_scale = FP8 ? scale * kv_scale : scale;
  const float o_scale = FP8 ? kv_scale : 1.f;
  constexpr int OSTR = D + 4;   // padded row of the merge buffer (conflict-free fragment stores)
  __shared__ float sm_m[kWarps][GMAX], sm_l[kWarps][GMAX];
  __shared__ float sm_o[kWarps][GMAX][OSTR];
  const int kvh = blockIdx.x, seq = blockIdx.y, part = blockIdx.z, num_parts = gridDim.z;
  const int G = num_heads / num_kv_heads;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int* btab = block_tables + (int64_t)seq * max_blocks_per_seq;
  // L2 prefetch of this warp's first pages BEFORE the programmatic-dependency wait.  Nothing here depends on the previous
  // kernel of the step: seq_lens / block_tables are step inputs, the pages hold tokens of earlier steps (the one slot written
  // this step is written through L2, which the prefetch cannot make stale).  A warp otherwise walks its pages as a chain of
  // dependent HBM round trips (all 64 loads of a page in flight, then the MMAs); at 16 pages per CTA (ctx 256) that chain,
  // not bandwidth, is the kernel time.  kPrefetch pages ahead are kept in flight further down the loop.
  constexpr int kPrefetch = 2;   // measured 1 .. 4: 1 is best stand-alone (85 % at 64 x 256), 2 best inside the decode step; 3-4 lose both ways
  constexpr int kPageLines = D * BS * (FP8 ? 1 : 2) / 128;   // 128-byte lines of one K (or V) page: 32 / 16
  const int seq_len_pf = seq_lens[seq];
  const int start_block_pf = partitioned ? (part * kPartition) / BS : 0;
  const int end_block_pf = ((partitioned ? min(part * kPartition + kPartition, seq_len_pf) : seq_len_pf) + BS - 1) / BS;
  auto prefetch_page = [&](int blk) {
    if (blk < end_block_pf && lane < kPageLines) {
      const int64_t pb = ((int64_t)btab[blk] * kv_block_stride + (int64_t)kvh * kv_head_stride) * (FP8 ? 1 : 2) + lane * 128;
      asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const char*>(k_cache_v) + pb));
      asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const char*>(v_cache_v) + pb));
    }
  };
#pragma unroll
  for (int i = 0; i < kPrefetch; ++i) prefetch_page(start_block_pf + warp + i * kWarps);
  pdl_wait();
  pdl_launch_dependents();
  const int seq_len = seq_len_pf;
  const int start_token = partitioned ? part * kPartition : 0;
  if (partitioned && start_token >= seq_len) return;   // attention_kernels.cu:116-119
  const int end_token = partitioned ? min(start_token + kPartition, seq_len) : seq_len;
  const int start_block = start_token / BS, end_block = (end_token + BS - 1) / BS;
  const int g = lane >> 2, qd = lane & 3;
  const int head0 = kvh * G;

  // Q^T fragments (B operand): b0 = Q[head g][16c + 2qd, +1], b1 = Q[head g][16c + 8 + 2qd, +1]; heads >= G are zero
  uint32_t qf[8][2];
  {
    // fp8 caches: one 4-byte load holds d = 16c + 4qd .. +3, so MMA k slots (2qd, 2qd+1) / (2qd+8, 2qd+9) carry
    // d = 16c + 4qd + {0,1} / {2,3}; Q uses the same assignment
    const T* qrow = q + (int64_t)seq * q_stride + (int64_t)(head0 + (g < G ? g : 0)) * D + (FP8 ? 4 : 2) * qd;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      qf[c][0] = g < G ? repack<T, MT>(*reinterpret_cast<const uint32_t*>(qrow + 16 * c)) : 0u;
      qf[c][1] = g < G ? repack<T, MT>(*reinterpret_cast<const uint32_t*>(qrow + 16 * c + (FP8 ? 2 : 8))) : 0u;
    }
  }
  if constexpr (FUSED) {
    const int64_t pos = rc.positions[seq];
    const T* cs = reinterpret_cast<const T*>(rc.cos_sin) + pos * D;
    // the new token's K (rotated) and V go into the cache: thread d of the CTA that owns the token's partition
    const int64_t slot = rc.slot_mapping[seq];
    const bool writer = slot >= 0 && (!partitioned || part == (seq_len - 1) / kPartition);
    if (writer)
    for (int d = threadIdx.x; d < D; d += kThreads) {
      const int dd = d & 63;
      const T* krow = reinterpret_cast<const T*>(rc.key) + (int64_t)seq * rc.key_stride + (int64_t)kvh * D;
      const T x = krow[dd], y = krow[dd + 64], c = cs[dd], sn = cs[64 + dd];
      const T kr = d < 64 ? rope_x<T>(x, y, c, sn) : rope_y<T>(x, y, c, sn);
      const T vv = reinterpret_cast<const T*>(rc.value)[(int64_t)seq * rc.value_stride + (int64_t)kvh * D + d];
      const int64_t wbase = (slot / BS) * kv_block_stride + (int64_t)kvh * kv_head_stride;
      const int off = (int)(slot % BS);
      if constexpr (!FP8) {
        reinterpret_cast<T*>(rc.k_cache_w)[wbase + (d >> 3) * (BS * 8) + off * 8 + (d & 7)] = kr;
        reinterpret_cast<T*>(rc.v_cache_w)[wbase + d * BS + off] = vv;
      } else {
        reinterpret_cast<uint8_t*>(rc.k_cache_w)[wbase + (d >> 4) * (BS * 16) + off * 16 + (d & 15)] =
            fp8_quant<FP8 ? KV : B200_KV_FP8_E4M3>(to_f32<T>(kr), kv_scale);
        reinterpret_cast<uint8_t*>(rc.v_cache_w)[wbase + d * BS + off] = fp8_quant<FP8 ? KV : B200_KV_FP8_E4M3>(to_f32<T>(vv), kv_scale);
      }
    }
    // the query heads are rotated in their fragments: element d (chunk c < 4) and its partner d + 64 (chunk c + 4) sit in
    // the same lane
    if (g < G) {
      const T* qrow = q + (int64_t)seq * q_stride + (int64_t)(head0 + g) * D + (FP8 ? 4 : 2) * qd;
#pragma unroll
      for (int c = 0; c < 4; ++c)
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          const int doff = 16 * c + (i ? (FP8 ? 2 : 8) : 0);
          const int d0 = doff + (FP8 ? 4 : 2) * qd;
          uint32_t x2 = *reinterpret_cast<const uint32_t*>(qrow + doff);
          uint32_t y2 = *reinterpret_cast<const uint32_t*>(qrow + doff + 64);
          rope_pair<T>(x2, y2, *reinterpret_cast<const uint32_t*>(cs + d0), *reinterpret_cast<const uint32_t*>(cs + 64 + d0));
          qf[c][i] = repack<T, MT>(x2);
          qf[c + 4][i] = repack<T, MT>(y2);
        }
    }
    __syncthreads();   // the token just written is read back below by the warp that walks its page
  }
  float slope[2] = {0.f, 0.f};
  if (alibi_slopes) {
    if (2 * qd < G) slope[0] = alibi_slopes[head0 + 2 * qd];
    if (2 * qd + 1 < G) slope[1] = alibi_slopes[head0 + 2 * qd + 1];
  }

  constexpr float NEG = -3.402823466e38f;
  float m_run[2] = {NEG, NEG}, l_run[2] = {0.f, 0.f};
  float o[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;

  for (int blk = start_block + warp; blk < end_block; blk += kWarps) {
    prefetch_page(blk + kPrefetch * kWarps);
    const int64_t base = (int64_t)btab[blk] * kv_block_stride + (int64_t)kvh * kv_head_stride;
    const int tok0 = blk * BS;
    // MMA row r of S^T <-> token: auto: r; fp8: rows 0-7 -> 2r - r%2, rows 8-15 -> that + 2 (so that the transposed P
    // fragment lines up with the 4-token groups one fp8 V load delivers)
    const int t_lo = FP8 ? 2 * g - (g & 1) : g;
    const int t_hi = FP8 ? t_lo + 2 : g + 8;
    uint32_t kf[8][4], vf[8][4];
    if constexpr (!FP8) {
      const T* kp = reinterpret_cast<const T*>(k_cache_v) + base + g * 8 + 2 * qd;          // K [D/8][16][8]
      const T* vp = reinterpret_cast<const T*>(v_cache_v) + base + g * BS + 2 * qd;          // V [D][16]
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        kf[c][0] = ldg32<NC>(kp + (2 * c) * (BS * 8));
        kf[c][1] = ldg32<NC>(kp + (2 * c) * (BS * 8) + 64);
        kf[c][2] = ldg32<NC>(kp + (2 * c + 1) * (BS * 8));
        kf[c][3] = ldg32<NC>(kp + (2 * c + 1) * (BS * 8) + 64);
      }
#pragma unroll
      for (int dm = 0; dm < 8; ++dm) {
        vf[dm][0] = ldg32<NC>(vp + (16 * dm) * BS);
        vf[dm][1] = ldg32<NC>(vp + (16 * dm + 8) * BS);
        vf[dm][2] = ldg32<NC>(vp + (16 * dm) * BS + 8);
        vf[dm][3] = ldg32<NC>(vp + (16 * dm + 8) * BS + 8);
      }
    } else {
      const uint8_t* kp = reinterpret_cast<const uint8_t*>(k_cache_v) + base + 4 * qd;        // K [D/16][16][16]
      const uint8_t* vp = reinterpret_cast<const uint8_t*>(v_cache_v) + base + g * BS + 4 * qd;  // V [D][16]
      uint32_t kr[8][2], vr[8][2];
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        kr[c][0] = ldg32<NC>(kp + c * (BS * 16) + t_lo * 16);
        kr[c][1] = ldg32<NC>(kp + c * (BS * 16) + t_hi * 16);
      }
#pragma unroll
      for (int dm = 0; dm < 8; ++dm) {
        vr[dm][0] = ldg32<NC>(vp + (16 * dm) * BS);
        vr[dm][1] = ldg32<NC>(vp + (16 * dm + 8) * BS);
      }
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        fp8x4_to_h2<FP8 ? KV : B200_KV_FP8_E4M3>(kr[c][0], kf[c][0], kf[c][2]);
        fp8x4_to_h2<FP8 ? KV : B200_KV_FP8_E4M3>(kr[c][1], kf[c][1], kf[c][3]);
      }
#pragma unroll
      for (int dm = 0; dm < 8; ++dm) {
        fp8x4_to_h2<FP8 ? KV : B200_KV_FP8_E4M3>(vr[dm][0], vf[dm][0], vf[dm][2]);
        fp8x4_to_h2<FP8 ? KV : B200_KV_FP8_E4M3>(vr[dm][1], vf[dm][1], vf[dm][3]);
      }
    }
    float s[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int c = 0; c < 8; ++c) Mma<MT>::run(s, kf[c], qf[c][0], qf[c][1]);

    const int tk[2] = {tok0 + t_lo, tok0 + t_hi};
    const bool ok[2] = {tk[0] < end_token, tk[1] < end_token};
    float pmax[2] = {NEG, NEG};
#pragma unroll
    for (int r = 0; r < 2; ++r)
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        float v = s[2 * r + e] * s_scale;
        if (alibi_slopes) v += slope[e] * (float)(tk[r] - seq_len + 1);
        s[2 * r + e] = v;
        if (ok[r]) pmax[e] = fmaxf(pmax[e], v);
      }
    float alpha[2];
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      float v = pmax[e];
      v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 4));
      v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 8));
      v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 16));
      const float m_new = fmaxf(m_run[e], v);
      alpha[e] = __expf(m_run[e] - m_new);   // first page: exp(-huge) = 0
      m_run[e] = m_new;
    }
    float p[4];
#pragma unroll
    for (int r = 0; r < 2; ++r)
#pragma unroll
      for (int e = 0; e < 2; ++e) p[2 * r + e] = ok[r] ? __expf(s[2 * r + e] - m_run[e]) : 0.f;
#pragma unroll
    for (int e = 0; e < 2; ++e) l_run[e] = l_run[e] * alpha[e] + p[e] + p[2 + e];
    // P (tokens x heads, rounded to the query dtype) -> P^T fragments
    const uint32_t pb0 = movmatrix_trans(repack<T, MT>(Mma<T>::pack(p[0], p[1])));
    const uint32_t pb1 = movmatrix_trans(repack<T, MT>(Mma<T>::pack(p[2], p[3])));
    if (tok0 + BS > end_token) {
      // partial page: slots beyond the sequence may hold anything (NaN x 0 = NaN): zero them (attention_kernels.cu:420-430)
      const int v_lo = tok0 + (FP8 ? 4 * qd : 2 * qd), v_hi = tok0 + (FP8 ? 4 * qd + 2 : 2 * qd + 8);   // tokens of k slots
      const uint32_t mlo = ((v_lo < end_token) ? 0x0000ffffu : 0u) | ((v_lo + 1 < end_token) ? 0xffff0000u : 0u);
      const uint32_t mhi = ((v_hi < end_token) ? 0x0000ffffu : 0u) | ((v_hi + 1 < end_token) ? 0xffff0000u : 0u);
#pragma unroll
      for (int dm = 0; dm < 8; ++dm) { vf[dm][0] &= mlo; vf[dm][1] &= mlo; vf[dm][2] &= mhi; vf[dm][3] &= mhi; }
    }
#pragma unroll
    for (int dm = 0; dm < 8; ++dm) {
      o[dm][0] *= alpha[0]; o[dm][1] *= alpha[1]; o[dm][2] *= alpha[0]; o[dm][3] *= alpha[1];
      Mma<MT>::run(o[dm], vf[dm], pb0, pb1);
    }
  }

  // ---- merge the warps ----------------------------------------------------------------------------------------------------
#pragma unroll
  for (int e = 0; e < 2; ++e) {
    float v = l_run[e];
    v += __shfl_xor_sync(0xffffffffu, v, 4);
    v += __shfl_xor_sync(0xffffffffu, v, 8);
    v += __shfl_xor_sync(0xffffffffu, v, 16);
    l_run[e] = v;
  }
  if (g == 0) {
    sm_m[warp][2 * qd] = m_run[0]; sm_m[warp][2 * qd + 1] = m_run[1];
    sm_l[warp][2 * qd] = l_run[0]; sm_l[warp][2 * qd + 1] = l_run[1];
  }
#pragma unroll
  for (int dm = 0; dm < 8; ++dm) {
    sm_o[warp][2 * qd][16 * dm + g] = o[dm][0];
    sm_o[warp][2 * qd + 1][16 * dm + g] = o[dm][1];
    sm_o[warp][2 * qd][16 * dm + g + 8] = o[dm][2];
    sm_o[warp][2 * qd + 1][16 * dm + g + 8] = o[dm][3];
  }
  __syncthreads();
  for (int d = threadIdx.x; d < D; d += kThreads)
  for (int h = 0; h < G; ++h) {
    float M = NEG;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) M = fmaxf(M, sm_m[w][h]);
    float L = 0.f, acc = 0.f;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) {
      const float f = __expf(sm_m[w][h] - M);
      L += sm_l[w][h] * f;
      acc += sm_o[w][h][d] * f;
    }
    const float res = acc * __fdividef(o_scale, L + 1e-6f);
    int64_t oi = (int64_t)seq * num_heads + head0 + h;
    if (partitioned) {
      oi = oi * num_parts + part;
      if (d == 0) { max_logits[oi] = M; exp_sums[oi] = L; }
    }
    out[oi * D + d] = from_f32<T>(res);
  }
}

// returns 1 if handled, 0 if the shape is not covered by this path (the caller runs the generic kernel), < 0 on error
static int launch_mma_impl(float* exp_sums, float* max_logits, void* out, void* tmp_out, const void* q, const void* kc,
                           const void* vc, int num_seqs, int num_heads, int num_kv_heads, int head_size, float scale,
                           const int* block_tables, const int* seq_lens, int max_blocks_per_seq, int block_size,
                           int max_seq_len, const float* alibi, int64_t q_stride, int64_t kvbs, int64_t kvhs, int partitioned, int num_parts,
                           int dtype, int kv_dtype, float kv_scale, const amma::RopeCacheArgs* rcp, cudaStream_t st) {
  using namespace amma;
  static const bool disabled = getenv("B200_ATTN_GENERIC") != nullptr;
  if (disabled) return 0;
  const int G = num_heads / num_kv_heads;
  if (head_size != D || block_size != BS || G > GMAX) return 0;
  if (dtype != B200_F16 && dtype != B200_BF16) return 0;
  if (kv_dtype != B200_KV_AUTO && kv_dtype != B200_KV_FP8_E4M3 && kv_dtype != B200_KV_FP8_E5M2) return 0;
  // fragment loads are 4-byte: the cache strides (elements) and the query rows must keep them aligned
  const int64_t align_el = kv_dtype == B200_KV_AUTO ? 2 : 4;
  if ((kvbs % align_el) || (kvhs % align_el) || (q_stride & 1) || (reinterpret_cast<uintptr_t>(q) & 3) ||
      (reinterpret_cast<uintptr_t>(kc) & 3) || (reinterpret_cast<uintptr_t>(vc) & 3))
    return 0;
  dim3 grid(num_kv_heads, num_seqs, partitioned ? num_parts : 1);
  void* dst = partitioned ? tmp_out : out;
  cudaError_t e = cudaSuccess;
  RopeCacheArgs rc = {};
  if (rcp) rc = *rcp;
  // widest CTA whose grid fits in one wave (see the note at the top); beyond 6 CTAs of 2 warps per SM it is multi-wave anyway
  const long long ctas = (long long)grid.x * grid.y * grid.z;
  static const int force_w = getenv("B200_ATTN_WARPS") ? atoi(getenv("B200_ATTN_WARPS")) : 0;
  // ... but only when a CTA has enough pages for bandwidth, not latency, to be the limit (>= 16 per warp of the narrow CTA):
  // at 16 pages per CTA (ctx 256) the step measured slower with 3 warps (3.03 vs 2.91 ms), at 128 pages (ctx 2048) the
  // kernel goes from 78 % to 91 % of the HBM rate (profiles/r2_attn_perf_warps.txt)
  const int pages_per_cta = partitioned ? kPartition / BS : (max_seq_len + BS - 1) / BS;
  int nw = 4;
  // (or when the launch as a whole is long: >= 16k pages in total, e.g. one 32k-token sequence cut into 512 partitions)
  if (ctas > (long long)num_sms() * 3 && (pages_per_cta >= 64 || ctas * pages_per_cta >= 16384))
    nw = ctas <= (long long)num_sms() * 4 ? 3 : (ctas <= (long long)num_sms() * 6 ? 2 : 4);
  if (force_w >= 2 && force_w <= 4) nw = force_w;
#define GQA_GO3(TT, KVV, FF, WW)                                                                                        \
  e = launch_pdl(paged_attn_gqa_kernel<TT, KVV, FF, WW>, grid, dim3(WW * 32), 0, st, exp_sums, max_logits, (TT*)dst,     \
                 (const TT*)q, kc, vc, num_heads, num_kv_heads, scale, block_tables, seq_lens, max_blocks_per_seq, alibi, \
                 q_stride, kvbs, kvhs, partitioned, kv_scale, rc)
#define GQA_GO2(TT, KVV, FF) do { if (nw == 4) GQA_GO3(TT, KVV, FF, 4); else if (nw == 3) GQA_GO3(TT, KVV, FF, 3); else GQA_GO3(TT, KVV, FF, 2); } while (0)
#define GQA_GO(TT, KVV) do { if (rcp) GQA_GO2(TT, KVV, true); else GQA_GO2(TT, KVV, false); } while (0)
  if (dtype == B200_F16) {
    if (kv_dtype == B200_KV_AUTO) GQA_GO(__half, B200_KV_AUTO);
    else if (kv_dtype == B200_KV_FP8_E4M3) GQA_GO(__half, B200_KV_FP8_E4M3);
    else GQA_GO(__half, B200_KV_FP8_E5M2);
  } else {
    if (kv_dtype == B200_KV_AUTO) GQA_GO(__nv_bfloat16, B200_KV_AUTO);
    else if (kv_dtype == B200_KV_FP8_E4M3) GQA_GO(__nv_bfloat16, B200_KV_FP8_E4M3);
    else GQA_GO(__nv_bfloat16, B200_KV_FP8_E5M2);
  }
#undef GQA_GO
#undef GQA_GO2
#undef GQA_GO3
  if (e != cudaSuccess) return set_error(B200_ECUDA, "paged_attention (gqa path) launch: %s", cudaGetErrorString(e));
  return 1;
}

int launch_mma(float* exp_sums, float* max_logits, void* out, void* tmp_out, const void* q, const void* kc,
               const void* vc, int num_seqs, int num_heads, int num_kv_heads, int head_size, float scale,
               const int* block_tables, const int* seq_lens, int max_blocks_per_seq, int block_size, int max_seq_len,
               const float* alibi, int64_t q_stride, int64_t kvbs, int64_t kvhs, int partitioned, int num_parts,
               int dtype, int kv_dtype, float kv_scale, cudaStream_t st) {
  return launch_mma_impl(exp_sums, max_logits, out, tmp_out, q, kc, vc, num_seqs, num_heads, num_kv_heads, head_size, scale,
                         block_tables, seq_lens, max_blocks_per_seq, block_size, max_seq_len, alibi, q_stride, kvbs, kvhs, partitioned,
                         num_parts, dtype, kv_dtype, kv_scale, nullptr, st);
}

int launch_v2_reduce(void* out, const float* exp_sums, const float* max_logits, const void* tmp_out,
                     const int* seq_lens, int num_seqs, int num_heads, int head_size, int max_parts, int dtype,
                     cudaStream_t st);

}  // namespace b200

// Fused decode-step attention: rotary embedding (neox, rot_dim == head_size == 128) of q and k, the new token's K / V
// written into the paged cache, and paged_attention_v1 / v2 over the cache -- ONE launch (two with the v2 merge) instead
// of rotary_embedding + reshape_and_cache + paged_attention.  q is rotated in registers (the caller's query tensor is left
// unrotated), the rotated k only exists in the cache.  Contract beyond the three separate ops: slot_mapping[s] is the
// slot of token seq_lens[s] - 1 of sequence s (decode), and no two sequences of the batch write into / read from the
// same physical block (what the reference's block manager guarantees for a block that is being appended to).
// Returns B200_ENOTSUP when the shape is outside the GQA fast path (run the three ops instead).
extern "C" int b200_rope_cache_paged_attention(void* out, float* exp_sums, float* max_logits, void* tmp_out,
                                               const void* query, const void* key, const void* value, void* key_cache,
                                               void* value_cache, const int64_t* positions, const void* cos_sin_cache,
                                               int rot_dim, const int64_t* slot_mapping, int num_seqs, int num_heads,
                                               int num_kv_heads, int head_size, float scale, const int32_t* block_tables,
                                               const int32_t* seq_lens, int max_blocks_per_seq, int block_size,
                                               int max_seq_len, int64_t q_stride, int64_t k_stride, int64_t v_stride,
                                               int64_t kv_block_stride, int64_t kv_head_stride, int dtype, int kv_dtype,
                                               float kv_scale, int use_v2, void* stream) {
  using namespace b200;
  if (num_seqs == 0) return B200_OK;
  if (rot_dim != head_size || head_size != amma::D || block_size != amma::BS)
    return set_error(B200_ENOTSUP, "rope_cache_paged_attention: only head_size = rot_dim = 128, block_size 16");
  if (use_v2 && (exp_sums == nullptr || max_logits == nullptr || tmp_out == nullptr))
    return set_error(B200_EINVAL, "rope_cache_paged_attention: v2 needs exp_sums / max_logits / tmp_out");
  amma::RopeCacheArgs rc;
  rc.positions = positions; rc.cos_sin = cos_sin_cache; rc.key = key; rc.value = value; rc.slot_mapping = slot_mapping;
  rc.key_stride = k_stride; rc.value_stride = v_stride; rc.k_cache_w = key_cache; rc.v_cache_w = value_cache;
  cudaStream_t st = (cudaStream_t)stream;
  const int num_parts = use_v2 ? (max_seq_len + 511) / 512 : 1;
  int r = launch_mma_impl(exp_sums, max_logits, out, tmp_out, query, key_cache, value_cache, num_seqs, num_heads,
                          num_kv_heads, head_size, scale, block_tables, seq_lens, max_blocks_per_seq, block_size, max_seq_len,
                          nullptr, q_stride, kv_block_stride, kv_head_stride, use_v2 ? 1 : 0, num_parts, dtype, kv_dtype, kv_scale,
                          &rc, st);
  if (r < 0) return r;
  if (r == 0) return set_error(B200_ENOTSUP, "rope_cache_paged_attention: shape outside the GQA fast path");
  if (use_v2) {
    r = launch_v2_reduce(out, exp_sums, max_logits, tmp_out, seq_lens, num_seqs, num_heads, head_size, num_parts, dtype, st);
    if (r < 0) return r;
  }
  B200_LAUNCH_CHECK("rope_cache_paged_attention");
  return B200_OK;
}
