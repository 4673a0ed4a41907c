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

constexpr int kWarps = 4;
constexpr int kThreads = kWarps * 32;
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
__device__ __forceinline__ uint32_t ldg32(const void* p) {
  uint32_t v;
  asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
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

template <typename T, int KV>
__global__ void __launch_bounds__(amma::kThreads, 3)
paged_attn_gqa_kernel(float* __restrict__ exp_sums, float* __restrict__ max_logits, T* __restrict__ out,
                      const T* __restrict__ q, const void* __restrict__ k_cache_v, const void* __restrict__ v_cache_v,
                      int num_heads, int num_kv_heads, float scale, const int* __restrict__ block_tables,
                      const int* __restrict__ seq_lens, int max_blocks_per_seq, const float* __restrict__ alibi_slopes,
                      int64_t q_stride, int64_t kv_block_stride, int64_t kv_head_stride, int partitioned, float kv_scale) {
  using namespace amma;
  constexpr bool FP8 = KV != B200_KV_AUTO;
  using MT = typename std::conditional<FP8, __half, T>::type;   // MMA operand type
  const float s_scale = FP8 ? scale * kv_scale : scale;
  const float o_scale = FP8 ? kv_scale : 1.f;
  constexpr int OSTR = D + 4;   // padded row of the merge buffer (conflict-free fragment stores)
  __shared__ float sm_m[kWarps][GMAX], sm_l[kWarps][GMAX];
  __shared__ float sm_o[kWarps][GMAX][OSTR];
  pdl_wait();
  pdl_launch_dependents();
  const int kvh = blockIdx.x, seq = blockIdx.y, part = blockIdx.z, num_parts = gridDim.z;
  const int G = num_heads / num_kv_heads;
  const int seq_len = seq_lens[seq];
  const int start_token = partitioned ? part * kPartition : 0;
  if (partitioned && start_token >= seq_len) return;   // attention_kernels.cu:116-119
  const int end_token = partitioned ? min(start_token + kPartition, seq_len) : seq_len;
  const int start_block = start_token / BS, end_block = (end_token + BS - 1) / BS;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
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

  const int* btab = block_tables + (int64_t)seq * max_blocks_per_seq;
  for (int blk = start_block + warp; blk < end_block; blk += kWarps) {
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
        kf[c][0] = ldg32(kp + (2 * c) * (BS * 8));
        kf[c][1] = ldg32(kp + (2 * c) * (BS * 8) + 64);
        kf[c][2] = ldg32(kp + (2 * c + 1) * (BS * 8));
        kf[c][3] = ldg32(kp + (2 * c + 1) * (BS * 8) + 64);
      }
#pragma unroll
      for (int dm = 0; dm < 8; ++dm) {
        vf[dm][0] = ldg32(vp + (16 * dm) * BS);
        vf[dm][1] = ldg32(vp + (16 * dm + 8) * BS);
        vf[dm][2] = ldg32(vp + (16 * dm) * BS + 8);
        vf[dm][3] = ldg32(vp + (16 * dm + 8) * BS + 8);
      }
    } else {
      const uint8_t* kp = reinterpret_cast<const uint8_t*>(k_cache_v) + base + 4 * qd;        // K [D/16][16][16]
      const uint8_t* vp = reinterpret_cast<const uint8_t*>(v_cache_v) + base + g * BS + 4 * qd;  // V [D][16]
      uint32_t kr[8][2], vr[8][2];
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        kr[c][0] = ldg32(kp + c * (BS * 16) + t_lo * 16);
        kr[c][1] = ldg32(kp + c * (BS * 16) + t_hi * 16);
      }
#pragma unroll
      for (int dm = 0; dm < 8; ++dm) {
        vr[dm][0] = ldg32(vp + (16 * dm) * BS);
        vr[dm][1] = ldg32(vp + (16 * dm + 8) * BS);
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
  const int d = threadIdx.x;   // kThreads == D
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
int launch_mma(float* exp_sums, float* max_logits, void* out, void* tmp_out, const void* q, const void* kc,
               const void* vc, int num_seqs, int num_heads, int num_kv_heads, int head_size, float scale,
               const int* block_tables, const int* seq_lens, int max_blocks_per_seq, int block_size, int max_seq_len,
               const float* alibi, int64_t q_stride, int64_t kvbs, int64_t kvhs, int partitioned, int num_parts,
               int dtype, int kv_dtype, float kv_scale, cudaStream_t st) {
  using namespace amma;
  (void)max_seq_len;
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
#define GQA_GO(TT, KVV)                                                                                                 \
  e = launch_pdl(paged_attn_gqa_kernel<TT, KVV>, grid, dim3(kThreads), 0, st, exp_sums, max_logits, (TT*)dst,            \
                 (const TT*)q, kc, vc, num_heads, num_kv_heads, scale, block_tables, seq_lens, max_blocks_per_seq, alibi, \
                 q_stride, kvbs, kvhs, partitioned, kv_scale)
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
  if (e != cudaSuccess) return set_error(B200_ECUDA, "paged_attention (gqa path) launch: %s", cudaGetErrorString(e));
  return 1;
}

}  // namespace b200
