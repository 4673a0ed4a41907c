// Generic paged-attention decode (any head size, fp16/bf16/fp32 queries, auto/fp8 KV, block size 8/16/32, ALiBi).
// This is the all-shapes path; the GQA/tensor-core fast path for 16-bit queries lives in attention_mma.cu and the
// launchers in this file pick between them.
//
// Reference semantics restated from csrc/attention/attention_kernels.cu:90-496 (v1/v2 body), :567-669 (v2 reduce):
//   logit_j = scale * <q, k_j> (+ slope * (j - len + 1));  m = max_j logit_j;  e = sum_j exp(logit_j - m)
//   p_j     = exp(logit_j - m) / (e + 1e-6), rounded to the query dtype before P.V (attention_kernels.cu:398-400)
//   out     = sum_j p_j v_j  (fp32 accumulation)
//   v2: the same per 512-token partition -> tmp_out/exp_sums/max_logits, then the log-sum-exp merge.
// Mapping (differs from the reference's thread-group scheme): one warp walks KV blocks; in the QK phase lane ->
// (token in block, slice of D) so each lane issues 16-byte loads that are contiguous across the warp in the
// [D/x, BS, x] key layout; in the PV phase lane -> row d of the [D, BS] value layout.
#include "common.cuh"

namespace b200 {

constexpr int kPartition = 512;  // fixed by the caller-allocated v2 buffers (paged_attn.py:10)
constexpr int kAttnThreads = 128;
constexpr int kAttnWarps = kAttnThreads / 32;

template <typename cache_t, int KV>
__device__ __forceinline__ float cache_to_f32(cache_t v, float kv_scale) {
  if constexpr (KV == B200_KV_AUTO) return to_f32<cache_t>(v);
  else return fp8_dequant<KV>((uint8_t)v, kv_scale);
}

// dequantised K/V values are rounded to the query dtype in the reference (scaled_convert<K_vec, ...>) before use
template <typename scalar_t, typename cache_t, int KV>
__device__ __forceinline__ float cache_val(cache_t v, float kv_scale) {
  if constexpr (KV == B200_KV_AUTO) return to_f32<cache_t>(v);
  else return to_f32<scalar_t>(from_f32<scalar_t>(fp8_dequant<KV>((uint8_t)v, kv_scale)));
}

__device__ __forceinline__ float block_reduce_max(float v, float* red) {
  v = warp_max(v);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float r = red[0];
#pragma unroll
  for (int i = 1; i < kAttnWarps; ++i) r = fmaxf(r, red[i]);
  __syncthreads();
  return r;
}
__device__ __forceinline__ float block_reduce_sum(float v, float* red) {
  v = warp_sum(v);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float r = 0.f;
#pragma unroll
  for (int i = 0; i < kAttnWarps; ++i) r += red[i];
  __syncthreads();
  return r;
}

template <typename scalar_t, typename cache_t, int KV>
__global__ void __launch_bounds__(kAttnThreads)
paged_attn_generic_kernel(float* __restrict__ exp_sums, float* __restrict__ max_logits, scalar_t* __restrict__ out,
                          const scalar_t* __restrict__ q, const cache_t* __restrict__ k_cache,
                          const cache_t* __restrict__ v_cache, int num_kv_heads, float scale,
                          const int* __restrict__ block_tables, const int* __restrict__ seq_lens,
                          int max_blocks_per_seq, const float* __restrict__ alibi_slopes, int64_t q_stride,
                          int64_t kv_block_stride, int64_t kv_head_stride, int head_size, int block_size,
                          int partition_size /* 0: whole sequence (v1) */, float kv_scale) {
  constexpr int X = 16 / sizeof(cache_t);
  extern __shared__ float smem[];
  pdl_wait();
  pdl_launch_dependents();
  const int head = blockIdx.x, seq = blockIdx.y, part = blockIdx.z;
  const int num_heads = gridDim.x, num_parts = gridDim.z;
  const int seq_len = seq_lens[seq];
  const bool partitioned = partition_size > 0;
  const int start_token = partitioned ? part * partition_size : 0;
  if (partitioned && start_token >= seq_len) return;  // attention_kernels.cu:116-119
  const int end_token = partitioned ? min(start_token + partition_size, seq_len) : seq_len;
  const int ntok = max(end_token - start_token, 0);
  const int start_block = start_token / block_size;
  const int end_block = (end_token + block_size - 1) / block_size;
  const int kv_head = head / (num_heads / num_kv_heads);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  float* q_s = smem;                                   // [head_size]
  float* red = q_s + ((head_size + 3) & ~3);           // [8]
  float* acc_s = red + 8;                              // [warps][head_size]
  float* logits = acc_s + kAttnWarps * head_size;      // [ntok rounded to block]

  for (int d = threadIdx.x; d < head_size; d += kAttnThreads)
    q_s[d] = to_f32<scalar_t>(q[(int64_t)seq * q_stride + (int64_t)head * head_size + d]);
  __syncthreads();

  const int* btab = block_tables + (int64_t)seq * max_blocks_per_seq;
  const float slope = alibi_slopes ? alibi_slopes[head] : 0.f;

  // ---- phase 1: logits ----
  const int nparts = 32 / block_size;  // lanes sharing one token (block_size in {8,16,32})
  const int t = lane % block_size, dpart = lane / block_size;
  const int nchunks = head_size / X;   // 16-byte chunks along D
  float qk_max = -3.402823466e38f;
  for (int blk = start_block + warp; blk < end_block; blk += kAttnWarps) {
    const int64_t phys = btab[blk];
    const cache_t* kp = k_cache + phys * kv_block_stride + (int64_t)kv_head * kv_head_stride + (int64_t)t * X;
    float dot = 0.f;
    for (int c = dpart; c < nchunks; c += nparts) {
      const int4 raw = *reinterpret_cast<const int4*>(kp + (int64_t)c * block_size * X);
      const cache_t* e = reinterpret_cast<const cache_t*>(&raw);
#pragma unroll
      for (int j = 0; j < X; ++j) dot += q_s[c * X + j] * cache_val<scalar_t, cache_t, KV>(e[j], kv_scale);
    }
    for (int o = block_size; o < 32; o <<= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
    const int token = blk * block_size + t;
    float logit = dot * scale;
    if (alibi_slopes) logit += slope * (float)(token - seq_len + 1);
    const bool valid = token < end_token;
    if (dpart == 0) logits[token - start_token] = valid ? logit : 0.f;
    if (valid) qk_max = fmaxf(qk_max, logit);
  }
  qk_max = block_reduce_max(qk_max, red);  // also publishes logits[] (barrier inside)

  // ---- phase 2: softmax ----
  float esum = 0.f;
  for (int i = threadIdx.x; i < ntok; i += kAttnThreads) {
    const float v = __expf(logits[i] - qk_max);
    logits[i] = v;
    esum += v;
  }
  esum = block_reduce_sum(esum, red);
  const float inv = __fdividef(1.f, esum + 1e-6f);
  if (partitioned && threadIdx.x == 0) {
    const int64_t o = ((int64_t)seq * num_heads + head) * num_parts + part;
    max_logits[o] = qk_max;
    exp_sums[o] = esum;
  }

  // ---- phase 3: P.V ----
  constexpr int RMAX = 8;  // head_size <= 256
  float acc[RMAX];
#pragma unroll
  for (int r = 0; r < RMAX; ++r) acc[r] = 0.f;
  for (int blk = start_block + warp; blk < end_block; blk += kAttnWarps) {
    const int64_t phys = btab[blk];
    const cache_t* vp = v_cache + phys * kv_block_stride + (int64_t)kv_head * kv_head_stride;
    const int tok0 = blk * block_size;
    const int nvalid = min(block_size, end_token - tok0);
#pragma unroll
    for (int r = 0; r < RMAX; ++r) {
      const int d = lane + 32 * r;
      if (d < head_size) {
        const cache_t* row = vp + (int64_t)d * block_size;
        if (block_size % X == 0) {
          for (int v0 = 0; v0 < block_size; v0 += X) {
            if (v0 >= nvalid) break;
            const int4 raw = *reinterpret_cast<const int4*>(row + v0);
            const cache_t* e = reinterpret_cast<const cache_t*>(&raw);
#pragma unroll
            for (int j = 0; j < X; ++j) {
              if (v0 + j < nvalid) {
                const float p = to_f32<scalar_t>(from_f32<scalar_t>(logits[tok0 - start_token + v0 + j] * inv));
                acc[r] += p * cache_val<scalar_t, cache_t, KV>(e[j], kv_scale);
              }
            }
          }
        } else {
          for (int j = 0; j < nvalid; ++j) {
            const float p = to_f32<scalar_t>(from_f32<scalar_t>(logits[tok0 - start_token + j] * inv));
            acc[r] += p * cache_val<scalar_t, cache_t, KV>(row[j], kv_scale);
          }
        }
      }
    }
  }
#pragma unroll
  for (int r = 0; r < RMAX; ++r) {
    const int d = lane + 32 * r;
    if (d < head_size) acc_s[warp * head_size + d] = acc[r];
  }
  __syncthreads();
  for (int d = threadIdx.x; d < head_size; d += kAttnThreads) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < kAttnWarps; ++w) s += acc_s[w * head_size + d];
    int64_t o = ((int64_t)seq * num_heads + head);
    if (partitioned) o = o * num_parts + part;
    out[o * head_size + d] = from_f32<scalar_t>(s);
  }
}

// v2 merge: reference attention_kernels.cu:567-669
template <typename scalar_t>
__global__ void __launch_bounds__(kAttnThreads)
paged_attn_v2_reduce_kernel(scalar_t* __restrict__ out, const float* __restrict__ exp_sums,
                            const float* __restrict__ max_logits, const scalar_t* __restrict__ tmp_out,
                            const int* __restrict__ seq_lens, int head_size, int max_parts) {
  extern __shared__ float smem[];
  pdl_wait();
  pdl_launch_dependents();
  const int head = blockIdx.x, seq = blockIdx.y, num_heads = gridDim.x;
  const int seq_len = seq_lens[seq];
  const int nparts = (seq_len + kPartition - 1) / kPartition;
  const int64_t base = ((int64_t)seq * num_heads + head);
  scalar_t* o = out + base * head_size;
  const scalar_t* tp = tmp_out + base * max_parts * head_size;
  if (nparts == 1) {  // single partition: plain copy (attention_kernels.cu:582-594)
    for (int d = threadIdx.x; d < head_size; d += blockDim.x) o[d] = tp[d];
    return;
  }
  float* w = smem;           // [max_parts]
  float* red = w + max_parts;  // [8]
  const float* ml = max_logits + base * max_parts;
  const float* es = exp_sums + base * max_parts;
  float m = -3.402823466e38f;
  for (int i = threadIdx.x; i < nparts; i += blockDim.x) m = fmaxf(m, ml[i]);
  m = block_reduce_max(m, red);
  float gs = 0.f;
  for (int i = threadIdx.x; i < nparts; i += blockDim.x) {
    const float v = es[i] * expf(ml[i] - m);
    w[i] = v;
    gs += v;
  }
  gs = block_reduce_sum(gs, red);
  const float inv = __fdividef(1.f, gs + 1e-6f);
  for (int d = threadIdx.x; d < head_size; d += blockDim.x) {
    float a = 0.f;
    for (int p = 0; p < nparts; ++p) a += to_f32<scalar_t>(tp[(int64_t)p * head_size + d]) * w[p] * inv;
    o[d] = from_f32<scalar_t>(a);
  }
}

static size_t generic_smem_bytes(int head_size, int block_size, int max_tokens) {
  const int padded = ((max_tokens + block_size - 1) / block_size) * block_size;
  return sizeof(float) * (size_t)(((head_size + 3) & ~3) + 8 + kAttnWarps * head_size + padded);
}

template <typename scalar_t, typename cache_t, int KV>
static int launch_generic_t(float* exp_sums, float* max_logits, void* out, const void* q, const void* kc,
                            const void* vc, int num_seqs, int num_heads, int num_kv_heads, int head_size, float scale,
                            const int* block_tables, const int* seq_lens, int max_blocks_per_seq, int block_size,
                            int max_seq_len, const float* alibi, int64_t q_stride, int64_t kvbs, int64_t kvhs,
                            int partition_size, int num_parts, float kv_scale, cudaStream_t st) {
  auto kern = paged_attn_generic_kernel<scalar_t, cache_t, KV>;
  const int max_tokens = partition_size > 0 ? partition_size : max_seq_len;
  const size_t smem = generic_smem_bytes(head_size, block_size, max_tokens);
  if (smem > 227 * 1024)
    return set_error(B200_ENOTSUP, "paged_attention_v1: max_seq_len %d needs %zu B of shared memory (> 227 KB); use v2",
                     max_seq_len, smem);
  if (smem > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  dim3 grid(num_heads, num_seqs, num_parts);
  launch_pdl(kern, grid, dim3(kAttnThreads), smem, st, exp_sums, max_logits, (scalar_t*)out, (const scalar_t*)q,
                                         (const cache_t*)kc, (const cache_t*)vc, num_kv_heads, scale, block_tables,
                                         seq_lens, max_blocks_per_seq, alibi, q_stride, kvbs, kvhs, head_size,
                                         block_size, partition_size, kv_scale);
  return B200_OK;
}

int launch_generic(float* exp_sums, float* max_logits, void* out, const void* q, const void* kc, const void* vc,
                   int num_seqs, int num_heads, int num_kv_heads, int head_size, float scale, const int* block_tables,
                   const int* seq_lens, int max_blocks_per_seq, int block_size, int max_seq_len, const float* alibi,
                   int64_t q_stride, int64_t kvbs, int64_t kvhs, int partition_size, int num_parts, int dtype,
                   int kv_dtype, float kv_scale, cudaStream_t st) {
#define GO(CT, KV)                                                                                                  \
  return launch_generic_t<scalar_t, CT, KV>(exp_sums, max_logits, out, q, kc, vc, num_seqs, num_heads, num_kv_heads, \
                                            head_size, scale, block_tables, seq_lens, max_blocks_per_seq, block_size, \
                                            max_seq_len, alibi, q_stride, kvbs, kvhs, partition_size, num_parts,     \
                                            kv_scale, st)
  B200_DISPATCH_FLOAT(dtype, "paged_attention", {
    if (kv_dtype == B200_KV_AUTO) { GO(scalar_t, B200_KV_AUTO); }
    else if (kv_dtype == B200_KV_FP8_E4M3) { GO(uint8_t, B200_KV_FP8_E4M3); }
    else if (kv_dtype == B200_KV_FP8_E5M2) { GO(uint8_t, B200_KV_FP8_E5M2); }
    else return set_error(B200_EINVAL, "Unsupported data type of kv cache: %d", kv_dtype);
  });
#undef GO
  return B200_OK;
}

int launch_v2_reduce(void* out, const float* exp_sums, const float* max_logits, const void* tmp_out,
                     const int* seq_lens, int num_seqs, int num_heads, int head_size, int max_parts, int dtype,
                     cudaStream_t st) {
  dim3 grid(num_heads, num_seqs);
  const size_t smem = sizeof(float) * (max_parts + 8);
  B200_DISPATCH_FLOAT(dtype, "paged_attention_v2", {
    launch_pdl(paged_attn_v2_reduce_kernel<scalar_t>, grid, dim3(kAttnThreads), smem, st,
        (scalar_t*)out, exp_sums, max_logits, (const scalar_t*)tmp_out, seq_lens, head_size, max_parts);
  });
  return B200_OK;
}

}  // namespace b200
