// C-ABI entry points for paged attention decode: validation + choice between the GQA tensor-core path
// (attention_mma.cu) and the all-shapes generic path (attention_generic.cu).
// Reference ops: csrc/attention/attention_kernels.cu:805-826 (v1), :966-990 (v2); dispatch preconditions
// :738-766, :789-803 (head sizes {64,80,96,112,128,192,256}, block sizes {8,16,32}).
#include "common.cuh"

namespace b200 {
int launch_generic(float* exp_sums, float* max_logits, void* out, const void* q, const void* kc, const void* vc,
                   int num_seqs, int num_heads, int num_kv_heads, int head_size, float scale, const int* block_tables,
                   const int* seq_lens, int max_blocks_per_seq, int block_size, int max_seq_len, const float* alibi,
                   int64_t q_stride, int64_t kvbs, int64_t kvhs, int partition_size, int num_parts, int dtype,
                   int kv_dtype, float kv_scale, cudaStream_t st);
int launch_v2_reduce(void* out, const float* exp_sums, const float* max_logits, const void* tmp_out,
                     const int* seq_lens, int num_seqs, int num_heads, int head_size, int max_parts, int dtype,
                     cudaStream_t st);
// returns 1 if handled, 0 if the shape is not covered by the fast path, <0 on error
int launch_mma(float* exp_sums, float* max_logits, void* out, void* tmp_out, const void* q, const void* kc,
               const void* vc, int num_seqs, int num_heads, int num_kv_heads, int head_size, float scale,
               const int* block_tables, const int* seq_lens, int max_blocks_per_seq, int block_size, int max_seq_len,
               const float* alibi, int64_t q_stride, int64_t kvbs, int64_t kvhs, int partitioned, int num_parts,
               int dtype, int kv_dtype, float kv_scale, cudaStream_t st);

static int check_common(const char* name, int num_heads, int num_kv_heads, int head_size, int block_size) {
  switch (head_size) {
    case 64: case 80: case 96: case 112: case 128: case 192: case 256: break;
    default: return set_error(B200_EINVAL, "Unsupported head size: %d", head_size);
  }
  if (block_size != 8 && block_size != 16 && block_size != 32)
    return set_error(B200_EINVAL, "Unsupported block size: %d", block_size);
  if (num_kv_heads <= 0 || num_heads % num_kv_heads != 0)
    return set_error(B200_EINVAL, "%s: num_heads %d not a multiple of num_kv_heads %d", name, num_heads, num_kv_heads);
  return B200_OK;
}
}  // namespace b200

extern "C" int b200_paged_attention_v1(void* out, const void* query, const void* key_cache, const void* value_cache,
                                       int num_seqs, int num_heads, int num_kv_heads, int head_size, float scale,
                                       const int32_t* block_tables, const int32_t* seq_lens, int max_blocks_per_seq,
                                       int block_size, int max_seq_len, const float* alibi_slopes, int64_t q_stride,
                                       int64_t kv_block_stride, int64_t kv_head_stride, int dtype, int kv_dtype,
                                       float kv_scale, void* stream) {
  using namespace b200;
  if (int e = check_common("paged_attention_v1", num_heads, num_kv_heads, head_size, block_size)) return e;
  if (num_seqs == 0) return B200_OK;
  cudaStream_t st = (cudaStream_t)stream;
  int r = launch_mma(nullptr, nullptr, out, nullptr, query, key_cache, value_cache, num_seqs, num_heads, num_kv_heads,
                     head_size, scale, block_tables, seq_lens, max_blocks_per_seq, block_size, max_seq_len,
                     alibi_slopes, q_stride, kv_block_stride, kv_head_stride, 0, 1, dtype, kv_dtype, kv_scale, st);
  if (r < 0) return r;
  if (r == 0) {
    r = launch_generic(nullptr, nullptr, out, query, key_cache, value_cache, num_seqs, num_heads, num_kv_heads,
                       head_size, scale, block_tables, seq_lens, max_blocks_per_seq, block_size, max_seq_len,
                       alibi_slopes, q_stride, kv_block_stride, kv_head_stride, 0, 1, dtype, kv_dtype, kv_scale, st);
    if (r < 0) return r;
  }
  B200_LAUNCH_CHECK("paged_attention_v1");
  return B200_OK;
}

extern "C" int b200_paged_attention_v2(void* out, float* exp_sums, float* max_logits, void* tmp_out,
                                       const void* query, const void* key_cache, const void* value_cache,
                                       int num_seqs, int num_heads, int num_kv_heads, int head_size, float scale,
                                       const int32_t* block_tables, const int32_t* seq_lens, int max_blocks_per_seq,
                                       int block_size, int max_seq_len, const float* alibi_slopes, int64_t q_stride,
                                       int64_t kv_block_stride, int64_t kv_head_stride, int dtype, int kv_dtype,
                                       float kv_scale, void* stream) {
  using namespace b200;
  if (int e = check_common("paged_attention_v2", num_heads, num_kv_heads, head_size, block_size)) return e;
  if (num_seqs == 0) return B200_OK;
  const int num_parts = (max_seq_len + 511) / 512;
  B200_CHECK(num_parts >= 1, "paged_attention_v2: max_seq_len must be positive");
  cudaStream_t st = (cudaStream_t)stream;
  int r = launch_mma(exp_sums, max_logits, out, tmp_out, query, key_cache, value_cache, num_seqs, num_heads,
                     num_kv_heads, head_size, scale, block_tables, seq_lens, max_blocks_per_seq, block_size,
                     max_seq_len, alibi_slopes, q_stride, kv_block_stride, kv_head_stride, 1, num_parts, dtype,
                     kv_dtype, kv_scale, st);
  if (r < 0) return r;
  if (r == 0) {
    r = launch_generic(exp_sums, max_logits, tmp_out, query, key_cache, value_cache, num_seqs, num_heads, num_kv_heads,
                       head_size, scale, block_tables, seq_lens, max_blocks_per_seq, block_size, max_seq_len,
                       alibi_slopes, q_stride, kv_block_stride, kv_head_stride, 512, num_parts, dtype, kv_dtype,
                       kv_scale, st);
    if (r < 0) return r;
  }
  r = launch_v2_reduce(out, exp_sums, max_logits, tmp_out, seq_lens, num_seqs, num_heads, head_size, num_parts, dtype, st);
  if (r < 0) return r;
  B200_LAUNCH_CHECK("paged_attention_v2");
  return B200_OK;
}
