/*
 * b200_ops.h -- C ABI of the B200-native decode hot path (libb200ops.so).
 *
 * Every entry point replaces one operator of nm-vllm's native extension for this path; the reference
 * passes torch::Tensor& straight into templated launchers, so the "interface replaced" is the op
 * schema in /root/reference/csrc/torch_bindings.cpp and the declaration in csrc/ops.h / csrc/cache.h
 * (file:line cited per function).  This ABI is what a binding in the reference (its torch_bindings.cpp,
 * or a ctypes stub in vllm/_custom_ops.py) would call: plain pointers, sizes, strides (in ELEMENTS),
 * enums and a cudaStream_t passed as void*.  No torch types.
 *
 * Conventions
 *   - all pointers are DEVICE pointers on the current CUDA device unless stated otherwise;
 *   - `stream` is a cudaStream_t (may be 0); every call is asynchronous on it, never synchronises the
 *     host, never allocates while a CUDA graph is being captured (scratch is allocated by the first
 *     eager call, see b200_gemm_scratch_reserve);
 *   - return value 0 = launched; negative = rejected before any launch (B200_E*), message via
 *     b200_last_error() (thread-local);
 *   - integer results (repack, cache slot writes for kv dtype "auto") are bit-exact w.r.t. the reference;
 *     floating point results carry the tolerance stated in DESIGN.md.
 */
#ifndef B200_OPS_H_
#define B200_OPS_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_OK 0
#define B200_EINVAL (-1)   /* shape / alignment / dtype precondition violated (reference: TORCH_CHECK) */
#define B200_ENOTSUP (-2)  /* combination the reference rejects or this build does not cover           */
#define B200_ECUDA (-3)    /* CUDA runtime reported an error at launch                                 */

/* element type of activations / outputs */
enum b200_dtype { B200_F16 = 0, B200_BF16 = 1, B200_F32 = 2 };
/* KV-cache storage: "auto" (= same as activations), "fp8" == "fp8_e4m3", "fp8_e5m2"
 * (reference: csrc/quantization/fp8/nvidia/quant_utils.cuh:529-566) */
enum b200_kv_dtype { B200_KV_AUTO = 0, B200_KV_FP8_E4M3 = 1, B200_KV_FP8_E5M2 = 2 };
/* 8-bit operand type of scaled_mm */
enum b200_q8_dtype { B200_Q8_INT8 = 0, B200_Q8_FP8_E4M3 = 1 };

const char* b200_last_error(void);
/* library self-description: "sm_100a;<git-less build tag>" */
const char* b200_build_info(void);

/* ---- device queries: csrc/cuda_utils_kernels.cu (torch_bindings.cpp:250-258) ---------------------- */
int b200_get_device_attribute(int attribute, int device, int* value_out);
int b200_get_max_shared_memory_per_block_device_attribute(int device, int* value_out);

/* ---- KV cache: csrc/cache.h:18-32, csrc/cache_kernels.cu:253-278 (op), :153-204 (kernel) ---------- */
/* key,value [T,Hkv,D] with row strides key_stride/value_stride (elements); key_cache
 * [NB,Hkv,D/x,BS,x], value_cache [NB,Hkv,D,BS]; slot_mapping int64 [T], slot<0 skipped. */
int b200_reshape_and_cache(const void* key, const void* value, void* key_cache, void* value_cache,
                           const int64_t* slot_mapping, int64_t num_tokens, int num_kv_heads,
                           int head_size, int block_size, int x, int64_t key_stride,
                           int64_t value_stride, int dtype, int kv_dtype, float kv_scale,
                           void* stream);

/* convert_fp8: csrc/cache_kernels.cu:339-389.  Exactly one side is fp8 (uint8 storage).
 * to_fp8 != 0: src is `dtype`, dst is fp8; else src is fp8, dst is `dtype`. */
int b200_convert_fp8(void* dst, const void* src, int64_t numel, int dtype, int kv_dtype, float scale,
                     int to_fp8, void* stream);

/* ---- paged attention: csrc/ops.h:6-29, csrc/attention/attention_kernels.cu:805-826 / :966-990 ----- */
/* query/out [S,H,D] (q_stride = query.stride(0), out is contiguous), caches as above with
 * kv_block_stride = key_cache.stride(0), kv_head_stride = key_cache.stride(1); block_tables int32
 * [S,max_blocks_per_seq]; seq_lens int32 [S]; alibi_slopes fp32 [H] or NULL. */
int b200_paged_attention_v1(void* out, const void* query, const void* key_cache,
                            const void* value_cache, int num_seqs, int num_heads, int num_kv_heads,
                            int head_size, float scale, const int32_t* block_tables,
                            const int32_t* seq_lens, int max_blocks_per_seq, int block_size,
                            int max_seq_len, const float* alibi_slopes, int64_t q_stride,
                            int64_t kv_block_stride, int64_t kv_head_stride, int dtype, int kv_dtype,
                            float kv_scale, void* stream);
/* exp_sums,max_logits fp32 [S,H,P]; tmp_out [S,H,P,D] of `dtype`; P = ceil(max_seq_len/512)
 * (PARTITION_SIZE is fixed on both sides: attention_kernels.cu:847, vllm/attention/ops/paged_attn.py:10) */
int b200_paged_attention_v2(void* out, float* exp_sums, float* max_logits, void* tmp_out,
                            const void* query, const void* key_cache, const void* value_cache,
                            int num_seqs, int num_heads, int num_kv_heads, int head_size, float scale,
                            const int32_t* block_tables, const int32_t* seq_lens,
                            int max_blocks_per_seq, int block_size, int max_seq_len,
                            const float* alibi_slopes, int64_t q_stride, int64_t kv_block_stride,
                            int64_t kv_head_stride, int dtype, int kv_dtype, float kv_scale,
                            void* stream);

/* ---- activation quantizers: csrc/ops.h:127-141, fp8/common.cu:129-165, int8_quant_kernels.cu:75-115 */
int b200_static_scaled_fp8_quant(void* out, const void* input, const float* scale, int64_t numel,
                                 int dtype, void* stream);
/* *scale must be zero on entry (reference: vllm/_custom_ops.py:316); written as max|x|/448 */
int b200_dynamic_scaled_fp8_quant(void* out, const void* input, float* scale, int64_t numel,
                                  int dtype, void* stream);
int b200_static_scaled_int8_quant(int8_t* out, const void* input, const float* scale,
                                  int64_t num_tokens, int hidden_size, int dtype, void* stream);
int b200_dynamic_scaled_int8_quant(int8_t* out, const void* input, float* scales,
                                   int64_t num_tokens, int hidden_size, int dtype, void* stream);

/* ---- Marlin weight repack: csrc/ops.h:92-94, gptq_marlin_repack.cu:276-348 ------------------------ */
/* b_q_weight int32 [K/pf,N] (GPTQ row-packed), perm int32 [K] or NULL, out int32 [K/16, N*16/pf] */
int b200_gptq_marlin_repack(const int32_t* b_q_weight, const int32_t* perm, int32_t* out,
                            int64_t size_k, int64_t size_n, int num_bits, void* stream);

/* ---- Marlin W4A16 / W8A16 GEMM: csrc/ops.h:75-77 (marlin_gemm), :86-90 (gptq_marlin_gemm) ---------
 * c[M,N] = a[M,K] . dequant(b_q_weight) ; a,c,b_scales of `dtype` (F16|BF16); b_q_weight in the Marlin
 * tile layout produced by b200_gptq_marlin_repack; b_scales [num_groups,N] permuted by
 * marlin_permute_scales; g_idx/perm int32 [K] or NULL (act-order); workspace int32 >= (N/64)*16 zeros,
 * left zero; num_groups = b_scales.size(0).  a row stride = lda elements. */
int b200_gptq_marlin_gemm(void* c, const void* a, const int32_t* b_q_weight, const void* b_scales,
                          const int32_t* g_idx, const int32_t* perm, int32_t* workspace,
                          int64_t workspace_numel, int num_bits, int64_t size_m, int64_t size_n,
                          int64_t size_k, int64_t lda, int num_groups, int is_k_full, int dtype,
                          void* stream);

/* Fused gate_up GEMM + SiluAndMul (decode step: MergedColumnParallelLinear gate_up_proj followed by SiluAndMul,
 * vllm/model_executor/models/llama.py:74-90; replaces gptq_marlin_gemm + silu_and_mul, csrc/activation_kernels.cu:14-76,
 * with one launch): out[M, N/2] = silu(g) * u where [g | u] = a . dequant(b_q_weight) is rounded to `dtype` first, so the
 * result is bit-identical to the two-op chain.  Stored layouts are unchanged (the kernel pairs column n with n + N/2).
 * size_n is the GEMM's N (= 2 x intermediate size, N/2 % 64 == 0); no act-order. */
int b200_gptq_marlin_gemm_silu_and_mul(void* out, const void* a, const int32_t* b_q_weight, const void* b_scales,
                                       int32_t* workspace, int64_t workspace_numel, int num_bits, int64_t size_m,
                                       int64_t size_n, int64_t size_k, int64_t lda, int num_groups, int dtype,
                                       void* stream);

/* ---- Sparse-Marlin 2:4: csrc/ops.h:79-84, marlin/sparse/marlin_24_cuda_kernel.cu:1017-1125 --------
 * b_q_weight int32 [K/32, N*16/pf]; b_meta int16 [K/32, 2N] (viewed [K/16... see DESIGN.md);
 * b_scales fp16 [num_groups,N]; fp16 only (gptq_marlin_24.py:145-147). */
int b200_gptq_marlin_24_gemm(void* c, const void* a, const int32_t* b_q_weight, const int16_t* b_meta,
                             const void* b_scales, int32_t* workspace, int64_t workspace_numel,
                             int num_bits, int64_t size_m, int64_t size_n, int64_t size_k,
                             int64_t lda, int num_groups, int dtype, void* stream);

/* ---- Sparse-Marlin 2:4 without any allocation ---------------------------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arguments,
 * preconditions, kernel and results as b200_gptq_marlin_24_gemm; the K-split scratch is a file-scope
 * `static __device__` array of the sizes below (20 MiB + 16 KiB, held by every GPU that loads the library
 * until exit), so the call performs no cudaMalloc / cudaFree / allocating library routine.  A problem whose
 * K-split needs more (the decomposition is sized from the SM count: every shape fits on a 148-SM B200)
 * returns B200_EINVAL with the required and available sizes.  All calls share the array: calls that overlap
 * on the GPU (different streams or threads on one device) race; calls on one stream are safe. */
#define B200_M24_STATIC_MAX_PARTIAL_FLOATS (5u * 1024u * 1024u)
#define B200_M24_STATIC_MAX_COUNTERS 4096
int b200_gptq_marlin_24_gemm_static(void* c, const void* a, const int32_t* b_q_weight,
                                    const int16_t* b_meta, const void* b_scales, int32_t* workspace,
                                    int64_t workspace_numel, int num_bits, int64_t size_m,
                                    int64_t size_n, int64_t size_k, int64_t lda, int num_groups,
                                    int dtype, void* stream);

/* ---- W8A8 scaled GEMM: csrc/ops.h:96-102, cutlass_w8a8/scaled_mm_entry.cu:48-100 -------------------
 * out[M,N] = a_scales (.) (b_scales (.) (a[M,K] . b[K,N])) (+ bias); a row-major (lda elements),
 * b COLUMN-major (ldb = b.stride(1) elements), out row-major (ldc); a_scales fp32 numel 1 or M,
 * b_scales fp32 numel 1 or N; bias [N] of out dtype or NULL. */
int b200_scaled_mm(void* out, const void* a, const void* b, const float* a_scales,
                   const float* b_scales, const void* bias, int64_t size_m, int64_t size_n,
                   int64_t size_k, int64_t lda, int64_t ldb, int64_t ldc, int a_scales_numel,
                   int b_scales_numel, int q8_dtype, int out_dtype, void* stream);
/* reference: scaled_mm_entry.cu:32-46 */
int b200_scaled_mm_supports_fp8(int cuda_device_capability);

/* ---- fused decode-step attention (new): rotary_embedding (csrc/pos_encoding_kernels.cu:71-203, neox,
 * rot_dim == head_size == 128) + reshape_and_cache (csrc/cache_kernels.cu:153-204) + paged_attention_v1/v2
 * (csrc/attention/attention_kernels.cu:805-990) in one launch (two with the v2 merge).  query is NOT modified (rotated in
 * registers); the rotated key only exists in the cache.  slot_mapping[s] must be the slot of token seq_lens[s]-1.
 * B200_ENOTSUP when the shape is outside the GQA fast path: call the three ops instead. */
int b200_rope_cache_paged_attention(void* out, float* exp_sums, float* max_logits, void* tmp_out,
                                    const void* query, const void* key, const void* value, void* key_cache,
                                    void* value_cache, const int64_t* positions, const void* cos_sin_cache,
                                    int rot_dim, const int64_t* slot_mapping, int num_seqs, int num_heads,
                                    int num_kv_heads, int head_size, float scale, const int32_t* block_tables,
                                    const int32_t* seq_lens, int max_blocks_per_seq, int block_size,
                                    int max_seq_len, int64_t q_stride, int64_t k_stride, int64_t v_stride,
                                    int64_t kv_block_stride, int64_t kv_head_stride, int dtype, int kv_dtype,
                                    float kv_scale, int use_v2, void* stream);

/* ---- scratch management (new; see DESIGN.md "split-K partials") -------------------------------------
 * Ensures the per-device split-K scratch is allocated for GEMMs up to max_m rows.  Called implicitly
 * by the first eager GEMM; call explicitly before CUDA-graph capture. */
int b200_gemm_scratch_reserve(int64_t max_m, int64_t max_n);

/* ---- decode-step glue: layernorm/pos_encoding/activation kernels ------ */
int b200_rms_norm(void* out, const void* input, const void* weight, int64_t num_tokens,
                  int hidden_size, float epsilon, int dtype, void* stream);
int b200_fused_add_rms_norm(void* input, void* residual, const void* weight, int64_t num_tokens,
                            int hidden_size, float epsilon, int dtype, void* stream);
int b200_silu_and_mul(void* out, const void* input, int64_t num_tokens, int d, int dtype,
                      void* stream);
/* ---- glue fused with the activation quantiser of the following W8A8 linear (new): the 16-bit result of
 * rms_norm / fused_add_rms_norm (csrc/layernorm_kernels.cu:22-352) or silu_and_mul (csrc/activation_kernels.cu:14-162) is
 * quantised in the same kernel exactly like static_scaled_fp8_quant (csrc/quantization/fp8/common.cu:74-125) or
 * static / dynamic_scaled_int8_quant (compressed_tensors/int8_quant_kernels.cu:30-71) would -- bit-identical to the two-op
 * chain.  qkind 1: fp8 e4m3, static per-tensor scale *scales; 2: int8, static per-tensor scale *scales; 3: int8, dynamic
 * per-token scales WRITTEN to scales[num_tokens]; 4: fp8 e4m3 with the DYNAMIC per-tensor scale max|x| / 448 of
 * dynamic_scaled_fp8_quant (fp8/common.cu:39-72) computed inside the same launch and WRITTEN to scales[0] (grid-wide
 * maximum through a static device word and an in-kernel barrier: num_tokens <= 512 or B200_ENOTSUP; calls that overlap on
 * one GPU race; measured slower than the separate kernels at decode sizes, so nothing selects it by default).
 * residual != NULL: fused_add_rms_norm (residual updated in place). */
int b200_rms_norm_quant(void* out_q, float* scales, const void* input, void* residual, const void* weight,
                        int64_t num_tokens, int hidden_size, float epsilon, int dtype, int qkind, void* stream);
int b200_silu_and_mul_quant(void* out_q, float* scales, const void* input, int64_t num_tokens, int d, int dtype,
                            int qkind, void* stream);
/* positions int64 [T]; query [T, Hq*D] / key [T, Hkv*D] rotated in place; cos_sin_cache [max_pos, rot_dim] */
int b200_rotary_embedding(const int64_t* positions, void* query, void* key, int head_size,
                          const void* cos_sin_cache, int rot_dim, int is_neox, int64_t num_tokens,
                          int num_heads, int num_kv_heads, int64_t query_stride, int64_t key_stride,
                          int dtype, void* stream);

/* greedy sampling: out[t] = argmax_v logits[t, v] (lowest index on ties, like torch.argmax on CUDA); logits
 * [num_tokens, vocab] with row pitch row_stride (elements).  Reference: vllm/model_executor/layers/sampler.py
 * _greedy_sample (torch.argmax); here so that the whole decode step is graph-captured kernels of this library. */
int b200_greedy_sample(int64_t* out, const void* logits, int64_t num_tokens, int64_t vocab,
                       int64_t row_stride, int dtype, void* stream);

/* ---- KV block maintenance: csrc/cache.h:8-32 ----------------------------
 * copy_blocks (cache_kernels.cu:64-148): key_cache_ptrs / value_cache_ptrs are DEVICE arrays of
 * num_layers device pointers, block_mapping a DEVICE int64 [num_pairs, 2] of (src, dst) block numbers;
 * block_bytes = bytes of one block of one layer's key (== value) cache. */
int b200_copy_blocks(const int64_t* key_cache_ptrs, const int64_t* value_cache_ptrs,
                     const int64_t* block_mapping, int num_layers, int64_t num_pairs,
                     int64_t block_bytes, void* stream);
/* swap_blocks (cache_kernels.cu:24-62): block_mapping_host is a HOST int64 [num_pairs, 2];
 * direction 0 = device->device, 1 = device->host, 2 = host->device. */
int b200_swap_blocks(const void* src, void* dst, const int64_t* block_mapping_host, int64_t num_pairs,
                     int64_t block_bytes, int direction, void* stream);
/* reshape_and_cache_flash (cache_kernels.cu:207-236, 280-314): caches [NB, BS, H, D], kv dtype auto only. */
int b200_reshape_and_cache_flash(const void* key, const void* value, void* k_cache, void* v_cache,
                                 const int64_t* slot_mapping, int64_t num_tokens, int num_heads,
                                 int head_size, int block_size, int64_t block_stride,
                                 int64_t key_stride, int64_t value_stride, int dtype, int kv_dtype,
                                 void* stream);

/* ---- one-shot all-reduce over NVLink peer memory -----------------
 * Replaces csrc/custom_all_reduce.cu (init_custom_ar / all_reduce_reg / dispose, torch_bindings.cpp:262-293)
 * for the row-parallel GEMM outputs of TP decode.  Every rank allocates one region with b200_ar_alloc
 * (cudaMalloc + IPC handle, zero-filled), exchanges the 64-byte handles out of band (torch.distributed),
 * maps the peers' regions with b200_ar_open and builds a context with b200_ar_init(region_ptrs[world],
 * own region at index `rank`).  b200_ar_all_reduce sums `numel` elements of `in` over the ranks into `out`
 * (may alias `in`), fp32 accumulation in rank order (bit-identical on every rank); asynchronous on
 * `stream`, CUDA-graph capturable; B200_ENOTSUP when numel*sizeof > max_bytes (caller falls back to NCCL,
 * like should_custom_ar, custom_all_reduce.cu:32-50).  Every rank must issue the same sequence of calls. */
int b200_ar_region_bytes(int world, int64_t max_bytes, int64_t* bytes_out);
int b200_ar_alloc(int64_t bytes, void** ptr_out, void* handle_out64);
int b200_ar_open(const void* handle64, void** ptr_out);
int b200_ar_close(void* ptr);
int b200_ar_free(void* ptr);
int b200_ar_init(int rank, int world, void* const* region_ptrs, int64_t max_bytes, int64_t* ctx_out);
int b200_ar_all_reduce(int64_t ctx, const void* in, void* out, int64_t numel, int dtype, void* stream);
int b200_ar_dispose(int64_t ctx);
/* Fused GEMM + all-reduce: the NEXT b200_gptq_marlin_gemm issued by this host thread, whose output must be [m, n],
 * exchanges its output tiles with the peers inside its epilogue (flag-in-data cells in the peers' regions) and writes
 * the all-reduced result -- bit-identical to b200_gptq_marlin_gemm followed by b200_ar_all_reduce.  B200_ENOTSUP when
 * the shape does not qualify (m > 128, the cells do not fit the exchange area, or payload x (world-1) is beyond the
 * one-shot range where the stand-alone two-shot kernel is faster): run the two calls separately then.  Replaces RowParallelLinear.forward's GEMM -> tensor_model_parallel_all_reduce pair
 * (vllm/model_executor/layers/linear.py:779-803). */
int b200_ar_fuse_next_gemm(int64_t ctx, int64_t m, int64_t n);

/* ---- prefill attention over the paged cache: context_attention_fwd, vllm/attention/ops/prefix_prefill.py:674-799
 * (Triton kernels :14-672 in the reference; caller PagedAttention.forward_prefix, vllm/attention/ops/paged_attn.py:185-216).
 * o,q [tokens, num_heads, head_size]; k,v [tokens, num_kv_heads, head_size] (the call's new tokens); k_cache
 * [NB, Hkv, head_size/8, block_size, 8], v_cache [NB, Hkv, head_size, block_size] of the same 16-bit dtype;
 * block_tables int32 [batch, *] (row stride bt_stride); b_start_loc / b_seq_len / b_ctx_len int32 [batch]: first
 * query token, total length and cached-context length of each sequence.  Query token i of a sequence attends to its
 * b_ctx_len cached tokens and to new tokens 0..i; scale = 1/sqrt(head_size); sliding_window <= 0 disables the window;
 * alibi_slopes float [num_heads] or NULL.  qkvo_strides = element strides {q tok, q head, k tok, k head, v tok, v head,
 * o tok, o head}.  head_size % 8 == 0 and <= 128 (B200_ENOTSUP above). */
int b200_context_attention_fwd(void* o, const void* q, const void* k, const void* v, const void* k_cache,
                               const void* v_cache, const int32_t* block_tables, int64_t bt_stride,
                               const int32_t* b_start_loc, const int32_t* b_seq_len, const int32_t* b_ctx_len,
                               int batch, int max_input_len, int num_heads, int num_kv_heads, int head_size,
                               int block_size, int x, const int64_t* qkvo_strides, int64_t kc_block_stride,
                               int64_t kc_head_stride, int64_t vc_block_stride, int64_t vc_head_stride,
                               const float* alibi_slopes, int sliding_window, int dtype, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200_OPS_H_ */
