// Torch-side shim: registers the SAME operator schemas as the reference's csrc/torch_bindings.cpp:18-259 under a
// private namespace (_b200_C, _b200_C_cache_ops, _b200_C_cuda_utils; the name _C belongs to whichever vLLM is
// installed) and forwards to the C ABI of libb200ops.so (include/b200_ops.h).  Preconditions are the reference's
// TORCH_CHECKs restated (files cited per op); C-ABI error codes become c10::Error -> Python RuntimeError.
// No kernels here and no CPU fallback: every op is registered for torch::kCUDA only.
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <torch/all.h>
#include <torch/library.h>

#include <optional>
#include <string>

#include "../../include/b200_ops.h"

namespace {

inline void raise_if(int rc, const char* op) {
  TORCH_CHECK(rc == 0, op, ": ", b200_last_error());
}

inline int dtype_enum(at::ScalarType t, const char* what) {
  switch (t) {
    case at::ScalarType::Half: return B200_F16;
    case at::ScalarType::BFloat16: return B200_BF16;
    case at::ScalarType::Float: return B200_F32;
    default: TORCH_CHECK(false, what, ": unsupported dtype ", t);
  }
  return -1;
}

// reference: DISPATCH_BY_KV_CACHE_DTYPE, csrc/quantization/fp8/nvidia/quant_utils.cuh:529-566
inline int kv_enum(const std::string& s) {
  if (s == "auto") return B200_KV_AUTO;
  if (s == "fp8" || s == "fp8_e4m3") return B200_KV_FP8_E4M3;
  if (s == "fp8_e5m2") return B200_KV_FP8_E5M2;
  TORCH_CHECK(false, "Unsupported data type of kv cache: ", s);
  return -1;
}

inline void* cur_stream(const torch::Tensor& t) { return (void*)at::cuda::getCurrentCUDAStream(t.get_device()).stream(); }

// ---------------------------------------------------------------- attention (attention_kernels.cu:690-767, 805-826)
void paged_attention_v1(torch::Tensor& out, torch::Tensor& query, torch::Tensor& key_cache, torch::Tensor& value_cache,
                        int64_t num_kv_heads, double scale, torch::Tensor& block_tables, torch::Tensor& seq_lens,
                        int64_t block_size, int64_t max_seq_len, const std::optional<torch::Tensor>& alibi_slopes,
                        const std::string& kv_cache_dtype, double kv_scale, const int64_t tp_rank,
                        const int64_t blocksparse_local_blocks, const int64_t blocksparse_vert_stride,
                        const int64_t blocksparse_block_size, const int64_t blocksparse_head_sliding_step) {
  TORCH_CHECK(blocksparse_vert_stride <= 1, "blocksparse attention is outside this build's scope (SURVEY 2a #19)");
  TORCH_CHECK(block_tables.scalar_type() == at::ScalarType::Int && seq_lens.scalar_type() == at::ScalarType::Int,
              "block_tables and seq_lens must be int32");
  TORCH_CHECK(query.stride(-1) == 1 && query.stride(1) == query.size(2), "query heads must be contiguous");
  TORCH_CHECK(out.is_contiguous(), "out must be contiguous");
  const at::cuda::OptionalCUDAGuard device_guard(device_of(query));
  const float* alibi = alibi_slopes ? reinterpret_cast<const float*>(alibi_slopes->data_ptr()) : nullptr;
  raise_if(b200_paged_attention_v1(out.data_ptr(), query.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(),
                                   (int)query.size(0), (int)query.size(1), (int)num_kv_heads, (int)query.size(2),
                                   (float)scale, block_tables.data_ptr<int>(), seq_lens.data_ptr<int>(),
                                   (int)block_tables.size(1), (int)block_size, (int)max_seq_len, alibi,
                                   query.stride(0), key_cache.stride(0), key_cache.stride(1),
                                   dtype_enum(query.scalar_type(), "paged_attention_v1"), kv_enum(kv_cache_dtype),
                                   (float)kv_scale, cur_stream(query)),
           "paged_attention_v1");
}

void paged_attention_v2(torch::Tensor& out, torch::Tensor& exp_sums, torch::Tensor& max_logits, torch::Tensor& tmp_out,
                        torch::Tensor& query, torch::Tensor& key_cache, torch::Tensor& value_cache,
                        int64_t num_kv_heads, double scale, torch::Tensor& block_tables, torch::Tensor& seq_lens,
                        int64_t block_size, int64_t max_seq_len, const std::optional<torch::Tensor>& alibi_slopes,
                        const std::string& kv_cache_dtype, double kv_scale, const int64_t tp_rank,
                        const int64_t blocksparse_local_blocks, const int64_t blocksparse_vert_stride,
                        const int64_t blocksparse_block_size, const int64_t blocksparse_head_sliding_step) {
  TORCH_CHECK(blocksparse_vert_stride <= 1, "blocksparse attention is outside this build's scope (SURVEY 2a #19)");
  TORCH_CHECK(block_tables.scalar_type() == at::ScalarType::Int && seq_lens.scalar_type() == at::ScalarType::Int,
              "block_tables and seq_lens must be int32");
  TORCH_CHECK(query.stride(-1) == 1 && query.stride(1) == query.size(2), "query heads must be contiguous");
  TORCH_CHECK(out.is_contiguous() && tmp_out.is_contiguous() && exp_sums.is_contiguous() && max_logits.is_contiguous());
  const int64_t parts = (max_seq_len + 511) / 512;
  TORCH_CHECK(exp_sums.size(-1) == parts && max_logits.size(-1) == parts && tmp_out.size(2) == parts,
              "exp_sums/max_logits/tmp_out must have ceil(max_seq_len/512) = ", parts, " partitions");
  const at::cuda::OptionalCUDAGuard device_guard(device_of(query));
  const float* alibi = alibi_slopes ? reinterpret_cast<const float*>(alibi_slopes->data_ptr()) : nullptr;
  raise_if(b200_paged_attention_v2(out.data_ptr(), exp_sums.data_ptr<float>(), max_logits.data_ptr<float>(),
                                   tmp_out.data_ptr(), query.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(),
                                   (int)query.size(0), (int)query.size(1), (int)num_kv_heads, (int)query.size(2),
                                   (float)scale, block_tables.data_ptr<int>(), seq_lens.data_ptr<int>(),
                                   (int)block_tables.size(1), (int)block_size, (int)max_seq_len, alibi,
                                   query.stride(0), key_cache.stride(0), key_cache.stride(1),
                                   dtype_enum(query.scalar_type(), "paged_attention_v2"), kv_enum(kv_cache_dtype),
                                   (float)kv_scale, cur_stream(query)),
           "paged_attention_v2");
}

// norm / activation fused with the activation quantiser of the following W8A8 linear.  kind: "fp8" (static per-tensor scale),
// "int8" (static per-tensor scale), "int8_token" (dynamic per-token scales, written to `scales` [tokens, 1])
static int qkind_enum(const std::string& kind) {
  if (kind == "fp8") return 1;
  if (kind == "int8") return 2;
  if (kind == "int8_token") return 3;
  if (kind == "fp8_dynamic") return 4;
  TORCH_CHECK(false, "unknown quantisation kind: ", kind);
  return 0;
}
void rms_norm_quant(torch::Tensor& out_q, torch::Tensor& scales, torch::Tensor& input, const std::optional<torch::Tensor>& residual,
                    torch::Tensor& weight, double epsilon, const std::string& kind) {
  TORCH_CHECK(input.is_contiguous() && out_q.is_contiguous() && scales.scalar_type() == at::ScalarType::Float);
  TORCH_CHECK(out_q.element_size() == 1 && out_q.numel() == input.numel(), "out_q must be a 1-byte tensor of input's shape");
  const int qk = qkind_enum(kind);
  const int64_t hidden = input.size(-1), tokens = input.numel() / hidden;
  TORCH_CHECK(qk != 3 || scales.numel() >= tokens, "int8_token: scales must hold one float per token");
  const at::cuda::OptionalCUDAGuard device_guard(device_of(input));
  raise_if(b200_rms_norm_quant(out_q.data_ptr(), scales.data_ptr<float>(), input.data_ptr(),
                               residual ? residual->data_ptr() : nullptr, weight.data_ptr(), tokens, (int)hidden, (float)epsilon,
                               dtype_enum(input.scalar_type(), "rms_norm_quant"), qk, cur_stream(input)),
           "rms_norm_quant");
}
void silu_and_mul_quant(torch::Tensor& out_q, torch::Tensor& scales, torch::Tensor& input, const std::string& kind) {
  TORCH_CHECK(input.is_contiguous() && out_q.is_contiguous() && scales.scalar_type() == at::ScalarType::Float);
  const int qk = qkind_enum(kind);
  const int64_t d = input.size(-1) / 2, tokens = input.numel() / input.size(-1);
  TORCH_CHECK(out_q.element_size() == 1 && out_q.numel() == tokens * d, "out_q must be a 1-byte tensor [tokens, d]");
  TORCH_CHECK(qk != 3 || scales.numel() >= tokens, "int8_token: scales must hold one float per token");
  const at::cuda::OptionalCUDAGuard device_guard(device_of(input));
  raise_if(b200_silu_and_mul_quant(out_q.data_ptr(), scales.data_ptr<float>(), input.data_ptr(), tokens, (int)d,
                                   dtype_enum(input.scalar_type(), "silu_and_mul_quant"), qk, cur_stream(input)),
           "silu_and_mul_quant");
}

// fused rotary_embedding + reshape_and_cache + paged_attention_v1/v2 (decode); returns false when the shape is outside the
// fast path and NOTHING was launched (the caller then runs the three reference ops)
bool rope_cache_paged_attention(torch::Tensor& out, const std::optional<torch::Tensor>& exp_sums,
                                const std::optional<torch::Tensor>& max_logits, const std::optional<torch::Tensor>& tmp_out,
                                torch::Tensor& query, torch::Tensor& key, torch::Tensor& value, torch::Tensor& key_cache,
                                torch::Tensor& value_cache, torch::Tensor& positions, torch::Tensor& cos_sin_cache,
                                torch::Tensor& slot_mapping, int64_t num_kv_heads, double scale,
                                torch::Tensor& block_tables, torch::Tensor& seq_lens, int64_t block_size,
                                int64_t max_seq_len, const std::string& kv_cache_dtype, double kv_scale) {
  TORCH_CHECK(block_tables.scalar_type() == at::ScalarType::Int && seq_lens.scalar_type() == at::ScalarType::Int,
              "block_tables and seq_lens must be int32");
  TORCH_CHECK(slot_mapping.scalar_type() == at::ScalarType::Long && positions.scalar_type() == at::ScalarType::Long,
              "slot_mapping and positions must be int64");
  TORCH_CHECK(query.dim() == 3 && key.dim() == 3 && value.dim() == 3, "query/key/value must be [tokens, heads, head_size]");
  TORCH_CHECK(query.stride(-1) == 1 && query.stride(1) == query.size(2), "query heads must be contiguous");
  TORCH_CHECK(key.stride(-1) == 1 && key.stride(1) == key.size(2) && value.stride(-1) == 1 && value.stride(1) == value.size(2),
              "key/value heads must be contiguous");
  TORCH_CHECK(out.is_contiguous() && cos_sin_cache.is_contiguous());
  const bool v2 = exp_sums.has_value();
  const at::cuda::OptionalCUDAGuard device_guard(device_of(query));
  int rc = b200_rope_cache_paged_attention(
      out.data_ptr(), v2 ? exp_sums->data_ptr<float>() : nullptr, v2 ? max_logits->data_ptr<float>() : nullptr,
      v2 ? tmp_out->data_ptr() : nullptr, query.data_ptr(), key.data_ptr(), value.data_ptr(), key_cache.data_ptr(),
      value_cache.data_ptr(), positions.data_ptr<int64_t>(), cos_sin_cache.data_ptr(), (int)cos_sin_cache.size(1),
      slot_mapping.data_ptr<int64_t>(), (int)query.size(0), (int)query.size(1), (int)num_kv_heads, (int)query.size(2),
      (float)scale, block_tables.data_ptr<int>(), seq_lens.data_ptr<int>(), (int)block_tables.size(1), (int)block_size,
      (int)max_seq_len, query.stride(0), key.stride(0), value.stride(0), key_cache.stride(0), key_cache.stride(1),
      dtype_enum(query.scalar_type(), "rope_cache_paged_attention"), kv_enum(kv_cache_dtype), (float)kv_scale, v2 ? 1 : 0,
      cur_stream(query));
  if (rc == B200_ENOTSUP) return false;
  raise_if(rc, "rope_cache_paged_attention");
  return true;
}

// ---------------------------------------------------------------- cache (cache_kernels.cu:253-278, 339-389)
void reshape_and_cache(torch::Tensor& key, torch::Tensor& value, torch::Tensor& key_cache, torch::Tensor& value_cache,
                       torch::Tensor& slot_mapping, const std::string& kv_cache_dtype, const double kv_scale) {
  TORCH_CHECK(slot_mapping.scalar_type() == at::ScalarType::Long, "slot_mapping must be int64");
  const at::cuda::OptionalCUDAGuard device_guard(device_of(key));
  raise_if(b200_reshape_and_cache(key.data_ptr(), value.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(),
                                  slot_mapping.data_ptr<int64_t>(), slot_mapping.size(0), (int)key.size(1),
                                  (int)key.size(2), (int)key_cache.size(3), (int)key_cache.size(4), key.stride(0),
                                  value.stride(0), dtype_enum(key.scalar_type(), "reshape_and_cache"),
                                  kv_enum(kv_cache_dtype), (float)kv_scale, cur_stream(key)),
           "reshape_and_cache");
}

// ---------------------------------------------------------------- block maintenance (cache_kernels.cu:24-148, 280-314)
void swap_blocks(torch::Tensor& src, torch::Tensor& dst, const torch::Tensor& block_mapping) {
  int direction;
  if (src.device().is_cuda() && dst.device().is_cuda()) {
    TORCH_CHECK(src.device().index() == dst.device().index(), "src and dst must be on the same GPU");
    direction = 0;
  } else if (src.device().is_cuda() && dst.device().is_cpu()) {
    direction = 1;
  } else if (src.device().is_cpu() && dst.device().is_cuda()) {
    direction = 2;
  } else {
    TORCH_CHECK(false, "Invalid device combination");
  }
  TORCH_CHECK(block_mapping.device().is_cpu(), "block_mapping must be on CPU");
  TORCH_CHECK(block_mapping.scalar_type() == at::ScalarType::Long && block_mapping.dim() == 2 && block_mapping.size(1) == 2,
              "block_mapping must be int64 [num_pairs, 2]");
  const torch::Tensor bm = block_mapping.contiguous();
  const int64_t block_bytes = src.element_size() * src[0].numel();
  const torch::Tensor& on_gpu = src.device().is_cuda() ? src : dst;
  const at::cuda::OptionalCUDAGuard device_guard(device_of(on_gpu));
  raise_if(b200_swap_blocks(src.data_ptr(), dst.data_ptr(), bm.data_ptr<int64_t>(), bm.size(0), block_bytes, direction,
                            cur_stream(on_gpu)),
           "swap_blocks");
}

void copy_blocks(std::vector<torch::Tensor> const& key_caches, std::vector<torch::Tensor> const& value_caches,
                 const torch::Tensor& block_mapping) {
  const int num_layers = (int)key_caches.size();
  TORCH_CHECK(num_layers == (int)value_caches.size());
  if (num_layers == 0) return;
  TORCH_CHECK(key_caches[0].device().is_cuda());
  TORCH_CHECK(block_mapping.scalar_type() == at::ScalarType::Long, "block_mapping must be int64");
  std::vector<int64_t> ptrs(2 * num_layers);
  for (int i = 0; i < num_layers; ++i) {
    ptrs[i] = reinterpret_cast<int64_t>(key_caches[i].data_ptr());
    ptrs[num_layers + i] = reinterpret_cast<int64_t>(value_caches[i].data_ptr());
  }
  const at::cuda::OptionalCUDAGuard device_guard(device_of(key_caches[0]));
  torch::Tensor dptrs = torch::from_blob(ptrs.data(), {2 * num_layers}, torch::kInt64).to(key_caches[0].device());
  const torch::Tensor bm = block_mapping.contiguous();
  const int64_t block_bytes = key_caches[0].element_size() * key_caches[0][0].numel();
  raise_if(b200_copy_blocks(dptrs.data_ptr<int64_t>(), dptrs.data_ptr<int64_t>() + num_layers, bm.data_ptr<int64_t>(),
                            num_layers, bm.size(0), block_bytes, cur_stream(key_caches[0])),
           "copy_blocks");
}

void reshape_and_cache_flash(torch::Tensor& key, torch::Tensor& value, torch::Tensor& k_cache, torch::Tensor& v_cache,
                             torch::Tensor& slot_mapping, const std::string& kv_cache_dtype) {
  TORCH_CHECK(kv_cache_dtype == "auto", "Unsupported data type of kv cache: ", kv_cache_dtype);
  TORCH_CHECK(k_cache.stride(0) == v_cache.stride(0));
  TORCH_CHECK(slot_mapping.scalar_type() == at::ScalarType::Long, "slot_mapping must be int64");
  const at::cuda::OptionalCUDAGuard device_guard(device_of(key));
  raise_if(b200_reshape_and_cache_flash(key.data_ptr(), value.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(),
                                        slot_mapping.data_ptr<int64_t>(), key.size(0), (int)key.size(1), (int)key.size(2),
                                        (int)k_cache.size(1), k_cache.stride(0), key.stride(0), value.stride(0),
                                        dtype_enum(key.scalar_type(), "reshape_and_cache_flash"), B200_KV_AUTO,
                                        cur_stream(key)),
           "reshape_and_cache_flash");
}

void convert_fp8(torch::Tensor& dst_cache, torch::Tensor& src_cache, const double kv_scale,
                 const std::string& kv_cache_dtype) {
  TORCH_CHECK(src_cache.device().is_cuda(), "src must be on a GPU");
  TORCH_CHECK(dst_cache.device().is_cuda(), "dst must be on a GPU");
  TORCH_CHECK(src_cache.device().index() == dst_cache.device().index(), "src and dst must be on the same GPU");
  TORCH_CHECK(kv_cache_dtype == "auto" || kv_cache_dtype == "fp8" || kv_cache_dtype == "fp8_e4m3",
              "Unsupported data type: ", kv_cache_dtype);
  TORCH_CHECK(src_cache.is_contiguous() && dst_cache.is_contiguous() && src_cache.numel() == dst_cache.numel());
  const at::cuda::OptionalCUDAGuard device_guard(device_of(src_cache));
  const bool src_fp8 = src_cache.scalar_type() == at::ScalarType::Byte;
  const bool dst_fp8 = dst_cache.scalar_type() == at::ScalarType::Byte;
  TORCH_CHECK(src_fp8 != dst_fp8, "convert_fp8: exactly one of src/dst must be the uint8 (fp8) cache");
  const int dt = dtype_enum(src_fp8 ? dst_cache.scalar_type() : src_cache.scalar_type(), "convert_fp8");
  raise_if(b200_convert_fp8(dst_cache.data_ptr(), src_cache.data_ptr(), src_cache.numel(), dt, kv_enum(kv_cache_dtype),
                            (float)kv_scale, dst_fp8 ? 1 : 0, cur_stream(src_cache)),
           "convert_fp8");
}

// ---------------------------------------------------------------- quantizers (fp8/common.cu:129-165, int8_quant_kernels.cu:75-115)
void static_scaled_fp8_quant(torch::Tensor& out, torch::Tensor& input, torch::Tensor& scale) {
  TORCH_CHECK(input.is_contiguous() && out.is_contiguous());
  TORCH_CHECK(scale.scalar_type() == at::ScalarType::Float && scale.numel() == 1);
  const at::cuda::OptionalCUDAGuard device_guard(device_of(input));
  raise_if(b200_static_scaled_fp8_quant(out.data_ptr(), input.data_ptr(), scale.data_ptr<float>(), input.numel(),
                                        dtype_enum(input.scalar_type(), "static_scaled_fp8_quant"), cur_stream(input)),
           "static_scaled_fp8_quant");
}
void dynamic_scaled_fp8_quant(torch::Tensor& out, torch::Tensor& input, torch::Tensor& scale) {
  TORCH_CHECK(input.is_contiguous() && out.is_contiguous());
  TORCH_CHECK(scale.scalar_type() == at::ScalarType::Float && scale.numel() == 1);
  const at::cuda::OptionalCUDAGuard device_guard(device_of(input));
  raise_if(b200_dynamic_scaled_fp8_quant(out.data_ptr(), input.data_ptr(), scale.data_ptr<float>(), input.numel(),
                                         dtype_enum(input.scalar_type(), "dynamic_scaled_fp8_quant"),
                                         cur_stream(input)),
           "dynamic_scaled_fp8_quant");
}
void static_scaled_int8_quant(torch::Tensor& out, torch::Tensor const& input, torch::Tensor const& scale) {
  TORCH_CHECK(input.is_contiguous());
  TORCH_CHECK(out.is_contiguous());
  TORCH_CHECK(scale.numel() == 1);
  const at::cuda::OptionalCUDAGuard device_guard(device_of(input));
  const int hidden = (int)input.size(-1);
  raise_if(b200_static_scaled_int8_quant(out.data_ptr<int8_t>(), input.data_ptr(), scale.data_ptr<float>(),
                                         hidden ? input.numel() / hidden : 0, hidden,
                                         dtype_enum(input.scalar_type(), "static_scaled_int8_quant"),
                                         cur_stream(input)),
           "static_scaled_int8_quant");
}
void dynamic_scaled_int8_quant(torch::Tensor& out, torch::Tensor const& input, torch::Tensor& scales) {
  TORCH_CHECK(input.is_contiguous());
  TORCH_CHECK(out.is_contiguous());
  const at::cuda::OptionalCUDAGuard device_guard(device_of(input));
  const int hidden = (int)input.size(-1);
  raise_if(b200_dynamic_scaled_int8_quant(out.data_ptr<int8_t>(), input.data_ptr(), scales.data_ptr<float>(),
                                          hidden ? input.numel() / hidden : 0, hidden,
                                          dtype_enum(input.scalar_type(), "dynamic_scaled_int8_quant"),
                                          cur_stream(input)),
           "dynamic_scaled_int8_quant");
}

// ---------------------------------------------------------------- Marlin repack (gptq_marlin_repack.cu:276-348)
torch::Tensor gptq_marlin_repack(torch::Tensor& b_q_weight, torch::Tensor& perm, int64_t size_k, int64_t size_n,
                                 int64_t num_bits) {
  TORCH_CHECK(size_k % 16 == 0, "size_k = ", size_k, " is not divisible by tile_k_size = ", 16);
  TORCH_CHECK(size_n % 64 == 0, "size_n = ", size_n, " is not divisible by tile_n_size = ", 64);
  TORCH_CHECK(num_bits == 4 || num_bits == 8, "num_bits must be 4 or 8. Got = ", num_bits);
  const int pack_factor = 32 / (int)num_bits;
  TORCH_CHECK((size_k / pack_factor) == b_q_weight.size(0), "Shape mismatch: b_q_weight.size(0) = ",
              b_q_weight.size(0), ", size_k = ", size_k, ", pack_factor = ", pack_factor);
  TORCH_CHECK(b_q_weight.size(1) == size_n, "b_q_weight.size(1) = ", b_q_weight.size(1),
              " is not size_n = ", size_n);
  TORCH_CHECK(b_q_weight.device().is_cuda(), "b_q_weight is not on GPU");
  TORCH_CHECK(b_q_weight.is_contiguous(), "b_q_weight is not contiguous");
  TORCH_CHECK(b_q_weight.dtype() == at::kInt, "b_q_weight type is not kInt");
  TORCH_CHECK(perm.device().is_cuda(), "perm is not on GPU");
  TORCH_CHECK(perm.is_contiguous(), "perm is not contiguous");
  TORCH_CHECK(perm.dtype() == at::kInt, "perm type is not at::kInt");
  const at::cuda::OptionalCUDAGuard device_guard(device_of(b_q_weight));
  auto options = torch::TensorOptions().dtype(b_q_weight.dtype()).device(b_q_weight.device());
  torch::Tensor out = torch::empty({size_k / 16, size_n * 16 / pack_factor}, options);
  const bool has_perm = perm.size(0) != 0;
  if (has_perm) TORCH_CHECK(perm.size(0) == size_k, "perm.size(0) = ", perm.size(0), " is not size_k = ", size_k);
  raise_if(b200_gptq_marlin_repack(b_q_weight.data_ptr<int>(), has_perm ? perm.data_ptr<int>() : nullptr,
                                   out.data_ptr<int>(), size_k, size_n, (int)num_bits, cur_stream(b_q_weight)),
           "gptq_marlin_repack");
  return out;
}

// ---------------------------------------------------------------- prefill attention over the paged cache
// context_attention_fwd (vllm/attention/ops/prefix_prefill.py:674-799, Triton in the reference): same argument order
void context_attention_fwd(torch::Tensor& q, torch::Tensor& k, torch::Tensor& v, torch::Tensor& o, torch::Tensor& k_cache,
                           torch::Tensor& v_cache, torch::Tensor& b_loc, torch::Tensor& b_start_loc,
                           torch::Tensor& b_seq_len, torch::Tensor& b_ctx_len, int64_t max_input_len,
                           const c10::optional<torch::Tensor>& alibi_slopes, int64_t sliding_window) {
  TORCH_CHECK(q.dim() == 3 && k.dim() == 3 && v.dim() == 3 && o.dim() == 3, "q / k / v / o must be [tokens, heads, head_size]");
  TORCH_CHECK(q.size(2) == k.size(2) && k.size(2) == v.size(2), "head sizes of q, k, v differ");   // prefix_prefill.py:691-692
  TORCH_CHECK(k_cache.dim() == 5 && v_cache.dim() == 4, "k_cache must be 5-D [NB, Hkv, D/x, BS, x], v_cache 4-D [NB, Hkv, D, BS]");
  TORCH_CHECK(q.stride(2) == 1 && k.stride(2) == 1 && v.stride(2) == 1 && o.stride(2) == 1, "head dimension must be contiguous");
  TORCH_CHECK(k_cache.scalar_type() == q.scalar_type() && v_cache.scalar_type() == q.scalar_type(),
              "context_attention_fwd: the cache must hold the query dtype");
  TORCH_CHECK(k_cache.stride(4) == 1 && k_cache.stride(3) == k_cache.size(4) && k_cache.stride(2) == k_cache.size(3) * k_cache.size(4) &&
              v_cache.stride(3) == 1 && v_cache.stride(2) == v_cache.size(3), "cache blocks must be contiguous per kv head");
  const at::cuda::OptionalCUDAGuard device_guard(device_of(q));
  auto i32 = [&](torch::Tensor& t) { return t.scalar_type() == at::ScalarType::Int ? t : t.to(at::ScalarType::Int); };
  torch::Tensor bt = i32(b_loc), st = i32(b_start_loc).contiguous(), sl = i32(b_seq_len).contiguous(), cl = i32(b_ctx_len).contiguous();
  TORCH_CHECK(bt.dim() == 2 && bt.stride(1) == 1, "b_loc must be [batch, max_blocks] with contiguous rows");
  const int batch = (int)sl.size(0);
  TORCH_CHECK(st.size(0) >= batch && cl.size(0) == batch && bt.size(0) >= batch, "b_start_loc / b_ctx_len / b_loc do not match b_seq_len");
  const int64_t strides[8] = {q.stride(0), q.stride(1), k.stride(0), k.stride(1), v.stride(0), v.stride(1), o.stride(0), o.stride(1)};
  const float* alibi = nullptr;
  if (alibi_slopes.has_value()) {
    TORCH_CHECK(alibi_slopes->scalar_type() == at::ScalarType::Float && alibi_slopes->is_contiguous() && alibi_slopes->numel() == q.size(1),
                "alibi_slopes must be float32 [num_heads]");
    alibi = alibi_slopes->data_ptr<float>();
  }
  raise_if(b200_context_attention_fwd(o.data_ptr(), q.data_ptr(), k.data_ptr(), v.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(),
                                      bt.data_ptr<int>(), bt.stride(0), st.data_ptr<int>(), sl.data_ptr<int>(), cl.data_ptr<int>(), batch,
                                      (int)max_input_len, (int)q.size(1), (int)k.size(1), (int)q.size(2), (int)v_cache.size(3),
                                      (int)k_cache.size(4), strides, k_cache.stride(0), k_cache.stride(1), v_cache.stride(0),
                                      v_cache.stride(1), alibi, (int)sliding_window,
                                      dtype_enum(q.scalar_type(), "context_attention_fwd"), cur_stream(q)),
           "context_attention_fwd");
}

// ---------------------------------------------------------------- Marlin GEMMs
// gptq_marlin.cu:1735-1868
torch::Tensor gptq_marlin_gemm(torch::Tensor& a, torch::Tensor& b_q_weight, torch::Tensor& b_scales,
                               torch::Tensor& g_idx, torch::Tensor& perm, torch::Tensor& workspace, int64_t num_bits,
                               int64_t size_m, int64_t size_n, int64_t size_k, bool is_k_full) {
  TORCH_CHECK(num_bits == 4 || num_bits == 8, "num_bits must be 4 or 8. Got = ", num_bits);
  const int pack_factor = 32 / (int)num_bits;
  TORCH_CHECK(a.size(0) == size_m, "Shape mismatch: a.size(0) = ", a.size(0), ", size_m = ", size_m);
  TORCH_CHECK(a.size(1) == size_k, "Shape mismatch: a.size(1) = ", a.size(1), ", size_k = ", size_k);
  TORCH_CHECK(size_k % 16 == 0, "size_k = ", size_k, " is not divisible by tile_size = ", 16);
  TORCH_CHECK((size_k / 16) == b_q_weight.size(0), "Shape mismatch: b_q_weight.size(0) = ", b_q_weight.size(0),
              ", size_k = ", size_k, ", tile_size = ", 16);
  TORCH_CHECK(b_q_weight.size(1) % 16 == 0, "b_q_weight.size(1) = ", b_q_weight.size(1),
              " is not divisible by tile_size = ", 16);
  const int64_t actual_size_n = (b_q_weight.size(1) / 16) * pack_factor;
  TORCH_CHECK(size_n == actual_size_n, "size_n = ", size_n, ", actual_size_n = ", actual_size_n);
  TORCH_CHECK(a.device().is_cuda(), "A is not on GPU");
  TORCH_CHECK(a.is_contiguous(), "A is not contiguous");
  TORCH_CHECK(b_q_weight.device().is_cuda(), "b_q_weight is not on GPU");
  TORCH_CHECK(b_q_weight.is_contiguous(), "b_q_weight is not contiguous");
  TORCH_CHECK(b_scales.device().is_cuda(), "b_scales is not on GPU");
  TORCH_CHECK(b_scales.is_contiguous(), "b_scales is not contiguous");
  TORCH_CHECK(g_idx.device().is_cuda(), "g_idx is not on GPU");
  TORCH_CHECK(g_idx.is_contiguous(), "g_idx is not contiguous");
  TORCH_CHECK(perm.device().is_cuda(), "perm is not on GPU");
  TORCH_CHECK(perm.is_contiguous(), "perm is not contiguous");
  TORCH_CHECK((g_idx.size(0) == 0 && perm.size(0) == 0) || (g_idx.size(0) == size_k && perm.size(0) == size_k),
              "Unexpected g_idx.size(0) = ", g_idx.size(0), " and perm.size(0) = ", perm.size(0),
              ", where size_k = ", size_k);
  TORCH_CHECK(b_scales.dim() == 2, "b_scales rank = ", b_scales.dim(), " is not 2");
  TORCH_CHECK(b_scales.size(1) == size_n, "b_scales dim 1 = ", b_scales.size(1), " is not size_n = ", size_n);
  TORCH_CHECK(b_scales.scalar_type() == a.scalar_type(), "b_scales must have the dtype of a");
  TORCH_CHECK(size_n % 64 == 0, "size_n = ", size_n, ", is not divisible by min_thread_n = ", 64);
  const int64_t min_workspace_size = (size_n / 64) * 16;
  TORCH_CHECK(workspace.numel() >= min_workspace_size, "workspace.numel = ", workspace.numel(),
              " is below min_workspace_size = ", min_workspace_size);
  TORCH_CHECK(a.scalar_type() == at::ScalarType::Half || a.scalar_type() == at::ScalarType::BFloat16,
              "gpt_marlin_gemm only supports bfloat16 and float16");
  const at::cuda::OptionalCUDAGuard device_guard(device_of(a));
  torch::Tensor c = torch::empty({size_m, size_n}, a.options());
  const bool act_order = g_idx.size(0) != 0;
  raise_if(b200_gptq_marlin_gemm(c.data_ptr(), a.data_ptr(), b_q_weight.data_ptr<int>(), b_scales.data_ptr(),
                                 act_order ? g_idx.data_ptr<int>() : nullptr,
                                 act_order ? perm.data_ptr<int>() : nullptr, workspace.data_ptr<int>(),
                                 workspace.numel(), (int)num_bits, size_m, size_n, size_k, a.stride(0),
                                 (int)b_scales.size(0), is_k_full ? 1 : 0,
                                 dtype_enum(a.scalar_type(), "gptq_marlin_gemm"), cur_stream(a)),
           "gptq_marlin_gemm");
  return c;
}

// gate_up_proj GEMM with SiluAndMul in its epilogue (no counterpart op in the reference: it launches gptq_marlin_gemm and
// silu_and_mul, llama.py:74-90); same argument checks as gptq_marlin_gemm, output [size_m, size_n / 2]
torch::Tensor gptq_marlin_gemm_silu_and_mul(torch::Tensor& a, torch::Tensor& b_q_weight, torch::Tensor& b_scales,
                                            torch::Tensor& workspace, int64_t num_bits, int64_t size_m, int64_t size_n,
                                            int64_t size_k) {
  TORCH_CHECK(num_bits == 4 || num_bits == 8, "num_bits must be 4 or 8. Got = ", num_bits);
  const int pack_factor = 32 / (int)num_bits;
  TORCH_CHECK(a.dim() == 2 && a.size(0) == size_m && a.size(1) == size_k, "Shape mismatch: a is not [size_m, size_k]");
  TORCH_CHECK(size_k % 16 == 0 && (size_k / 16) == b_q_weight.size(0), "Shape mismatch: b_q_weight.size(0) = ",
              b_q_weight.size(0), ", size_k = ", size_k);
  TORCH_CHECK(b_q_weight.size(1) % 16 == 0 && size_n == (b_q_weight.size(1) / 16) * pack_factor, "size_n = ", size_n,
              " does not match b_q_weight");
  TORCH_CHECK(a.device().is_cuda() && a.stride(1) == 1, "A is not on GPU / not row-major");
  TORCH_CHECK(b_q_weight.device().is_cuda() && b_q_weight.is_contiguous(), "b_q_weight is not on GPU / not contiguous");
  TORCH_CHECK(b_scales.device().is_cuda() && b_scales.is_contiguous(), "b_scales is not on GPU / not contiguous");
  TORCH_CHECK(b_scales.dim() == 2 && b_scales.size(1) == size_n, "b_scales dim 1 = ", b_scales.size(1), " is not size_n = ", size_n);
  TORCH_CHECK(b_scales.scalar_type() == a.scalar_type(), "b_scales must have the dtype of a");
  TORCH_CHECK(size_n % 128 == 0, "size_n / 2 = ", size_n / 2, " is not divisible by 64");
  TORCH_CHECK(workspace.numel() >= (size_n / 64) * 16, "workspace.numel = ", workspace.numel(), " is below min_workspace_size");
  TORCH_CHECK(a.scalar_type() == at::ScalarType::Half || a.scalar_type() == at::ScalarType::BFloat16,
              "gptq_marlin_gemm_silu_and_mul only supports bfloat16 and float16");
  const at::cuda::OptionalCUDAGuard device_guard(device_of(a));
  torch::Tensor out = torch::empty({size_m, size_n / 2}, a.options());
  raise_if(b200_gptq_marlin_gemm_silu_and_mul(out.data_ptr(), a.data_ptr(), b_q_weight.data_ptr<int>(), b_scales.data_ptr(),
                                              workspace.data_ptr<int>(), workspace.numel(), (int)num_bits, size_m, size_n,
                                              size_k, a.stride(0), (int)b_scales.size(0),
                                              dtype_enum(a.scalar_type(), "gptq_marlin_gemm_silu_and_mul"), cur_stream(a)),
           "gptq_marlin_gemm_silu_and_mul");
  return out;
}

// marlin/dense/marlin_cuda_kernel.cu:1044-1136 (W4A16, fp16, groups -1/128): same layout family, same kernel.
torch::Tensor marlin_gemm(torch::Tensor& a, torch::Tensor& b_q_weight, torch::Tensor& b_scales,
                          torch::Tensor& workspace, int64_t size_m, int64_t size_n, int64_t size_k) {
  TORCH_CHECK(a.size(0) == size_m, "Shape mismatch: a.size(0) = ", a.size(0), ", size_m = ", size_m);
  TORCH_CHECK(a.size(1) == size_k, "Shape mismatch: a.size(1) = ", a.size(1), ", size_k = ", size_k);
  TORCH_CHECK(size_k % 16 == 0, "size_k = ", size_k, " is not divisible by tile_size = ", 16);
  TORCH_CHECK((size_k / 16) == b_q_weight.size(0), "Shape mismatch: b_q_weight.size(0) = ", b_q_weight.size(0),
              ", size_k = ", size_k, ", tile_size = ", 16);
  TORCH_CHECK(b_q_weight.size(1) % 16 == 0, "b_q_weight.size(1) = ", b_q_weight.size(1),
              " is not divisible by tile_size = ", 16);
  const int64_t actual_size_n = (b_q_weight.size(1) / 16) * 8;
  TORCH_CHECK(size_n == actual_size_n, "size_n = ", size_n, ", actual_size_n = ", actual_size_n);
  TORCH_CHECK(a.device().is_cuda() && a.is_contiguous(), "A is not on GPU / not contiguous");
  TORCH_CHECK(b_q_weight.device().is_cuda() && b_q_weight.is_contiguous(), "b_q_weight is not on GPU / not contiguous");
  TORCH_CHECK(b_scales.device().is_cuda() && b_scales.is_contiguous(), "b_scales is not on GPU / not contiguous");
  TORCH_CHECK(a.scalar_type() == at::ScalarType::Half, "marlin_gemm only supports float16");
  TORCH_CHECK(b_scales.dim() == 2 && b_scales.size(1) == size_n, "b_scales dim 1 = ", b_scales.size(1),
              " is not size_n = ", size_n);
  const int64_t groups = b_scales.size(0);
  TORCH_CHECK(groups == 1 || (size_k % groups == 0 && size_k / groups == 128),
              "Unexpected groupsize = ", groups ? size_k / groups : 0);  // marlin_cuda_kernel.cu:1112-1116
  TORCH_CHECK(size_n % 64 == 0, "size_n = ", size_n, ", is not divisible by min_thread_n = ", 64);
  TORCH_CHECK(workspace.numel() >= (size_n / 64) * 16, "workspace.numel = ", workspace.numel(),
              " is below min_workspace_size = ", (size_n / 64) * 16);
  const at::cuda::OptionalCUDAGuard device_guard(device_of(a));
  torch::Tensor c = torch::empty({size_m, size_n}, a.options());
  raise_if(b200_gptq_marlin_gemm(c.data_ptr(), a.data_ptr(), b_q_weight.data_ptr<int>(), b_scales.data_ptr(), nullptr,
                                 nullptr, workspace.data_ptr<int>(), workspace.numel(), 4, size_m, size_n, size_k,
                                 a.stride(0), (int)groups, 1, B200_F16, cur_stream(a)),
           "marlin_gemm");
  return c;
}

// marlin/sparse/marlin_24_cuda_kernel.cu:1017-1125
torch::Tensor gptq_marlin_24_gemm(torch::Tensor& a, torch::Tensor& b_q_weight, torch::Tensor& b_meta,
                                  torch::Tensor& b_scales, torch::Tensor& workspace, int64_t num_bits, int64_t size_m,
                                  int64_t size_n, int64_t size_k) {
  TORCH_CHECK(num_bits == 4 || num_bits == 8, "num_bits must be 4 or 8. Got = ", num_bits);
  const int pack_factor = 32 / (int)num_bits;
  TORCH_CHECK(a.size(0) == size_m, "Shape mismatch: a.size(0) = ", a.size(0), ", size_m = ", size_m);
  TORCH_CHECK(a.size(1) == size_k, "Shape mismatch: a.size(1) = ", a.size(1), ", size_k = ", size_k);
  TORCH_CHECK(size_k % 16 == 0, "size_k = ", size_k, " is not divisible by tile_size = ", 16);
  TORCH_CHECK((size_k / 16 / 2) == b_q_weight.size(0), "Shape mismatch: b_q_weight.size(0) = ", b_q_weight.size(0),
              ", size_k = ", size_k, ", tile_size = ", 16);
  TORCH_CHECK(b_meta.size(0) == size_k / 8 / 2 / 2, "Shape mismatch: b_meta.size(0) = ", b_meta.size(0),
              ", size_k = ", size_k);
  TORCH_CHECK(b_meta.size(1) == size_n * 2, "Shape mismatch: b_meta.size(1) = ", b_meta.size(1),
              ", size_n = ", size_n);
  TORCH_CHECK(b_q_weight.size(1) % 16 == 0, "b_q_weight.size(1) = ", b_q_weight.size(1),
              " is not divisible by tile_size = ", 16);
  const int64_t actual_size_n = (b_q_weight.size(1) / 16) * pack_factor;
  TORCH_CHECK(size_n == actual_size_n, "size_n = ", size_n, ", actual_size_n = ", actual_size_n);
  TORCH_CHECK(a.device().is_cuda() && a.is_contiguous(), "A is not on GPU / not contiguous");
  TORCH_CHECK(b_q_weight.device().is_cuda() && b_q_weight.is_contiguous(), "b_q_weight is not on GPU / not contiguous");
  TORCH_CHECK(b_scales.device().is_cuda() && b_scales.is_contiguous(), "b_scales is not on GPU / not contiguous");
  TORCH_CHECK(b_meta.device().is_cuda() && b_meta.is_contiguous(), "b_meta is not on GPU / not contiguous");
  TORCH_CHECK(b_meta.scalar_type() == at::ScalarType::Short, "b_meta must be int16");
  TORCH_CHECK(a.scalar_type() == at::ScalarType::Half, "gptq_marlin_24_gemm only supports float16");
  TORCH_CHECK(b_scales.dim() == 2 && b_scales.size(1) == size_n, "b_scales.size(1) = ", b_scales.size(1),
              ", size_n = ", size_n);
  const int64_t groups = b_scales.size(0);
  TORCH_CHECK(groups == 1 || (size_k % groups == 0 && size_k / groups == 128),
              "Unexpected groupsize = ", groups ? size_k / groups : 0);  // marlin_24_cuda_kernel.cu:1098-1104
  TORCH_CHECK(size_n % 128 == 0, "size_n = ", size_n, ", is not divisible by min_thread_n = ", 128);
  TORCH_CHECK(workspace.numel() >= (size_n / 128) * 64, "workspace.numel = ", workspace.numel(),
              " is below min_workspace_size = ", (size_n / 128) * 64);
  const at::cuda::OptionalCUDAGuard device_guard(device_of(a));
  torch::Tensor c = torch::empty({size_m, size_n}, a.options());
  raise_if(b200_gptq_marlin_24_gemm(c.data_ptr(), a.data_ptr(), b_q_weight.data_ptr<int>(),
                                    b_meta.data_ptr<int16_t>(), b_scales.data_ptr(), workspace.data_ptr<int>(),
                                    workspace.numel(), (int)num_bits, size_m, size_n, size_k, a.stride(0), (int)groups,
                                    B200_F16, cur_stream(a)),
           "gptq_marlin_24_gemm");
  return c;
}

// ---------------------------------------------------------------- scaled_mm (scaled_mm_entry.cu:32-100)
bool cutlass_scaled_mm_supports_fp8(int64_t cuda_device_capability) {
  return b200_scaled_mm_supports_fp8((int)cuda_device_capability) != 0;
}

void cutlass_scaled_mm(torch::Tensor& c, torch::Tensor const& a, torch::Tensor const& b, torch::Tensor const& a_scales,
                       torch::Tensor const& b_scales, std::optional<torch::Tensor> const& bias) {
  TORCH_CHECK(a.dim() == 2 && b.dim() == 2 && c.dim() == 2);
  TORCH_CHECK(c.size(0) == a.size(0) && a.size(1) == b.size(0) && b.size(1) == c.size(1));
  TORCH_CHECK(a_scales.numel() == 1 || a_scales.numel() == a.size(0));
  TORCH_CHECK(b_scales.numel() == 1 || b_scales.numel() == b.size(1));
  TORCH_CHECK(a.stride(1) == 1 && c.stride(1) == 1);             // Row-major
  TORCH_CHECK(b.stride(0) == 1);                                 // Column-major
  TORCH_CHECK(c.stride(0) % 16 == 0 && b.stride(1) % 16 == 0);   // 16 Byte Alignment
  TORCH_CHECK(a_scales.is_contiguous() && b_scales.is_contiguous());
  TORCH_CHECK(a_scales.scalar_type() == at::ScalarType::Float && b_scales.scalar_type() == at::ScalarType::Float);
  if (bias) {
    TORCH_CHECK(bias->numel() == b.size(1) && bias->is_contiguous() && bias->dim() == 1);
    TORCH_CHECK(bias->scalar_type() == c.scalar_type(), "bias must have the dtype of out");
  }
  int q8;
  if (a.scalar_type() == at::ScalarType::Char) {
    TORCH_CHECK(b.scalar_type() == at::ScalarType::Char);
    q8 = B200_Q8_INT8;
  } else {
    TORCH_CHECK(a.scalar_type() == at::ScalarType::Float8_e4m3fn);
    TORCH_CHECK(b.scalar_type() == at::ScalarType::Float8_e4m3fn);
    q8 = B200_Q8_FP8_E4M3;
  }
  TORCH_CHECK(c.scalar_type() == at::ScalarType::BFloat16 || c.scalar_type() == at::ScalarType::Half);
  const at::cuda::OptionalCUDAGuard device_guard(device_of(a));
  raise_if(b200_scaled_mm(c.data_ptr(), a.data_ptr(), b.data_ptr(), a_scales.data_ptr<float>(),
                          b_scales.data_ptr<float>(), bias ? bias->data_ptr() : nullptr, a.size(0), b.size(1),
                          a.size(1), a.stride(0), b.stride(1), c.stride(0), (int)a_scales.numel(),
                          (int)b_scales.numel(), q8, dtype_enum(c.scalar_type(), "cutlass_scaled_mm"), cur_stream(a)),
           "cutlass_scaled_mm");
}

// ---------------------------------------------------------------- glue (layernorm/activation/pos_encoding kernels)
void rms_norm(torch::Tensor& out, torch::Tensor& input, torch::Tensor& weight, double epsilon) {
  TORCH_CHECK(input.is_contiguous() && out.is_contiguous());
  const int hidden = (int)input.size(-1);
  const at::cuda::OptionalCUDAGuard device_guard(device_of(input));
  raise_if(b200_rms_norm(out.data_ptr(), input.data_ptr(), weight.data_ptr(), input.numel() / hidden, hidden,
                         (float)epsilon, dtype_enum(input.scalar_type(), "rms_norm"), cur_stream(input)),
           "rms_norm");
}
void fused_add_rms_norm(torch::Tensor& input, torch::Tensor& residual, torch::Tensor& weight, double epsilon) {
  TORCH_CHECK(input.is_contiguous() && residual.is_contiguous());
  const int hidden = (int)input.size(-1);
  const at::cuda::OptionalCUDAGuard device_guard(device_of(input));
  raise_if(b200_fused_add_rms_norm(input.data_ptr(), residual.data_ptr(), weight.data_ptr(), input.numel() / hidden,
                                   hidden, (float)epsilon, dtype_enum(input.scalar_type(), "fused_add_rms_norm"),
                                   cur_stream(input)),
           "fused_add_rms_norm");
}
void greedy_sample(torch::Tensor& out, torch::Tensor& logits) {
  TORCH_CHECK(logits.dim() == 2 && logits.stride(1) == 1, "greedy_sample: logits must be [tokens, vocab] with unit inner stride");
  TORCH_CHECK(out.scalar_type() == at::ScalarType::Long && out.is_contiguous() && out.numel() == logits.size(0),
              "greedy_sample: out must be contiguous int64 [tokens]");
  const at::cuda::OptionalCUDAGuard device_guard(device_of(logits));
  raise_if(b200_greedy_sample(out.data_ptr<int64_t>(), logits.data_ptr(), logits.size(0), logits.size(1), logits.stride(0),
                              dtype_enum(logits.scalar_type(), "greedy_sample"), cur_stream(logits)),
           "greedy_sample");
}

void silu_and_mul(torch::Tensor& out, torch::Tensor& input) {
  TORCH_CHECK(input.is_contiguous() && out.is_contiguous());
  const int d = (int)input.size(-1) / 2;
  const at::cuda::OptionalCUDAGuard device_guard(device_of(input));
  raise_if(b200_silu_and_mul(out.data_ptr(), input.data_ptr(), d ? input.numel() / (2 * d) : 0, d,
                             dtype_enum(input.scalar_type(), "silu_and_mul"), cur_stream(input)),
           "silu_and_mul");
}
void rotary_embedding(torch::Tensor& positions, torch::Tensor& query, torch::Tensor& key, int64_t head_size,
                      torch::Tensor& cos_sin_cache, bool is_neox) {
  const int64_t num_tokens = query.numel() / query.size(-1);
  const int rot_dim = (int)cos_sin_cache.size(1);
  const int num_heads = (int)(query.size(-1) / head_size);
  const int num_kv_heads = (int)(key.size(-1) / head_size);
  TORCH_CHECK(positions.scalar_type() == at::ScalarType::Long);
  const at::cuda::OptionalCUDAGuard device_guard(device_of(query));
  raise_if(b200_rotary_embedding(positions.data_ptr<int64_t>(), query.data_ptr(), key.data_ptr(), (int)head_size,
                                 cos_sin_cache.data_ptr(), rot_dim, is_neox ? 1 : 0, num_tokens, num_heads,
                                 num_kv_heads, query.stride(-2), key.stride(-2),
                                 dtype_enum(query.scalar_type(), "rotary_embedding"), cur_stream(query)),
           "rotary_embedding");
}

// ---------------------------------------------------------------- cuda utils (cuda_utils_kernels.cu)
int64_t get_device_attribute(int64_t attribute, int64_t device_id) {
  int v = 0;
  raise_if(b200_get_device_attribute((int)attribute, (int)device_id, &v), "get_device_attribute");
  return v;
}
int64_t get_max_shared_memory_per_block_device_attribute(int64_t device_id) {
  int v = 0;
  raise_if(b200_get_max_shared_memory_per_block_device_attribute((int)device_id, &v),
           "get_max_shared_memory_per_block_device_attribute");
  return v;
}

}  // namespace

// Schema strings: identical to csrc/torch_bindings.cpp:24-48, 146-150, 174-182, 194-204, 221-243; ops the reference
// defines from function pointers (marlin_gemm, gptq_marlin_24_gemm, gptq_marlin_gemm, gptq_marlin_repack,
// cutlass_scaled_mm_supports_fp8, cuda utils) are defined the same way so the inferred schema is the same.
TORCH_LIBRARY(_b200_C, ops) {
  ops.def(
      "paged_attention_v1("
      "    Tensor! out, Tensor query, Tensor key_cache,"
      "    Tensor value_cache, int num_kv_heads, float scale,"
      "    Tensor block_tables, Tensor seq_lens, int block_size,"
      "    int max_seq_len, Tensor? alibi_slopes,"
      "    str kv_cache_dtype, float kv_scale, int tp_rank,"
      "    int blocksparse_local_blocks,"
      "    int blocksparse_vert_stride, int blocksparse_block_size,"
      "    int blocksparse_head_sliding_step) -> ()");
  ops.impl("paged_attention_v1", torch::kCUDA, &paged_attention_v1);
  ops.def(
      "paged_attention_v2("
      "    Tensor! out, Tensor exp_sums, Tensor max_logits,"
      "    Tensor tmp_out, Tensor query, Tensor key_cache,"
      "    Tensor value_cache, int num_kv_heads, float scale,"
      "    Tensor block_tables, Tensor seq_lens, int block_size,"
      "    int max_seq_len, Tensor? alibi_slopes,"
      "    str kv_cache_dtype, float kv_scale, int tp_rank,"
      "    int blocksparse_local_blocks,"
      "    int blocksparse_vert_stride, int blocksparse_block_size,"
      "    int blocksparse_head_sliding_step) -> ()");
  ops.impl("paged_attention_v2", torch::kCUDA, &paged_attention_v2);

  ops.def(
      "rope_cache_paged_attention(Tensor! out, Tensor? exp_sums, Tensor? max_logits, Tensor? tmp_out, Tensor query,"
      " Tensor key, Tensor value, Tensor! key_cache, Tensor! value_cache, Tensor positions, Tensor cos_sin_cache,"
      " Tensor slot_mapping, int num_kv_heads, float scale, Tensor block_tables, Tensor seq_lens, int block_size,"
      " int max_seq_len, str kv_cache_dtype, float kv_scale) -> bool");
  ops.impl("rope_cache_paged_attention", torch::kCUDA, &rope_cache_paged_attention);
  ops.def("rms_norm_quant(Tensor! out_q, Tensor! scales, Tensor input, Tensor!? residual, Tensor weight, float epsilon, str kind) -> ()");
  ops.impl("rms_norm_quant", torch::kCUDA, &rms_norm_quant);
  ops.def("silu_and_mul_quant(Tensor! out_q, Tensor! scales, Tensor input, str kind) -> ()");
  ops.impl("silu_and_mul_quant", torch::kCUDA, &silu_and_mul_quant);
  ops.def("greedy_sample(Tensor! out, Tensor logits) -> ()");
  ops.impl("greedy_sample", torch::kCUDA, &greedy_sample);
  ops.def("silu_and_mul(Tensor! out, Tensor input) -> ()");
  ops.impl("silu_and_mul", torch::kCUDA, &silu_and_mul);
  ops.def("rms_norm(Tensor! out, Tensor input, Tensor weight, float epsilon) -> ()");
  ops.impl("rms_norm", torch::kCUDA, &rms_norm);
  ops.def("fused_add_rms_norm(Tensor! input, Tensor! residual, Tensor weight, float epsilon) -> ()");
  ops.impl("fused_add_rms_norm", torch::kCUDA, &fused_add_rms_norm);
  ops.def(
      "rotary_embedding(Tensor positions, Tensor! query,"
      "                 Tensor! key, int head_size,"
      "                 Tensor cos_sin_cache, bool is_neox) -> ()");
  ops.impl("rotary_embedding", torch::kCUDA, &rotary_embedding);

  ops.def("marlin_gemm", &marlin_gemm);
  ops.impl("marlin_gemm", torch::kCUDA, &marlin_gemm);
  ops.def("gptq_marlin_24_gemm", &gptq_marlin_24_gemm);
  ops.impl("gptq_marlin_24_gemm", torch::kCUDA, &gptq_marlin_24_gemm);
  ops.def("gptq_marlin_gemm", &gptq_marlin_gemm);
  ops.impl("gptq_marlin_gemm", torch::kCUDA, &gptq_marlin_gemm);
  ops.def(
      "context_attention_fwd(Tensor q, Tensor k, Tensor v, Tensor! o, Tensor k_cache, Tensor v_cache, Tensor b_loc,"
      "    Tensor b_start_loc, Tensor b_seq_len, Tensor b_ctx_len, int max_input_len, Tensor? alibi_slopes,"
      "    int sliding_window) -> ()");
  ops.impl("context_attention_fwd", torch::kCUDA, &context_attention_fwd);
  ops.def("gptq_marlin_gemm_silu_and_mul", &gptq_marlin_gemm_silu_and_mul);
  ops.impl("gptq_marlin_gemm_silu_and_mul", torch::kCUDA, &gptq_marlin_gemm_silu_and_mul);
  ops.def("gptq_marlin_repack", &gptq_marlin_repack);
  ops.impl("gptq_marlin_repack", torch::kCUDA, &gptq_marlin_repack);

  ops.def(
      "cutlass_scaled_mm(Tensor! out, Tensor a,"
      "                  Tensor b, Tensor a_scales,"
      "                  Tensor b_scales, Tensor? bias) -> ()");
  ops.impl("cutlass_scaled_mm", torch::kCUDA, &cutlass_scaled_mm);
  ops.def("cutlass_scaled_mm_supports_fp8", &cutlass_scaled_mm_supports_fp8);

  ops.def("static_scaled_fp8_quant(Tensor! out, Tensor input, Tensor scale) -> ()");
  ops.impl("static_scaled_fp8_quant", torch::kCUDA, &static_scaled_fp8_quant);
  ops.def("dynamic_scaled_fp8_quant(Tensor! out, Tensor input, Tensor! scale) -> ()");
  ops.impl("dynamic_scaled_fp8_quant", torch::kCUDA, &dynamic_scaled_fp8_quant);
  ops.def("static_scaled_int8_quant(Tensor! out, Tensor input, Tensor scale) -> ()");
  ops.impl("static_scaled_int8_quant", torch::kCUDA, &static_scaled_int8_quant);
  ops.def("dynamic_scaled_int8_quant(Tensor! out, Tensor input, Tensor! scale) -> ()");
  ops.impl("dynamic_scaled_int8_quant", torch::kCUDA, &dynamic_scaled_int8_quant);
}

TORCH_LIBRARY(_b200_C_cache_ops, cache_ops) {
  cache_ops.def(
      "reshape_and_cache(Tensor key, Tensor value,"
      "                  Tensor! key_cache, Tensor! value_cache,"
      "                  Tensor slot_mapping,"
      "                  str kv_cache_dtype,"
      "                  float kv_scale) -> ()");
  cache_ops.impl("reshape_and_cache", torch::kCUDA, &reshape_and_cache);
  cache_ops.def("swap_blocks(Tensor src, Tensor! dst, Tensor block_mapping) -> ()");
  cache_ops.impl("swap_blocks", torch::kCUDA, &swap_blocks);
  cache_ops.impl("swap_blocks", torch::kCPU, &swap_blocks);   // host -> device swaps dispatch on the CPU source
  cache_ops.def(
      "copy_blocks(Tensor[]! key_caches, Tensor[]! value_caches, Tensor "
      "block_mapping) -> ()");
  cache_ops.impl("copy_blocks", torch::kCUDA, &copy_blocks);
  cache_ops.def(
      "reshape_and_cache_flash(Tensor key, Tensor value,"
      "                        Tensor! key_cache,"
      "                        Tensor! value_cache,"
      "                        Tensor slot_mapping,"
      "                        str kv_cache_dtype) -> ()");
  cache_ops.impl("reshape_and_cache_flash", torch::kCUDA, &reshape_and_cache_flash);
  cache_ops.def(
      "convert_fp8(Tensor! dst_cache, Tensor src_cache, float scale, str "
      "kv_cache_dtype) -> ()");
  cache_ops.impl("convert_fp8", torch::kCUDA, &convert_fp8);
}

// ---------------------------------------------------------------- custom all-reduce (csrc/custom_all_reduce.cu:12-155)
// The reference's op surface over this build's kernels.  Differences that follow from the design: the
// peer-visible memory is ONE region per rank -- the `meta` tensor, which must be zero-initialised and hold
// meta_size() + max_size bytes exactly as the reference's Python allocates it -- and inputs are pushed into the peers'
// regions by the kernel itself, so nothing has to be registered: register_buffer / register_graph_buffers are accepted and
// ignored, get_graph_buffer_ipc_meta returns empty lists, all_reduce_unreg does not need its staging buffer.
using fptr_t = int64_t;
struct ArHandle {
  int64_t ctx;
  std::vector<void*> opened;
};

int64_t meta_size() {
  int64_t b = 0;
  raise_if(b200_ar_region_bytes(8, 8192 * 1024, &b), "meta_size");
  return b;   // covers every world <= 8 at the reference's default max_size of 8 MiB
}

fptr_t init_custom_ar(torch::Tensor& meta, torch::Tensor& rank_data, const std::vector<std::string>& handles,
                      const std::vector<int64_t>& offsets, int64_t rank, bool full_nvlink) {
  (void)rank_data; (void)full_nvlink;
  const int world = (int)offsets.size();
  if (world > 8) throw std::invalid_argument("world size > 8 is not supported");
  if (world % 2 != 0) throw std::invalid_argument("Odd num gpus is not supported for now");
  if (world != (int)handles.size()) throw std::invalid_argument("handles length should equal to offsets length");
  if (rank < 0 || rank >= world) throw std::invalid_argument("invalid rank passed in");
  const at::cuda::OptionalCUDAGuard device_guard(device_of(meta));
  // largest payload whose region fits into `meta`
  const int64_t have = meta.numel() * meta.element_size();
  int64_t lo = 4096, hi = have, need = 0;
  raise_if(b200_ar_region_bytes(world, lo, &need), "init_custom_ar");
  TORCH_CHECK(need <= have, "init_custom_ar: meta holds ", have, " bytes, the smallest region needs ", need);
  while (hi - lo > 4096) {
    const int64_t mid = (lo + hi) / 2 / 4096 * 4096;
    raise_if(b200_ar_region_bytes(world, mid, &need), "init_custom_ar");
    if (need <= have) lo = mid; else hi = mid;
  }
  auto* h = new ArHandle();
  std::vector<void*> regions(world);
  for (int r = 0; r < world; ++r) {
    if (r == rank) {
      regions[r] = meta.data_ptr();
    } else {
      TORCH_CHECK(handles[r].size() >= 64, "init_custom_ar: IPC handle of rank ", r, " is shorter than 64 bytes");
      void* base = nullptr;
      raise_if(b200_ar_open(handles[r].data(), &base), "init_custom_ar");
      h->opened.push_back(base);
      regions[r] = static_cast<char*>(base) + offsets[r];
    }
  }
  raise_if(b200_ar_init((int)rank, world, regions.data(), lo, &h->ctx), "init_custom_ar");
  return reinterpret_cast<fptr_t>(h);
}

bool should_custom_ar(torch::Tensor& inp, int64_t max_size, int64_t world_size, bool full_nvlink) {
  const int64_t bytes = inp.numel() * inp.element_size();
  if (bytes % 16 != 0) return false;
  const bool weak_contig = inp.is_contiguous() ||
      (int64_t)(inp.storage().nbytes() - inp.storage_offset() * inp.element_size()) == bytes;
  if (!weak_contig) return false;
  if (world_size == 2 || full_nvlink) return bytes <= max_size;
  return false;
}

void all_reduce_reg(fptr_t fa, torch::Tensor& inp, torch::Tensor& out) {
  const at::cuda::OptionalCUDAGuard device_guard(device_of(inp));
  TORCH_CHECK(inp.scalar_type() == out.scalar_type() && inp.numel() == out.numel());
  auto* h = reinterpret_cast<ArHandle*>(fa);
  raise_if(b200_ar_all_reduce(h->ctx, inp.data_ptr(), out.data_ptr(), inp.numel(), dtype_enum(inp.scalar_type(), "all_reduce_reg"),
                              cur_stream(inp)),
           "all_reduce_reg");
}

void all_reduce_unreg(fptr_t fa, torch::Tensor& inp, torch::Tensor& reg_buffer, torch::Tensor& out) {
  (void)reg_buffer;   // the kernel reads `inp` locally and pushes it to the peers: no staging copy
  all_reduce_reg(fa, inp, out);
}

void dispose(fptr_t fa) {
  auto* h = reinterpret_cast<ArHandle*>(fa);
  b200_ar_dispose(h->ctx);
  for (void* p : h->opened) b200_ar_close(p);
  delete h;
}

void register_buffer(fptr_t fa, torch::Tensor& t, const std::vector<std::string>& handles, const std::vector<int64_t>& offsets) {
  (void)fa; (void)t; (void)handles; (void)offsets;
}

std::tuple<torch::Tensor, std::vector<int64_t>> get_graph_buffer_ipc_meta(fptr_t fa) {
  (void)fa;
  return {torch::empty({0}, torch::TensorOptions().dtype(torch::kUInt8).device(torch::kCPU)), std::vector<int64_t>()};
}

void register_graph_buffers(fptr_t fa, const std::vector<std::string>& handles, const std::vector<std::vector<int64_t>>& offsets) {
  (void)fa; (void)handles; (void)offsets;
}

TORCH_LIBRARY(_b200_C_cuda_utils, cuda_utils) {
  // no tensor arguments -> no backend key to dispatch on: the def() catch-all kernel is the implementation
  cuda_utils.def("get_device_attribute", &get_device_attribute);
  cuda_utils.def("get_max_shared_memory_per_block_device_attribute", &get_max_shared_memory_per_block_device_attribute);
}

TORCH_LIBRARY(_b200_C_custom_ar, custom_ar) {
  // same names / schema strings as csrc/torch_bindings.cpp:262-294
  custom_ar.def("init_custom_ar", &init_custom_ar);
  custom_ar.impl("init_custom_ar", torch::kCUDA, &init_custom_ar);
  custom_ar.def("should_custom_ar", &should_custom_ar);
  custom_ar.impl("should_custom_ar", torch::kCUDA, &should_custom_ar);
  custom_ar.def("all_reduce_reg(int fa, Tensor inp, Tensor! out) -> ()");
  custom_ar.impl("all_reduce_reg", torch::kCUDA, &all_reduce_reg);
  custom_ar.def("all_reduce_unreg(int fa, Tensor inp, Tensor reg_buffer, Tensor! out) -> ()");
  custom_ar.impl("all_reduce_unreg", torch::kCUDA, &all_reduce_unreg);
  custom_ar.def("dispose", &dispose);
  custom_ar.def("meta_size", &meta_size);
  custom_ar.def("register_buffer", &register_buffer);
  custom_ar.impl("register_buffer", torch::kCUDA, &register_buffer);
  custom_ar.def("get_graph_buffer_ipc_meta", &get_graph_buffer_ipc_meta);
  custom_ar.def("register_graph_buffers", &register_graph_buffers);
}
