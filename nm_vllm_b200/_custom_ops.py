"""Mirror of vllm/_custom_ops.py (reference lines cited per function) for the hot-path operators.

Same names, positional order, defaults and return conventions; each function forwards to torch.ops._b200_C* which the
in-tree extension registers with the reference's schema strings.  `install()` rebinds these onto an importable
`vllm._custom_ops` so the reference's LinearMethod / PagedAttention classes call the sm_100a kernels unchanged.
"""
from typing import Optional, Tuple, Type

import torch

from . import load_extension


def _ops():
    load_extension()
    return torch.ops._b200_C


def _cache_ops():
    load_extension()
    return torch.ops._b200_C_cache_ops


# ---- paged attention: vllm/_custom_ops.py:74-129 -----------------------------------------------------------------
def paged_attention_v1(out: torch.Tensor, query: torch.Tensor, key_cache: torch.Tensor, value_cache: torch.Tensor,
                       num_kv_heads: int, scale: float, block_tables: torch.Tensor, seq_lens: torch.Tensor,
                       block_size: int, max_seq_len: int, alibi_slopes: Optional[torch.Tensor], kv_cache_dtype: str,
                       kv_scale: float, tp_rank: int = 0, blocksparse_local_blocks: int = 0,
                       blocksparse_vert_stride: int = 0, blocksparse_block_size: int = 64,
                       blocksparse_head_sliding_step: int = 0) -> None:
    _ops().paged_attention_v1(out, query, key_cache, value_cache, num_kv_heads, scale, block_tables, seq_lens,
                              block_size, max_seq_len, alibi_slopes, kv_cache_dtype, kv_scale, tp_rank,
                              blocksparse_local_blocks, blocksparse_vert_stride, blocksparse_block_size,
                              blocksparse_head_sliding_step)


def paged_attention_v2(out: torch.Tensor, exp_sum: torch.Tensor, max_logits: torch.Tensor, tmp_out: torch.Tensor,
                       query: torch.Tensor, key_cache: torch.Tensor, value_cache: torch.Tensor, num_kv_heads: int,
                       scale: float, block_tables: torch.Tensor, seq_lens: torch.Tensor, block_size: int,
                       max_seq_len: int, alibi_slopes: Optional[torch.Tensor], kv_cache_dtype: str, kv_scale: float,
                       tp_rank: int = 0, blocksparse_local_blocks: int = 0, blocksparse_vert_stride: int = 0,
                       blocksparse_block_size: int = 64, blocksparse_head_sliding_step: int = 0) -> None:
    _ops().paged_attention_v2(out, exp_sum, max_logits, tmp_out, query, key_cache, value_cache, num_kv_heads, scale,
                              block_tables, seq_lens, block_size, max_seq_len, alibi_slopes, kv_cache_dtype, kv_scale,
                              tp_rank, blocksparse_local_blocks, blocksparse_vert_stride, blocksparse_block_size,
                              blocksparse_head_sliding_step)


# ---- glue: vllm/_custom_ops.py:48-50 (silu_and_mul), :132-160 (rotary_embedding, rms_norm, fused_add_rms_norm) ----
def greedy_sample(logits: torch.Tensor) -> torch.Tensor:
    """argmax over the vocabulary (lowest index on ties) -- the temperature-0 branch of the reference's Sampler
    (vllm/model_executor/layers/sampler.py _greedy_sample)."""
    out = torch.empty((logits.shape[0],), dtype=torch.int64, device=logits.device)
    _ops().greedy_sample(out, logits)
    return out


def rope_cache_paged_attention(out: torch.Tensor, exp_sums: Optional[torch.Tensor], max_logits: Optional[torch.Tensor],
                               tmp_out: Optional[torch.Tensor], query: torch.Tensor, key: torch.Tensor,
                               value: torch.Tensor, key_cache: torch.Tensor, value_cache: torch.Tensor,
                               positions: torch.Tensor, cos_sin_cache: torch.Tensor, slot_mapping: torch.Tensor,
                               num_kv_heads: int, scale: float, block_tables: torch.Tensor, seq_lens: torch.Tensor,
                               block_size: int, max_seq_len: int, kv_cache_dtype: str, kv_scale: float) -> bool:
    """Decode-step fusion of rotary_embedding (neox) + reshape_and_cache + paged_attention_v1 (exp_sums is None) / v2 in
    one launch; `query` is left unrotated.  Returns False (nothing launched) when the shape is outside the fast path:
    the caller then runs the three reference ops."""
    return _ops().rope_cache_paged_attention(out, exp_sums, max_logits, tmp_out, query, key, value, key_cache,
                                             value_cache, positions, cos_sin_cache, slot_mapping, num_kv_heads, scale,
                                             block_tables, seq_lens, block_size, max_seq_len, kv_cache_dtype, kv_scale)


def context_attention_fwd(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, o: torch.Tensor, k_cache: torch.Tensor,
                          v_cache: torch.Tensor, b_loc: torch.Tensor, b_start_loc: torch.Tensor, b_seq_len: torch.Tensor,
                          b_ctx_len: torch.Tensor, max_input_len: int, alibi_slopes: Optional[torch.Tensor] = None,
                          sliding_window: Optional[int] = None) -> None:
    """Prefill attention over the paged cache, same signature as vllm/attention/ops/prefix_prefill.py:674-686 (Triton in
    the reference, one CUDA kernel here): query token i of sequence b attends to the b_ctx_len[b] cached tokens of its
    block-table row and to the call's new tokens 0..i (k, v); writes `o` in place."""
    _ops().context_attention_fwd(q, k, v, o, k_cache, v_cache, b_loc, b_start_loc, b_seq_len, b_ctx_len, max_input_len,
                                 alibi_slopes, 0 if sliding_window is None or sliding_window <= 0 else sliding_window)


_Q_DTYPE = {"fp8": torch.float8_e4m3fn, "fp8_dynamic": torch.float8_e4m3fn, "int8": torch.int8, "int8_token": torch.int8}
FUSED_DYNAMIC_FP8_MAX_TOKENS = 512   # kind "fp8_dynamic": all token CTAs must be co-resident (csrc/glue_quant.cu kDynMaxTokens)


def rms_norm_quant(input: torch.Tensor, weight: torch.Tensor, epsilon: float, kind: str, scale: Optional[torch.Tensor] = None,
                   residual: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """rms_norm (residual is None) or fused_add_rms_norm (residual updated in place) fused with the activation quantiser of the
    W8A8 linear that consumes it; returns (quantised activations, scales) like scaled_fp8_quant / scaled_int8_quant.
    kind "fp8" / "int8": `scale` is the static per-tensor scale; "int8_token": dynamic per-token scales are produced;
    "fp8_dynamic": the dynamic per-TENSOR scale max|x| / 448 of scaled_fp8_quant(x, None) is computed in the same launch
    (at most FUSED_DYNAMIC_FP8_MAX_TOKENS rows)."""
    out = torch.empty(input.shape, dtype=_Q_DTYPE[kind], device=input.device)
    if kind == "int8_token":
        scale = torch.empty((input.numel() // input.shape[-1], 1), dtype=torch.float32, device=input.device)
    elif kind == "fp8_dynamic":
        scale = torch.empty(1, dtype=torch.float32, device=input.device)
    _ops().rms_norm_quant(out, scale, input, residual, weight, epsilon, kind)
    return out, scale


def silu_and_mul_quant(x: torch.Tensor, kind: str, scale: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """silu_and_mul fused with the activation quantiser of the following W8A8 linear (see rms_norm_quant)."""
    d = x.shape[-1] // 2
    out = torch.empty(x.shape[:-1] + (d,), dtype=_Q_DTYPE[kind], device=x.device)
    if kind == "int8_token":
        scale = torch.empty((x.numel() // x.shape[-1], 1), dtype=torch.float32, device=x.device)
    elif kind == "fp8_dynamic":
        scale = torch.empty(1, dtype=torch.float32, device=x.device)
    _ops().silu_and_mul_quant(out, scale, x, kind)
    return out, scale


def silu_and_mul(out: torch.Tensor, x: torch.Tensor) -> None:
    _ops().silu_and_mul(out, x)


def rotary_embedding(positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor, head_size: int,
                     cos_sin_cache: torch.Tensor, is_neox: bool) -> None:
    _ops().rotary_embedding(positions, query, key, head_size, cos_sin_cache, is_neox)


def rms_norm(out: torch.Tensor, input: torch.Tensor, weight: torch.Tensor, epsilon: float) -> None:
    _ops().rms_norm(out, input, weight, epsilon)


def fused_add_rms_norm(input: torch.Tensor, residual: torch.Tensor, weight: torch.Tensor, epsilon: float) -> None:
    _ops().fused_add_rms_norm(input, residual, weight, epsilon)


# ---- marlin: vllm/_custom_ops.py:201-220 --------------------------------------------------------------------------
def marlin_gemm(a: torch.Tensor, b_q_weight: torch.Tensor, b_scales: torch.Tensor, workspace: torch.Tensor,
                size_m: int, size_n: int, size_k: int) -> torch.Tensor:
    return _ops().marlin_gemm(a, b_q_weight, b_scales, workspace, size_m, size_n, size_k)


def gptq_marlin_24_gemm(a: torch.Tensor, b_q_weight: torch.Tensor, b_meta: torch.Tensor, b_scales: torch.Tensor,
                        workspace: torch.Tensor, num_bits: int, size_m: int, size_n: int,
                        size_k: int) -> torch.Tensor:
    return _ops().gptq_marlin_24_gemm(a, b_q_weight, b_meta, b_scales, workspace, num_bits, size_m, size_n, size_k)


# ---- cutlass_scaled_mm: vllm/_custom_ops.py:223-238 -----------------------------------------------------------------
def cutlass_scaled_mm_supports_fp8(cuda_device_capability: int) -> bool:
    return _ops().cutlass_scaled_mm_supports_fp8(cuda_device_capability)


def cutlass_scaled_mm(a: torch.Tensor, b: torch.Tensor, scale_a: torch.Tensor, scale_b: torch.Tensor,
                      out_dtype: Type[torch.dtype], bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    assert (b.shape[0] % 16 == 0 and b.shape[1] % 16 == 0)
    assert (out_dtype is torch.bfloat16 or out_dtype is torch.float16)
    m = a.shape[0]
    n = b.shape[1]
    out = torch.empty((m, n), dtype=out_dtype, device=a.device)
    _ops().cutlass_scaled_mm(out, a, b, scale_a, scale_b, bias)
    return out


# ---- gptq_marlin: vllm/_custom_ops.py:257-280 ----------------------------------------------------------------------
def gptq_marlin_repack(b_q_weight: torch.Tensor, perm: torch.Tensor, size_k: int, size_n: int,
                       num_bits: int) -> torch.Tensor:
    return _ops().gptq_marlin_repack(b_q_weight, perm, size_k, size_n, num_bits)


def gptq_marlin_gemm(a: torch.Tensor, b_q_weight: torch.Tensor, b_scales: torch.Tensor, g_idx: torch.Tensor,
                     perm: torch.Tensor, workspace: torch.Tensor, num_bits: int, size_m: int, size_n: int,
                     size_k: int, is_k_full: bool) -> torch.Tensor:
    return _ops().gptq_marlin_gemm(a, b_q_weight, b_scales, g_idx, perm, workspace, num_bits, size_m, size_n, size_k,
                                   is_k_full)


def gptq_marlin_gemm_silu_and_mul(a: torch.Tensor, b_q_weight: torch.Tensor, b_scales: torch.Tensor,
                                  workspace: torch.Tensor, num_bits: int, size_m: int, size_n: int,
                                  size_k: int) -> torch.Tensor:
    """gate_up_proj + SiluAndMul in one launch (llama.py:74-90 issues gptq_marlin_gemm then silu_and_mul): returns
    [size_m, size_n // 2], bit-identical to the two-op chain; the stored weight / scale layouts are untouched."""
    return _ops().gptq_marlin_gemm_silu_and_mul(a, b_q_weight, b_scales, workspace, num_bits, size_m, size_n, size_k)


# ---- fp8 / int8 activation quantizers: vllm/_custom_ops.py:284-350 --------------------------------------------------
def scaled_fp8_quant(input: torch.Tensor, scale: Optional[torch.Tensor] = None,
                     batch_dim_padding: Optional[int] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    if batch_dim_padding:
        shape = (max(batch_dim_padding, input.shape[0]), *input.shape[1:])
        output = torch.empty(shape, device=input.device, dtype=torch.float8_e4m3fn)
    else:
        output = torch.empty_like(input, dtype=torch.float8_e4m3fn)
    if scale is None:
        scale = torch.zeros(1, device=input.device, dtype=torch.float32)
        _ops().dynamic_scaled_fp8_quant(output, input, scale)
    else:
        _ops().static_scaled_fp8_quant(output, input, scale)
    return output, scale


def scaled_int8_quant(input: torch.Tensor, scale: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    output = torch.empty_like(input, dtype=torch.int8)
    if scale is not None:
        _ops().static_scaled_int8_quant(output, input, scale)
        return output, scale
    input_scales = torch.empty((input.numel() // input.shape[-1], 1), device=input.device, dtype=torch.float32)
    _ops().dynamic_scaled_int8_quant(output, input, input_scales)
    return output, input_scales


# ---- cache ops: vllm/_custom_ops.py:370-412 ---------------------------------------------------------------------------
def reshape_and_cache(key: torch.Tensor, value: torch.Tensor, key_cache: torch.Tensor, value_cache: torch.Tensor,
                      slot_mapping: torch.Tensor, kv_cache_dtype: str, kv_scale: float) -> None:
    _cache_ops().reshape_and_cache(key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype, kv_scale)


def reshape_and_cache_flash(key: torch.Tensor, value: torch.Tensor, key_cache: torch.Tensor, value_cache: torch.Tensor,
                            slot_mapping: torch.Tensor, kv_cache_dtype: str) -> None:
    _cache_ops().reshape_and_cache_flash(key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype)


def copy_blocks(key_caches, value_caches, block_mapping: torch.Tensor) -> None:
    _cache_ops().copy_blocks(key_caches, value_caches, block_mapping)


def swap_blocks(src: torch.Tensor, dst: torch.Tensor, block_mapping: torch.Tensor) -> None:
    _cache_ops().swap_blocks(src, dst, block_mapping)


def convert_fp8(output: torch.Tensor, input: torch.Tensor, scale: float = 1.0, kv_dtype: str = "fp8") -> None:
    _cache_ops().convert_fp8(output, input, scale, kv_dtype)


def get_device_attribute(attribute: int, device: int) -> int:
    load_extension()
    return torch.ops._b200_C_cuda_utils.get_device_attribute(attribute, device)


def get_max_shared_memory_per_block_device_attribute(device: int) -> int:
    load_extension()
    return torch.ops._b200_C_cuda_utils.get_max_shared_memory_per_block_device_attribute(device)


# ---- custom all-reduce: vllm/_custom_ops.py:353-398 ---------------------------------------------------------------
def _ar_ops():
    load_extension()
    return torch.ops._b200_C_custom_ar


def init_custom_ar(meta: torch.Tensor, rank_data: torch.Tensor, handles, offsets, rank: int, full_nvlink: bool) -> int:
    return _ar_ops().init_custom_ar(meta, rank_data, handles, offsets, rank, full_nvlink)


def should_custom_ar(inp: torch.Tensor, max_size: int, world_size: int, full_nvlink: bool) -> bool:
    return _ar_ops().should_custom_ar(inp, max_size, world_size, full_nvlink)


def all_reduce_reg(fa: int, inp: torch.Tensor, out: torch.Tensor) -> None:
    _ar_ops().all_reduce_reg(fa, inp, out)


def all_reduce_unreg(fa: int, inp: torch.Tensor, reg_buffer: torch.Tensor, out: torch.Tensor) -> None:
    _ar_ops().all_reduce_unreg(fa, inp, reg_buffer, out)


def dispose(fa: int) -> None:
    _ar_ops().dispose(fa)


def meta_size() -> int:
    return _ar_ops().meta_size()


def register_buffer(fa: int, t: torch.Tensor, handles, offsets) -> None:
    return _ar_ops().register_buffer(fa, t, handles, offsets)


def get_graph_buffer_ipc_meta(fa: int):
    return _ar_ops().get_graph_buffer_ipc_meta(fa)


def register_graph_buffers(fa: int, handles, offsets) -> None:
    _ar_ops().register_graph_buffers(fa, handles, offsets)


_HOT_PATH_OPS = [
    "paged_attention_v1", "paged_attention_v2", "silu_and_mul", "rotary_embedding", "rms_norm", "fused_add_rms_norm",
    "marlin_gemm", "gptq_marlin_24_gemm", "cutlass_scaled_mm_supports_fp8", "cutlass_scaled_mm", "gptq_marlin_repack",
    "gptq_marlin_gemm", "scaled_fp8_quant", "scaled_int8_quant", "reshape_and_cache", "convert_fp8",
    "reshape_and_cache_flash", "copy_blocks", "swap_blocks",
    "get_device_attribute", "get_max_shared_memory_per_block_device_attribute",
    "init_custom_ar", "should_custom_ar", "all_reduce_reg", "all_reduce_unreg", "dispose", "meta_size", "register_buffer",
    "get_graph_buffer_ipc_meta", "register_graph_buffers",
]


def install(target=None) -> list:
    """Rebind the hot-path functions of `vllm._custom_ops` (or `target`) to this module's.  Returns the names bound."""
    if target is None:
        import vllm._custom_ops as target  # type: ignore
    g = globals()
    for name in _HOT_PATH_OPS:
        setattr(target, name, g[name])
    return list(_HOT_PATH_OPS)
