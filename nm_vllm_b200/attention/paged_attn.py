"""`PagedAttention` + `PagedAttentionMetadata`: mirror of vllm/attention/ops/paged_attn.py:13-239 -- cache shape and
views (:37-59), `write_to_paged_cache` (:62-79), `forward_decode` with the v1 / v2 heuristic and the
tmp_out / exp_sums / max_logits allocation (:82-182), block maintenance (:213-239).  `forward_prefix` (prefill over a
cached prefix, Triton in the reference) runs the CUDA prefill kernel of csrc/attention_prefill.cu.  New: `forward_decode_fused` runs the
step's rotary embedding, cache write and attention as one launch (csrc/attention_mma.cu) when the shape allows."""
from dataclasses import dataclass
from typing import List, Optional, Tuple

import torch

from .. import _custom_ops as ops

_PARTITION_SIZE = 512   # must equal the partition of paged_attention_v2 (csrc/attention_mma.cu kPartition)


@dataclass
class PagedAttentionMetadata:
    seq_lens_tensor: Optional[torch.Tensor]    # [batch] int32: tokens seen so far per sequence
    max_decode_seq_len: int                    # 0 for a prefill-only batch
    block_tables: Optional[torch.Tensor]       # [batch, max_blocks_per_seq] int32, zero padded under CUDA graphs


def _use_v1(max_seq_len: int, num_seqs: int, num_heads: int) -> Tuple[bool, int]:
    parts = (max_seq_len + _PARTITION_SIZE - 1) // _PARTITION_SIZE
    # one partition: no reduction needed; many (seq, head) pairs: enough parallelism without splitting; beyond 8192 tokens
    # the reference's v1 runs out of shared memory (paged_attn.py:113-121)
    return max_seq_len <= 8192 and (parts == 1 or num_seqs * num_heads > 512), parts


class PagedAttention:
    @staticmethod
    def get_supported_head_sizes() -> List[int]:
        return [64, 80, 96, 112, 128, 192, 256]

    @staticmethod
    def get_kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int, head_size: int) -> Tuple[int, ...]:
        return (2, num_blocks, block_size * num_kv_heads * head_size)

    @staticmethod
    def split_kv_cache(kv_cache: torch.Tensor, num_kv_heads: int, head_size: int) -> Tuple[torch.Tensor, torch.Tensor]:
        x = 16 // kv_cache.element_size()      # 8 for 16-bit caches, 16 for fp8 (uint8)
        nb = kv_cache.shape[1]
        return kv_cache[0].view(nb, num_kv_heads, head_size // x, -1, x), kv_cache[1].view(nb, num_kv_heads, head_size, -1)

    @staticmethod
    def write_to_paged_cache(key: torch.Tensor, value: torch.Tensor, key_cache: torch.Tensor, value_cache: torch.Tensor,
                             slot_mapping: torch.Tensor, kv_cache_dtype: str, kv_scale: float) -> None:
        ops.reshape_and_cache(key, value, key_cache, value_cache, slot_mapping.flatten(), kv_cache_dtype, kv_scale)

    @staticmethod
    def forward_decode(query: torch.Tensor, key_cache: torch.Tensor, value_cache: torch.Tensor, block_tables: torch.Tensor,
                       seq_lens: torch.Tensor, max_seq_len: int, kv_cache_dtype: str, num_kv_heads: int, scale: float,
                       alibi_slopes: Optional[torch.Tensor], kv_scale: float, tp_rank: int = 0,
                       blocksparse_local_blocks: int = 0, blocksparse_vert_stride: int = 0, blocksparse_block_size: int = 64,
                       blocksparse_head_sliding_step: int = 0) -> torch.Tensor:
        if blocksparse_vert_stride is not None and blocksparse_vert_stride > 1:
            block_size = value_cache.size(-1)
            assert blocksparse_block_size > 0 and blocksparse_block_size % block_size == 0, \
                f"{blocksparse_block_size=} needs to be a multiple of{block_size=} used in block_tables."
        output = torch.empty_like(query)
        block_size = value_cache.shape[3]
        num_seqs, num_heads, head_size = query.shape
        use_v1, parts = _use_v1(max_seq_len, num_seqs, num_heads)
        if use_v1:
            ops.paged_attention_v1(output, query, key_cache, value_cache, num_kv_heads, scale, block_tables, seq_lens, block_size,
                                   max_seq_len, alibi_slopes, kv_cache_dtype, kv_scale, tp_rank, blocksparse_local_blocks,
                                   blocksparse_vert_stride, blocksparse_block_size, blocksparse_head_sliding_step)
        else:
            assert _PARTITION_SIZE % block_size == 0
            tmp_output = torch.empty(size=(num_seqs, num_heads, parts, head_size), dtype=output.dtype, device=output.device)
            exp_sums = torch.empty(size=(num_seqs, num_heads, parts), dtype=torch.float32, device=output.device)
            max_logits = torch.empty_like(exp_sums)
            ops.paged_attention_v2(output, exp_sums, max_logits, tmp_output, query, key_cache, value_cache, num_kv_heads, scale,
                                   block_tables, seq_lens, block_size, max_seq_len, alibi_slopes, kv_cache_dtype, kv_scale, tp_rank,
                                   blocksparse_local_blocks, blocksparse_vert_stride, blocksparse_block_size,
                                   blocksparse_head_sliding_step)
        return output

    @staticmethod
    def forward_decode_fused(query: torch.Tensor, key: torch.Tensor, value: torch.Tensor, key_cache: torch.Tensor,
                             value_cache: torch.Tensor, positions: torch.Tensor, cos_sin_cache: torch.Tensor,
                             slot_mapping: torch.Tensor, block_tables: torch.Tensor, seq_lens: torch.Tensor, max_seq_len: int,
                             kv_cache_dtype: str, num_kv_heads: int, scale: float, kv_scale: float) -> Optional[torch.Tensor]:
        """rotary_embedding (neox) + write_to_paged_cache + forward_decode in one launch.  query / key are the UNROTATED
        projections [tokens, heads, head_size]; query is left untouched.  Returns None -- nothing launched -- when the shape
        is outside the fast path (head 128, block 16, <= 8 query heads per kv head): run the three steps instead."""
        num_seqs, num_heads, head_size = query.shape
        block_size = value_cache.shape[3]
        if head_size != 128 or block_size != 16:
            return None
        output = torch.empty_like(query)
        use_v1, parts = _use_v1(max_seq_len, num_seqs, num_heads)
        tmp = es = ml = None
        if not use_v1:
            tmp = torch.empty(size=(num_seqs, num_heads, parts, head_size), dtype=output.dtype, device=output.device)
            es = torch.empty(size=(num_seqs, num_heads, parts), dtype=torch.float32, device=output.device)
            ml = torch.empty_like(es)
        ok = ops.rope_cache_paged_attention(output, es, ml, tmp, query, key, value, key_cache, value_cache, positions, cos_sin_cache,
                                            slot_mapping.flatten(), num_kv_heads, scale, block_tables, seq_lens, block_size,
                                            max_seq_len, kv_cache_dtype, kv_scale)
        return output if ok else None

    @staticmethod
    def forward_prefix(query: torch.Tensor, key: torch.Tensor, value: torch.Tensor, key_cache: torch.Tensor,
                       value_cache: torch.Tensor, block_tables: torch.Tensor, query_start_loc: torch.Tensor,
                       seq_lens_tensor: torch.Tensor, context_lens: torch.Tensor, max_query_len: int,
                       alibi_slopes: Optional[torch.Tensor], sliding_window: Optional[int], kv_cache_dtype: str = "auto",
                       kv_scale: float = 1.0, scale: Optional[float] = None) -> torch.Tensor:
        """Prefill over a cached prefix (paged_attn.py:185-211 -> context_attention_fwd, prefix_prefill.py:674): query token i of
        a sequence attends to the sequence's first context_len + i + 1 tokens.  16-bit caches go through the prefill kernel
        (`ops.context_attention_fwd`, csrc/attention_prefill.cu: context from the block-table cache, new tokens from `key` /
        `value`, sliding window and ALiBi like the reference).  fp8 caches (which the reference's Triton kernel does not take
        either) fall back to the observation that the new tokens' K / V are already in the paged cache at this point
        (write_to_paged_cache runs first, xformers.py:262-270), so causal prefill IS paged decode attention with one row per
        QUERY TOKEN -- exact, but every query token re-reads its KV prefix.  The reference's trailing arguments are extended
        with the cache dtype / scale it leaves implicit."""
        num_tokens, num_heads, head_size = query.shape
        if (kv_cache_dtype == "auto" and key is not None and value is not None and head_size <= 128 and head_size % 8 == 0
                and (scale is None or abs(scale - float(head_size) ** -0.5) < 1e-7 * scale)):
            # the reference's path: one prefill kernel over (cached context, new tokens), csrc/attention_prefill.cu
            output = torch.empty_like(query)
            ops.context_attention_fwd(query, key, value, output, key_cache, value_cache, block_tables, query_start_loc[:-1],
                                      seq_lens_tensor, context_lens, max_query_len, alibi_slopes, sliding_window)
            return output
        del key, value, max_query_len
        if sliding_window is not None and sliding_window > 0:
            raise NotImplementedError("sliding-window prefix attention over an fp8 cache")
        num_kv_heads = value_cache.shape[1]
        if scale is None:
            scale = float(head_size) ** -0.5
        starts = query_start_loc.to(torch.int64)
        q_lens = (seq_lens_tensor - context_lens).to(torch.int64)
        seq_of_tok = torch.repeat_interleave(torch.arange(q_lens.numel(), device=query.device), q_lens, output_size=num_tokens)
        pos_in_q = torch.arange(num_tokens, device=query.device) - starts[:-1][seq_of_tok] if starts.numel() == q_lens.numel() + 1 \
            else torch.arange(num_tokens, device=query.device) - starts[seq_of_tok]
        tok_lens = (context_lens.to(torch.int64)[seq_of_tok] + pos_in_q + 1).to(torch.int32)
        tok_tables = block_tables.index_select(0, seq_of_tok).contiguous()
        max_len = int(seq_lens_tensor.max().item())
        return PagedAttention.forward_decode(query, key_cache, value_cache, tok_tables, tok_lens, max_len, kv_cache_dtype,
                                             num_kv_heads, scale, alibi_slopes, kv_scale)

    @staticmethod
    def swap_blocks(src_kv_cache: torch.Tensor, dst_kv_cache: torch.Tensor, src_to_dst: torch.Tensor) -> None:
        ops.swap_blocks(src_kv_cache[0], dst_kv_cache[0], src_to_dst)
        ops.swap_blocks(src_kv_cache[1], dst_kv_cache[1], src_to_dst)

    @staticmethod
    def copy_blocks(kv_caches: List[torch.Tensor], src_to_dists: torch.Tensor) -> None:
        ops.copy_blocks([kv[0] for kv in kv_caches], [kv[1] for kv in kv_caches], src_to_dists)
