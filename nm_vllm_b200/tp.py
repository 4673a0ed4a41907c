"""Tensor-parallel sharding of the decode path's tensors (host logic, no device code).

Mirrors how the reference cuts a Llama layer over TP ranks:
  * ColumnParallelLinear / MergedColumnParallelLinear / QKVParallelLinear shard the OUTPUT dim N
    (vllm/model_executor/layers/linear.py:229-330, :333-462 gate_up = two stacked shards, :465-640 qkv = per-head shards,
    kv heads replicated when tp > num_kv_heads);
  * RowParallelLinear shards the INPUT dim K and all-reduces the partial outputs (linear.py:670-803); for a grouped
    quantised weight the scale rows go with their K rows, so K shards must be multiples of the group size
    (gptq_marlin.py:206-232 `is_k_full` / scales partitioning);
  * attention heads and the KV cache follow the qkv shard (llama.py:107-117).

Every function takes/returns plain tensors in [K, N] (input-major) orientation like the quantisation helpers.
"""
from typing import Tuple

import torch


def shard_range(total: int, tp_size: int, tp_rank: int) -> Tuple[int, int]:
    assert total % tp_size == 0, "%d is not divisible by tp=%d" % (total, tp_size)
    per = total // tp_size
    return tp_rank * per, (tp_rank + 1) * per


def kv_head_range(num_kv_heads: int, tp_size: int, tp_rank: int) -> Tuple[int, int]:
    """kv heads of this rank; replicated (one head shared by tp/num_kv_heads ranks) when tp > num_kv_heads
    (linear.py:497-506)."""
    if tp_size >= num_kv_heads:
        assert tp_size % num_kv_heads == 0
        h = tp_rank // (tp_size // num_kv_heads)
        return h, h + 1
    return shard_range(num_kv_heads, tp_size, tp_rank)


def shard_column(w: torch.Tensor, tp_size: int, tp_rank: int) -> torch.Tensor:
    """Column-parallel: rank r keeps output columns [r N/tp, (r+1) N/tp) of w [K, N]."""
    lo, hi = shard_range(w.shape[1], tp_size, tp_rank)
    return w[:, lo:hi].contiguous()


def shard_merged_column(w: torch.Tensor, sizes, tp_size: int, tp_rank: int) -> torch.Tensor:
    """MergedColumnParallelLinear (gate_up): each stacked sub-matrix is column-sharded on its own, shards re-stacked."""
    parts, off = [], 0
    for n in sizes:
        parts.append(shard_column(w[:, off:off + n], tp_size, tp_rank))
        off += n
    assert off == w.shape[1]
    return torch.cat(parts, dim=1).contiguous()


def shard_qkv(w: torch.Tensor, num_heads: int, num_kv_heads: int, head_dim: int, tp_size: int,
              tp_rank: int) -> torch.Tensor:
    """QKVParallelLinear: q heads sharded, kv heads sharded or replicated; w is [K, (H + 2 Hkv) D]."""
    q, k, v = w.split([num_heads * head_dim, num_kv_heads * head_dim, num_kv_heads * head_dim], dim=1)
    qlo, qhi = shard_range(num_heads, tp_size, tp_rank)
    klo, khi = kv_head_range(num_kv_heads, tp_size, tp_rank)
    D = head_dim
    return torch.cat([q[:, qlo * D:qhi * D], k[:, klo * D:khi * D], v[:, klo * D:khi * D]], dim=1).contiguous()


def shard_row(w: torch.Tensor, tp_size: int, tp_rank: int, group_size: int = -1) -> torch.Tensor:
    """Row-parallel: rank r keeps input rows [r K/tp, (r+1) K/tp) of w [K, N]; with grouped quantisation the shard
    must hold whole groups so that its scale rows are exactly rows [lo/g, hi/g) of the full scale tensor."""
    lo, hi = shard_range(w.shape[0], tp_size, tp_rank)
    if group_size > 0:
        assert (hi - lo) % group_size == 0, "K shard %d does not hold whole groups of %d" % (hi - lo, group_size)
    return w[lo:hi].contiguous()


def shard_row_input(x: torch.Tensor, tp_size: int, tp_rank: int) -> torch.Tensor:
    """The activation slice a row-parallel layer consumes (the previous column-parallel layer's local output)."""
    lo, hi = shard_range(x.shape[-1], tp_size, tp_rank)
    return x[..., lo:hi].contiguous()


def all_reduce_(t: torch.Tensor, group=None) -> torch.Tensor:
    """The one collective of the path (linear.py:779-803 tensor_model_parallel_all_reduce)."""
    if torch.distributed.is_initialized() and torch.distributed.get_world_size(group) > 1:
        torch.distributed.all_reduce(t, group=group)
    return t
