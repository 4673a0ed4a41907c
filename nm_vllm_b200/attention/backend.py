"""The decode half of the attention backend contract: `AttentionBackend` / `AttentionMetadata` / `AttentionImpl`
(vllm/attention/backends/abstract.py:9-132) and a paged-attention backend whose `forward` is the decode branch of
`XFormersImpl.forward` (vllm/attention/backends/xformers.py:238-338): cache write through
`PagedAttention.write_to_paged_cache`, then `PagedAttention.forward_decode` over `decode_metadata`.  Prefill tokens are
outside this build's scope: a batch that contains any raises."""
from abc import ABC, abstractmethod
from dataclasses import dataclass, fields
from typing import Any, Dict, Generic, List, Optional, Set, Tuple, Type, TypeVar

import torch

from .paged_attn import PagedAttention, PagedAttentionMetadata


class AttentionBackend(ABC):
    @staticmethod
    @abstractmethod
    def get_name() -> str:
        raise NotImplementedError

    @staticmethod
    @abstractmethod
    def get_impl_cls() -> Type["AttentionImpl"]:
        raise NotImplementedError

    @staticmethod
    @abstractmethod
    def get_metadata_cls() -> Type["AttentionMetadata"]:
        raise NotImplementedError

    @classmethod
    def make_metadata(cls, *args, **kwargs) -> "AttentionMetadata":
        return cls.get_metadata_cls()(*args, **kwargs)

    @staticmethod
    @abstractmethod
    def get_kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int, head_size: int) -> Tuple[int, ...]:
        raise NotImplementedError

    @staticmethod
    @abstractmethod
    def swap_blocks(src_kv_cache: torch.Tensor, dst_kv_cache: torch.Tensor, src_to_dst: torch.Tensor) -> None:
        raise NotImplementedError

    @staticmethod
    @abstractmethod
    def copy_blocks(kv_caches: List[torch.Tensor], src_to_dists: torch.Tensor) -> None:
        raise NotImplementedError


@dataclass
class AttentionMetadata:
    num_prefills: int
    num_prefill_tokens: int
    num_decode_tokens: int
    slot_mapping: torch.Tensor     # [tokens] int64 slot of every token of the step, -1 = padding

    @property
    def prefill_metadata(self) -> Optional["AttentionMetadata"]:
        return None

    @property
    def decode_metadata(self) -> Optional["AttentionMetadata"]:
        return None

    def asdict_zerocopy(self, skip_fields: Optional[Set[str]] = None) -> Dict[str, Any]:
        skip_fields = skip_fields or set()
        return {f.name: getattr(self, f.name) for f in fields(self) if f.name not in skip_fields}


T = TypeVar("T", bound=AttentionMetadata)


class AttentionImpl(ABC, Generic[T]):
    @abstractmethod
    def __init__(self, num_heads: int, head_size: int, scale: float, num_kv_heads: Optional[int] = None,
                 alibi_slopes: Optional[List[float]] = None, sliding_window: Optional[int] = None,
                 kv_cache_dtype: str = "auto", blocksparse_params: Optional[Dict[str, Any]] = None) -> None:
        raise NotImplementedError

    @abstractmethod
    def forward(self, query: torch.Tensor, key: torch.Tensor, value: torch.Tensor, kv_cache: torch.Tensor, attn_metadata: T,
                kv_scale: float = 1.0) -> torch.Tensor:
        raise NotImplementedError


@dataclass
class PagedDecodeMetadata(AttentionMetadata, PagedAttentionMetadata):
    """The fields the decode branch consumes (XFormersMetadata, xformers.py:66-175, minus the prefill-only ones)."""
    use_cuda_graph: bool = False

    @property
    def prefill_metadata(self) -> Optional["PagedDecodeMetadata"]:
        return None

    @property
    def decode_metadata(self) -> Optional["PagedDecodeMetadata"]:
        if self.num_decode_tokens == 0:
            return None
        assert self.block_tables is not None and self.seq_lens_tensor is not None
        return self


class PagedDecodeBackend(AttentionBackend):
    @staticmethod
    def get_name() -> str:
        return "b200-paged-decode"

    @staticmethod
    def get_impl_cls() -> Type["PagedDecodeImpl"]:
        return PagedDecodeImpl

    @staticmethod
    def get_metadata_cls() -> Type["PagedDecodeMetadata"]:
        return PagedDecodeMetadata

    @staticmethod
    def get_kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int, head_size: int) -> Tuple[int, ...]:
        return PagedAttention.get_kv_cache_shape(num_blocks, block_size, num_kv_heads, head_size)

    @staticmethod
    def swap_blocks(src_kv_cache: torch.Tensor, dst_kv_cache: torch.Tensor, src_to_dst: torch.Tensor) -> None:
        PagedAttention.swap_blocks(src_kv_cache, dst_kv_cache, src_to_dst)

    @staticmethod
    def copy_blocks(kv_caches: List[torch.Tensor], src_to_dists: torch.Tensor) -> None:
        PagedAttention.copy_blocks(kv_caches, src_to_dists)


class PagedDecodeImpl(AttentionImpl[PagedDecodeMetadata]):
    def __init__(self, num_heads: int, head_size: int, scale: float, num_kv_heads: Optional[int] = None,
                 alibi_slopes: Optional[List[float]] = None, sliding_window: Optional[int] = None,
                 kv_cache_dtype: str = "auto", blocksparse_params: Optional[Dict[str, Any]] = None) -> None:
        assert blocksparse_params is None, "blocksparse attention is outside this build's scope"
        if sliding_window is not None:
            raise ValueError("sliding window is handled by the block manager in the reference; not part of the decode kernels")
        self.num_heads, self.head_size, self.scale = num_heads, head_size, float(scale)
        self.num_kv_heads = num_heads if num_kv_heads is None else num_kv_heads
        self.alibi_slopes = torch.tensor(alibi_slopes, dtype=torch.float32) if alibi_slopes is not None else None
        self.kv_cache_dtype = kv_cache_dtype
        assert self.num_heads % self.num_kv_heads == 0
        if head_size not in PagedAttention.get_supported_head_sizes():
            raise ValueError(f"Head size {head_size} is not supported by PagedAttention. "
                             f"Supported head sizes are: {PagedAttention.get_supported_head_sizes()}.")

    def forward(self, query: torch.Tensor, key: torch.Tensor, value: torch.Tensor, kv_cache: Optional[torch.Tensor],
                attn_metadata: PagedDecodeMetadata, kv_scale: float = 1.0) -> torch.Tensor:
        """query [tokens, H * D], key / value [tokens, Hkv * D], kv_cache [2, NB, BS * Hkv * D] -> [tokens, H * D]."""
        if attn_metadata.num_prefill_tokens != 0:
            raise NotImplementedError("prefill tokens in the batch: prefill attention is outside this build's scope")
        query = query.view(-1, self.num_heads, self.head_size)
        key = key.view(-1, self.num_kv_heads, self.head_size)
        value = value.view(-1, self.num_kv_heads, self.head_size)
        assert kv_cache is not None, "decode needs the paged KV cache"
        key_cache, value_cache = PagedAttention.split_kv_cache(kv_cache, self.num_kv_heads, self.head_size)
        PagedAttention.write_to_paged_cache(key, value, key_cache, value_cache, attn_metadata.slot_mapping, self.kv_cache_dtype, kv_scale)
        assert key.shape[0] == attn_metadata.num_decode_tokens and value.shape[0] == attn_metadata.num_decode_tokens
        output = torch.empty_like(query)
        if decode_meta := attn_metadata.decode_metadata:
            if self.alibi_slopes is not None and self.alibi_slopes.device != query.device:
                self.alibi_slopes = self.alibi_slopes.to(query.device)
            output = PagedAttention.forward_decode(query, key_cache, value_cache, decode_meta.block_tables, decode_meta.seq_lens_tensor,
                                                   decode_meta.max_decode_seq_len, self.kv_cache_dtype, self.num_kv_heads, self.scale,
                                                   self.alibi_slopes, kv_scale)
        return output.view(-1, self.num_heads * self.head_size)
