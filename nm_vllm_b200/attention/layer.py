"""`Attention` module: mirror of vllm/attention/layer.py:16-100 -- owns the backend impl and the per-layer `_kv_scale`
(python float, 1.0 unless an fp8 checkpoint brings one through `Fp8KVCacheMethod`), forwards to `impl.forward`."""
from typing import Any, Dict, List, Optional

import torch
import torch.nn as nn

from .backend import AttentionMetadata, PagedDecodeBackend


class Attention(nn.Module):
    def __init__(self, num_heads: int, head_size: int, scale: float, num_kv_heads: Optional[int] = None,
                 alibi_slopes: Optional[List[float]] = None, cache_config: Optional[Any] = None,
                 quant_config: Optional[Any] = None, blocksparse_params: Optional[Dict[str, Any]] = None) -> None:
        super().__init__()
        kv_cache_dtype = getattr(cache_config, "cache_dtype", "auto") if cache_config is not None else "auto"
        sliding_window = getattr(cache_config, "sliding_window", None) if cache_config is not None else None
        if num_kv_heads is None:
            num_kv_heads = num_heads
        self.kv_cache_dtype = kv_cache_dtype
        self._kv_scale = 1.0   # ignored unless the cache is fp8 (layer.py:50-56)
        quant_method = quant_config.get_quant_method(self) if quant_config is not None else None
        if quant_method is not None:
            if self.kv_cache_dtype == "fp8_e5m2":
                raise ValueError("fp8_e5m2 kv-cache is not supported with fp8 checkpoints.")
            # a scalar parameter "kv_scale" the checkpoint loader fills; process_weights_after_loading turns it into _kv_scale
            self.quant_method = quant_method
            self.quant_method.create_weights(self)
        self.impl = PagedDecodeBackend.get_impl_cls()(num_heads, head_size, scale, num_kv_heads, alibi_slopes, sliding_window,
                                                      kv_cache_dtype, blocksparse_params)

    def forward(self, query: torch.Tensor, key: torch.Tensor, value: torch.Tensor, kv_cache: Optional[torch.Tensor],
                attn_metadata: AttentionMetadata) -> torch.Tensor:
        return self.impl.forward(query, key, value, kv_cache, attn_metadata, self._kv_scale)

    def extra_repr(self) -> str:
        return (f"head_size={self.impl.head_size}, num_heads={self.impl.num_heads}, num_kv_heads={self.impl.num_kv_heads}, "
                f"scale={self.impl.scale}")
