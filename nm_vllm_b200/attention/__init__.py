"""PagedAttention wrapper, metadata and the decode half of the attention backend contract."""
from .backend import (AttentionBackend, AttentionImpl, AttentionMetadata, PagedDecodeBackend, PagedDecodeImpl,
                      PagedDecodeMetadata)
from .layer import Attention
from .paged_attn import PagedAttention, PagedAttentionMetadata
