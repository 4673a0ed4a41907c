"""`context_attention_fwd`: mirror of vllm/attention/ops/prefix_prefill.py:674-799 (the reference's Triton prefill-over-prefix
kernels) over the CUDA kernel in csrc/attention_prefill.cu.  Same positional arguments and in-place output."""
from typing import Optional

import torch

from .. import _custom_ops as ops


def context_attention_fwd(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, o: torch.Tensor, k_cache: torch.Tensor,
                          v_cache: torch.Tensor, b_loc: torch.Tensor, b_start_loc: torch.Tensor, b_seq_len: torch.Tensor,
                          b_ctx_len: torch.Tensor, max_input_len: int, alibi_slopes: Optional[torch.Tensor] = None,
                          sliding_window: Optional[int] = None) -> None:
    ops.context_attention_fwd(q, k, v, o, k_cache, v_cache, b_loc, b_start_loc, b_seq_len, b_ctx_len, max_input_len,
                              alibi_slopes, sliding_window)
