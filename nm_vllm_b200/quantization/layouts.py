"""Product-side layout helpers of the Marlin formats (the one-off transformations the LinearMethods run at load time).

`marlin_permute_scales` / `get_scale_perms`: vllm/model_executor/layers/quantization/gptq_marlin.py:30-56 -- a thread of
the reference kernel reads the 8 scales of its fragment with one 16-byte load, so within every chunk of 64 (grouped) or
32 (channelwise) columns position 8 i + j holds column i + 8 j, resp. position 8 i + jj holds column
2 i + {0,1,8,9,16,17,24,25}[jj].  Pure permutations: exact."""
from typing import List, Tuple

import torch

GPTQ_MARLIN_TILE = 16
GPTQ_MARLIN_MIN_THREAD_N = 64
GPTQ_MARLIN_MIN_THREAD_K = 128
GPTQ_MARLIN_MAX_PARALLEL = 16
GPTQ_MARLIN_SUPPORTED_NUM_BITS = [4, 8]
GPTQ_MARLIN_SUPPORTED_GROUP_SIZES = [-1, 32, 64, 128]
GPTQ_MARLIN_SUPPORTED_SYM = [True]


def get_scale_perms(num_bits: int = 4) -> Tuple[List[int], List[int]]:
    del num_bits   # the tables do not depend on it (kept for signature parity)
    grouped = [i + 8 * j for i in range(8) for j in range(8)]
    single = [2 * i + j for i in range(4) for j in (0, 1, 8, 9, 16, 17, 24, 25)]
    return grouped, single


def get_pack_factor(num_bits: int) -> int:
    assert num_bits in GPTQ_MARLIN_SUPPORTED_NUM_BITS, f"Unsupported num_bits = {num_bits}"
    return 32 // num_bits


def marlin_permute_scales(s: torch.Tensor, size_k: int, size_n: int, group_size: int, num_bits: int) -> torch.Tensor:
    grouped, single = get_scale_perms(num_bits)
    perm = grouped if (group_size < size_k and group_size != -1) else single
    idx = torch.tensor(perm, dtype=torch.long, device=s.device)
    return s.reshape(-1, len(perm)).index_select(1, idx).reshape(-1, size_n).contiguous()


def marlin_workspace(size_n: int, device) -> torch.Tensor:
    """int32 zeros the reference kernel uses as locks (gptq_marlin.py:355-360); ours leaves it untouched."""
    return torch.zeros((size_n // GPTQ_MARLIN_MIN_THREAD_N) * GPTQ_MARLIN_MAX_PARALLEL, dtype=torch.int, device=device)


def gptq_pack(q: torch.Tensor, num_bits: int) -> torch.Tensor:
    """[K, N] integers in [0, 2^bits) -> GPTQ qweight [K / pf, N] int32, nibble i of a word = row pf r + i
    (quant_utils.py:125-146).  Used by the synthetic harness; checkpoints arrive already packed."""
    pf = 32 // num_bits
    K, N = q.shape
    q = q.to(torch.int64).reshape(K // pf, pf, N)
    out = torch.zeros((K // pf, N), dtype=torch.int64, device=q.device)
    for i in range(pf):
        out |= q[:, i, :] << (num_bits * i)
    out = out & 0xFFFFFFFF
    return torch.where(out >= 2 ** 31, out - 2 ** 32, out).to(torch.int32)
