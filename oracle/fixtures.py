"""ORACLE / test infrastructure: seeded synthetic inputs shared by make_golden.py, tests/ and bench.py.
Recipes follow the reference's fixtures: vllm/utils.py:515-563 (create_kv_caches_with_random),
tests/kernels/test_attention.py:152-190, benchmarks/kernels/benchmark_paged_attention.py:30-80."""
import random

import torch


def kv_caches(num_blocks, block_size, num_kv_heads, head_size, dtype, seed, device="cpu"):
    """K [NB,Hkv,D/x,BS,x], V [NB,Hkv,D,BS] uniform(-s, s), s = head_size**-0.5; x = 16 / sizeof(dtype)."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    x = 16 // torch.tensor([], dtype=dtype).element_size()
    scale = head_size**-0.5
    k = (torch.rand((num_blocks, num_kv_heads, head_size // x, block_size, x), generator=g) * 2 - 1) * scale
    v = (torch.rand((num_blocks, num_kv_heads, head_size, block_size), generator=g) * 2 - 1) * scale
    return k.to(dtype).to(device), v.to(dtype).to(device)


def attention_case(num_seqs, num_heads, num_kv_heads, head_size, block_size, num_blocks, max_len, dtype, seed,
                   use_alibi=False, fixed_len=None, device="cpu"):
    rnd = random.Random(seed)
    g = torch.Generator(device="cpu").manual_seed(seed + 1)
    scale = float(head_size**-0.5)
    q = ((torch.rand((num_seqs, num_heads, head_size), generator=g) * 2 - 1) * scale).to(dtype)
    if fixed_len is not None:
        lens = [fixed_len] * num_seqs
    else:
        lens = [rnd.randint(1, max_len) for _ in range(num_seqs)]
        lens[-1] = max_len
    mb = (max(lens) + block_size - 1) // block_size
    bt = torch.tensor([[rnd.randint(0, num_blocks - 1) for _ in range(mb)] for _ in range(num_seqs)], dtype=torch.int32)
    kc, vc = kv_caches(num_blocks, block_size, num_kv_heads, head_size, dtype, seed + 2)
    alibi = torch.randn(num_heads, generator=g, dtype=torch.float32) if use_alibi else None
    d = dict(query=q, key_cache=kc, value_cache=vc, block_tables=bt, seq_lens=torch.tensor(lens, dtype=torch.int32),
             scale=scale, alibi_slopes=alibi, max_seq_len=max(lens), block_size=block_size, num_kv_heads=num_kv_heads)
    if device != "cpu":
        d = {k: (v.to(device) if isinstance(v, torch.Tensor) else v) for k, v in d.items()}
    return d


def cache_write_case(num_tokens, num_kv_heads, head_size, block_size, num_blocks, dtype, seed, with_padding=True):
    rnd = random.Random(seed)
    g = torch.Generator(device="cpu").manual_seed(seed)
    slots = rnd.sample(range(num_blocks * block_size), num_tokens)
    if with_padding and num_tokens > 2:
        slots[1] = -1  # padding token, model_runner.py:60
    qkv = torch.randn((num_tokens, 3, num_kv_heads, head_size), generator=g).to(dtype)
    key, value = qkv[:, 1], qkv[:, 2]  # strided views, like slices of a fused qkv output
    kc, vc = kv_caches(num_blocks, block_size, num_kv_heads, head_size, dtype, seed + 5)
    return dict(key=key, value=value, key_cache=kc, value_cache=vc, slot_mapping=torch.tensor(slots, dtype=torch.long))


def gemm_weights(K, N, seed, dtype=torch.float16):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return torch.randn((K, N), generator=g).to(dtype)


def gemm_acts(M, K, seed, dtype=torch.float16):
    g = torch.Generator(device="cpu").manual_seed(seed + 1000)
    return torch.randn((M, K), generator=g).to(dtype)
