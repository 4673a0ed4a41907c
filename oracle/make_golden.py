"""Generates tests/golden/*.npz from the REFERENCE ITSELF, run in the authoring container:
  * the reference's Python quant/pack helpers imported by file path (oracle/refload.py), and
  * the reference's CPU operators compiled unmodified (oracle/_ref/_refcpu_C.so, see oracle/Makefile).
Inputs are regenerated from seeds by oracle/fixtures.py, so only outputs (and the random permutations the reference
drew) are stored.  Re-run:  python -m oracle.make_golden
"""
import os

import numpy as np
import torch

from . import fixtures, refcpu, refload

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

MARLIN_CASES = [  # (K, N, bits, group, act_order)
    (128, 64, 4, -1, False), (128, 64, 4, 128, False), (256, 128, 4, 128, False), (256, 128, 4, 32, False),
    (256, 192, 4, 64, False), (256, 128, 8, 128, False), (128, 64, 8, -1, False), (256, 128, 4, 64, True),
    (256, 128, 8, 32, True),
]
MARLIN24_CASES = [(128, 128, 4, -1), (256, 128, 4, 128), (128, 256, 8, -1), (256, 128, 8, 128)]
ATTN_CASES = [  # (S, H, Hkv, D, BS, NB, max_len, dtype, alibi)
    (3, 8, 2, 128, 16, 64, 300, "bfloat16", False), (2, 4, 4, 64, 16, 32, 100, "float32", False),
    (4, 8, 1, 128, 16, 96, 1100, "bfloat16", False), (2, 6, 3, 80, 16, 32, 77, "float32", True),
]


def _np(t):
    if t.dtype == torch.bfloat16:
        return t.view(torch.int16).numpy()
    return t.numpy()


def main():
    os.makedirs(OUT, exist_ok=True)
    ref = refload.load()
    mu, qu, f24 = ref.marlin_utils, ref.quant_utils, ref.format_24

    # ---- permutation tables -------------------------------------------------------------------------------------
    tabs = {}
    for bits in (4, 8):
        tabs[f"perm{bits}"] = ref.marlin_perms.marlin_perm[bits].numpy()
        tabs[f"scale_perm{bits}"] = np.asarray(ref.marlin_perms.marlin_scale_perm[bits])
        tabs[f"scale_perm_single{bits}"] = np.asarray(ref.marlin_perms.marlin_scale_perm_single[bits])
        tabs[f"perm24_{bits}"] = ref.marlin_24_perms.marlin_24_perm[bits].numpy()
        tabs[f"scale_perm24_{bits}"] = np.asarray(ref.marlin_24_perms.marlin_24_scale_perm[bits])
        tabs[f"scale_perm_single24_{bits}"] = np.asarray(ref.marlin_24_perms.marlin_24_scale_perm_single[bits])
    np.savez_compressed(os.path.join(OUT, "perm_tables.npz"), **tabs)

    # ---- dense Marlin: quantize, gptq pack, marlin pack, scale permutation --------------------------------------------
    for i, (K, N, bits, group, act) in enumerate(MARLIN_CASES):
        w = fixtures.gemm_weights(K, N, seed=100 + i)
        torch.manual_seed(7 + i)
        gs = K if group == -1 else group
        w_ref, q_w, s, g_idx, rand_perm = qu.quantize_weights(w, bits, gs, act)
        gptq = qu.gptq_pack(q_w, bits, K, N)
        sort_idx = torch.empty(0, dtype=torch.int32)
        q_sorted, g_sorted = q_w, g_idx
        if act:
            q_sorted, g_sorted, sort_idx = qu.sort_weights(q_w, g_idx)
        mq = mu.marlin_weights(q_sorted, K, N, bits, ref.marlin_perms.marlin_perm[bits])
        ms = mu.marlin_permute_scales(s, K, N, gs, ref.marlin_perms.marlin_scale_perm[bits],
                                      ref.marlin_perms.marlin_scale_perm_single[bits])
        np.savez_compressed(os.path.join(OUT, f"marlin_{i}.npz"), K=K, N=N, bits=bits, group=group, act=int(act),
                            seed=100 + i, w_ref=_np(w_ref), q_w=_np(q_w), s=_np(s), g_idx=_np(g_idx),
                            rand_perm=_np(rand_perm), gptq=_np(gptq), sort_idx=_np(sort_idx), g_sorted=_np(g_sorted),
                            marlin_q=_np(mq), marlin_s=_np(ms))

    # ---- Marlin-24 ---------------------------------------------------------------------------------------------
    for i, (K, N, bits, group) in enumerate(MARLIN24_CASES):
        w = fixtures.gemm_weights(K, N, seed=200 + i)
        w24_ref, q_comp, meta, s24 = mu.marlin_24_quantize(w, bits, group)
        # intermediate stages for layer-by-layer pinning
        w_24, mask = mu.inject_24(w, K, N)
        gs = K if group == -1 else group
        _, q_w_24, s_plain, _, _ = qu.quantize_weights(w_24, bits, gs, False)
        comp_plain, meta_plain = mu.compress_quantized_24_weight(q_w_24, K, N, bits)
        dense_back = f24.sparse_semi_structured_to_dense_cutlass(
            (comp_plain - (1 << bits) // 2).t().contiguous(), meta_plain.reshape(N, K // 16))
        np.savez_compressed(os.path.join(OUT, f"marlin24_{i}.npz"), K=K, N=N, bits=bits, group=group, seed=200 + i,
                            w24_ref=_np(w24_ref), q_comp=_np(q_comp), meta=_np(meta), s24=_np(s24), mask=_np(mask),
                            q_w_24=_np(q_w_24), comp_plain=_np(comp_plain), dense_back=_np(dense_back))

    # ---- attention / cache through the reference's CPU operators ---------------------------------------------------
    assert refcpu.available(), refcpu.why_unavailable()
    for i, (S, H, Hkv, D, BS, NB, max_len, dts, alibi) in enumerate(ATTN_CASES):
        dt = getattr(torch, dts)
        c = fixtures.attention_case(S, H, Hkv, D, BS, NB, max_len, dt, seed=300 + i, use_alibi=alibi)
        out1 = torch.empty_like(c["query"])
        refcpu.paged_attention_v1(out1, c["query"], c["key_cache"], c["value_cache"], Hkv, c["scale"], c["block_tables"],
                                  c["seq_lens"], BS, c["max_seq_len"], c["alibi_slopes"])
        P = (c["max_seq_len"] + 511) // 512
        out2 = torch.empty_like(c["query"])
        tmp = torch.zeros((S, H, P, D), dtype=dt)
        es = torch.zeros((S, H, P), dtype=torch.float32)
        ml = torch.zeros((S, H, P), dtype=torch.float32)
        refcpu.paged_attention_v2(out2, es, ml, tmp, c["query"], c["key_cache"], c["value_cache"], Hkv, c["scale"],
                                  c["block_tables"], c["seq_lens"], BS, c["max_seq_len"], c["alibi_slopes"])
        np.savez_compressed(os.path.join(OUT, f"attn_{i}.npz"), params=np.asarray([S, H, Hkv, D, BS, NB, max_len]),
                            dtype=dts, alibi=int(alibi), seed=300 + i, out_v1=_np(out1), out_v2=_np(out2))
    for i, (T, Hkv, D, BS, NB, dts) in enumerate([(7, 2, 128, 16, 16, "bfloat16"), (5, 3, 64, 16, 8, "float32")]):
        dt = getattr(torch, dts)
        c = fixtures.cache_write_case(T, Hkv, D, BS, NB, dt, seed=400 + i)
        refcpu.reshape_and_cache(c["key"], c["value"], c["key_cache"], c["value_cache"], c["slot_mapping"])
        np.savez_compressed(os.path.join(OUT, f"cache_{i}.npz"), params=np.asarray([T, Hkv, D, BS, NB]), dtype=dts,
                            seed=400 + i, key_cache=_np(c["key_cache"]), value_cache=_np(c["value_cache"]))
    print("golden vectors written to", OUT)


if __name__ == "__main__":
    main()
