"""CPU: the oracle restatements (oracle/) replay the golden vectors produced by the reference itself
(oracle/make_golden.py: reference Python helpers + reference CPU operators compiled unmodified)."""
import numpy as np
import pytest
import torch

from oracle import fixtures, ops_ref, quant_layouts as ql
from tests.util import as_torch, golden, golden_files


def test_perm_tables_match_reference():
    g = golden("perm_tables.npz")
    for bits in (4, 8):
        perm, sp, sps = ql.marlin_perm_tables(bits)
        assert np.array_equal(perm, g[f"perm{bits}"])
        assert np.array_equal(np.asarray(sp), g[f"scale_perm{bits}"])
        assert np.array_equal(np.asarray(sps), g[f"scale_perm_single{bits}"])
        perm24, sp24, sps24 = ql.marlin24_perm_tables(bits)
        assert np.array_equal(perm24, g[f"perm24_{bits}"])
        assert np.array_equal(np.asarray(sp24), g[f"scale_perm24_{bits}"])
        assert np.array_equal(np.asarray(sps24), g[f"scale_perm_single24_{bits}"])


@pytest.mark.parametrize("fname", golden_files("marlin_"))
def test_marlin_quantize_pack_match_reference(fname):
    g = golden(fname)
    K, N, bits, group, act, seed = (int(g[k]) for k in ("K", "N", "bits", "group", "act", "seed"))
    w = fixtures.gemm_weights(K, N, seed=seed)
    rp = as_torch(g["rand_perm"]) if act else None
    w_ref, q_w, s, g_idx, _ = ql.quantize_weights(w, bits, group, bool(act), rand_perm=rp)
    assert torch.equal(q_w, as_torch(g["q_w"]).to(q_w.dtype))
    assert torch.equal(s, as_torch(g["s"]))
    assert torch.equal(w_ref, as_torch(g["w_ref"]))
    assert torch.equal(ql.gptq_pack(q_w, bits), as_torch(g["gptq"]))
    sort_idx = torch.empty(0, dtype=torch.int32)
    if act:
        assert torch.equal(g_idx, as_torch(g["g_idx"]))
        q_w, g_sorted, sort_idx = ql.sort_weights(q_w, g_idx)
        assert torch.equal(g_sorted, as_torch(g["g_sorted"]))
    packed = ql.marlin_pack(q_w, bits)
    assert torch.equal(packed, as_torch(g["marlin_q"]))
    perm, sp, sps = ql.marlin_perm_tables(bits)
    gs = K if group == -1 else group
    assert torch.equal(ql.permute_scales(s, K, N, gs, sp, sps), as_torch(g["marlin_s"]))
    # inverse maps used by the GEMM oracle
    assert torch.equal(ql.marlin_unpack(packed, K, N, bits), q_w)
    assert torch.equal(ql.unpermute_scales(as_torch(g["marlin_s"]), K, N, gs, sp, sps), s)


@pytest.mark.parametrize("fname", golden_files("marlin24_"))
def test_marlin24_match_reference(fname):
    g = golden(fname)
    K, N, bits, group, seed = (int(g[k]) for k in ("K", "N", "bits", "group", "seed"))
    w = fixtures.gemm_weights(K, N, seed=seed)
    w_24, mask = ql.inject_24(w)
    assert torch.equal(mask, as_torch(g["mask"]))
    gs = K if group == -1 else group
    _, q_w_24, _, _, _ = ql.quantize_weights(w_24, bits, gs, False)
    assert torch.equal(q_w_24, as_torch(g["q_w_24"]).to(q_w_24.dtype))
    comp, meta = ql.compress_quantized_24_weight(q_w_24, K, N, bits)
    assert torch.equal(comp, as_torch(g["comp_plain"]).to(comp.dtype))
    assert torch.equal(meta, as_torch(g["meta"]))
    w24_ref, q_comp, meta2, s24 = ql.marlin_24_quantize(w, bits, group)
    assert torch.equal(w24_ref, as_torch(g["w24_ref"]))
    assert torch.equal(q_comp, as_torch(g["q_comp"]))
    assert torch.equal(meta2, as_torch(g["meta"]))
    assert torch.equal(s24, as_torch(g["s24"]))
    # compress -> decompress is exact and equals the reference's own decompression
    zp = (1 << bits) // 2
    dense = ql.decompress_24((comp - zp).t().contiguous(), meta.reshape(N, K // 16))
    assert torch.equal(dense, as_torch(g["dense_back"]).to(dense.dtype))
    assert torch.equal(dense.t() + zp, q_w_24)


@pytest.mark.parametrize("fname", golden_files("attn_"))
def test_attention_restatement_vs_reference_cpu_op(fname):
    g = golden(fname)
    S, H, Hkv, D, BS, NB, max_len = (int(v) for v in g["params"])
    dts = str(g["dtype"])
    dt = getattr(torch, dts)
    c = fixtures.attention_case(S, H, Hkv, D, BS, NB, max_len, dt, seed=int(g["seed"]), use_alibi=bool(g["alibi"]))
    out = ops_ref.paged_attention(c["query"], c["key_cache"], c["value_cache"], c["block_tables"], c["seq_lens"],
                                  c["scale"], c["alibi_slopes"])
    for key in ("out_v1", "out_v2"):
        ref = as_torch(g[key], dts)
        # tolerance of the reference's own test (tests/kernels/test_attention.py:276-284)
        assert torch.allclose(out.float(), ref.float(), atol=1e-3, rtol=1e-5), (out.float() - ref.float()).abs().max()


@pytest.mark.parametrize("fname", golden_files("cache_"))
def test_cache_write_restatement_vs_reference_cpu_op(fname):
    g = golden(fname)
    T, Hkv, D, BS, NB = (int(v) for v in g["params"])
    dts = str(g["dtype"])
    c = fixtures.cache_write_case(T, Hkv, D, BS, NB, getattr(torch, dts), seed=int(g["seed"]))
    ops_ref.reshape_and_cache(c["key"], c["value"], c["key_cache"], c["value_cache"], c["slot_mapping"])
    assert torch.equal(c["key_cache"], as_torch(g["key_cache"], dts))
    assert torch.equal(c["value_cache"], as_torch(g["value_cache"], dts))


def test_fp8_codec_corner_cases():
    # SATFINITE: finite overflow saturates to +-448 (e4m3 has no inf); round-to-nearest-even; e5m2 saturates to 57344
    x = torch.tensor([0.0, 1.0, -1.0, 447.0, 448.0, 449.0, 1e6, -1e6, 0.0009765625, 3.3], dtype=torch.float32)
    q = ops_ref.fp8_quant(x, 1.0, "fp8")
    back = ops_ref.fp8_dequant(q, 1.0, torch.float32, "fp8")
    assert back[5] == 448 and back[6] == 448 and back[7] == -448
    assert back[1] == 1 and back[2] == -1 and back[4] == 448
    q5 = ops_ref.fp8_quant(torch.tensor([1e9, -1e9]), 1.0, "fp8_e5m2")
    assert ops_ref.fp8_dequant(q5, 1.0, torch.float32, "fp8_e5m2").tolist() == [57344.0, -57344.0]
    # scale semantics: store x/scale, load *scale
    x = torch.linspace(-2, 2, 101)
    r = ops_ref.fp8_dequant(ops_ref.fp8_quant(x, 0.05, "fp8"), 0.05, torch.float32, "fp8")
    assert torch.allclose(r, x, atol=1e-3, rtol=0.1)


def test_context_attention_oracle_agrees_with_the_paged_attention_oracle():
    """Two independent restatements meet: prefill over (cached context + new tokens) == one paged-attention row per query
    token once the new tokens' K / V sit in the cache (the paged-attention oracle is pinned to the reference's CPU op)."""
    import torch
    from oracle import ops_ref
    H, Hkv, D, BS, NB = 4, 2, 32, 16, 24
    ctx, ql = [20, 0, 33], [5, 17, 1]
    g = torch.Generator().manual_seed(4)
    T = sum(ql)
    q = torch.randn((T, H, D), generator=g).to(torch.bfloat16)
    k = torch.randn((T, Hkv, D), generator=g).to(torch.bfloat16)
    v = torch.randn((T, Hkv, D), generator=g).to(torch.bfloat16)
    kc = torch.randn((NB, Hkv, D // 8, BS, 8), generator=g).to(torch.bfloat16)
    vc = torch.randn((NB, Hkv, D, BS), generator=g).to(torch.bfloat16)
    maxb = 4
    bt = torch.randperm(NB, generator=g)[:3 * maxb].view(3, maxb).to(torch.int32)
    start = torch.tensor([0, 5, 22], dtype=torch.int32)
    seq = torch.tensor([c + n for c, n in zip(ctx, ql)], dtype=torch.int32)
    out = ops_ref.context_attention_ref(q, k, v, kc, vc, bt, start, seq, torch.tensor(ctx, dtype=torch.int32))
    # write the new tokens into a copy of the cache, then decode-style rows
    kc2, vc2 = kc.clone(), vc.clone()
    slots = []
    for b in range(3):
        for i in range(ql[b]):
            p = ctx[b] + i
            slots.append(int(bt[b, p // BS]) * BS + p % BS)
    ops_ref.reshape_and_cache(k, v, kc2, vc2, torch.tensor(slots, dtype=torch.int64))
    rows_bt = torch.stack([bt[b] for b in range(3) for _ in range(ql[b])])
    rows_len = torch.tensor([ctx[b] + i + 1 for b in range(3) for i in range(ql[b])], dtype=torch.int32)
    ref = ops_ref.paged_attention(q, kc2, vc2, rows_bt, rows_len, D ** -0.5)
    torch.testing.assert_close(out.float(), ref.float(), atol=2e-2, rtol=2e-2)
