"""GPU parity of the fused decode attention (rotary_embedding + reshape_and_cache + paged_attention_v1 / v2 in one
launch, csrc/attention_mma.cu) against the chain of the three reference ops on identical inputs -- outputs equal
(same arithmetic, same rounding points), cache contents bit-exact -- and against the oracle chain
(oracle/ops_ref.py: rotary_embedding -> reshape_and_cache -> paged_attention) at the reference's attention tolerance."""
import pytest
import torch

from oracle import ops_ref

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
D, BS = 128, 16


def make_case(S, H, Hkv, ctx, dtype, kv_dtype, seed):
    g = torch.Generator().manual_seed(seed)
    NB = S * ((max(ctx) + BS - 1) // BS) + 8
    maxb = (max(ctx) + BS - 1) // BS
    q = torch.empty((S, H * D)).uniform_(-1, 1, generator=g).to(dtype)
    k = torch.empty((S, Hkv * D)).uniform_(-1, 1, generator=g).to(dtype)
    v = torch.empty((S, Hkv * D)).uniform_(-1, 1, generator=g).to(dtype)
    cache_t = dtype if kv_dtype == "auto" else torch.uint8
    x = 16 // torch.tensor([], dtype=cache_t).element_size()
    if kv_dtype == "auto":
        kc = torch.empty((NB, Hkv, D // x, BS, x)).uniform_(-1, 1, generator=g).to(dtype)
        vc = torch.empty((NB, Hkv, D, BS)).uniform_(-1, 1, generator=g).to(dtype)
    else:
        kc = torch.randint(0, 120, (NB, Hkv, D // x, BS, x), generator=g, dtype=torch.uint8)
        vc = torch.randint(0, 120, (NB, Hkv, D, BS), generator=g, dtype=torch.uint8)
    bt = torch.zeros((S, maxb), dtype=torch.int32)
    perm = torch.randperm(NB, generator=g)
    o = 0
    for s in range(S):
        nb = (ctx[s] + BS - 1) // BS
        bt[s, :nb] = perm[o:o + nb].to(torch.int32)
        o += nb
    lens = torch.tensor(ctx, dtype=torch.int32)
    pos = (lens - 1).long()
    slot = torch.tensor([int(bt[s, (ctx[s] - 1) // BS]) * BS + (ctx[s] - 1) % BS for s in range(S)], dtype=torch.long)
    inv = 1.0 / (500000.0 ** (torch.arange(0, D, 2, dtype=torch.float32) / D))
    fr = torch.outer(torch.arange(max(ctx) + 1, dtype=torch.float32), inv)
    cos_sin = torch.cat([fr.cos(), fr.sin()], dim=-1).to(dtype)
    return dict(q=q, k=k, v=v, kc=kc, vc=vc, bt=bt, lens=lens, pos=pos, slot=slot, cos_sin=cos_sin)


def run_both(ops, c, S, H, Hkv, max_len, kv_dtype, kv_scale, v2):
    scale = D ** -0.5
    dev = {n: t.to(DEV) for n, t in c.items()}
    P = (max_len + 511) // 512

    def bufs():
        if not v2:
            return None, None, None
        return (torch.empty((S, H, P), dtype=torch.float32, device=DEV), torch.empty((S, H, P), dtype=torch.float32, device=DEV),
                torch.empty((S, H, P, D), dtype=c["q"].dtype, device=DEV))

    # fused
    kc1, vc1 = dev["kc"].clone(), dev["vc"].clone()
    out1 = torch.full((S, H, D), float("nan"), dtype=c["q"].dtype, device=DEV)
    es, ml, tmp = bufs()
    q1 = dev["q"].clone()
    ok = ops.rope_cache_paged_attention(out1, es, ml, tmp, q1.view(S, H, D), dev["k"].view(S, Hkv, D), dev["v"].view(S, Hkv, D),
                                        kc1, vc1, dev["pos"], dev["cos_sin"], dev["slot"], Hkv, scale, dev["bt"], dev["lens"],
                                        BS, max_len, kv_dtype, kv_scale)
    assert ok
    assert torch.equal(q1, dev["q"]), "the fused op must leave the query untouched"
    # the three reference ops
    kc2, vc2 = dev["kc"].clone(), dev["vc"].clone()
    q2, k2 = dev["q"].clone(), dev["k"].clone()
    ops.rotary_embedding(dev["pos"], q2, k2, D, dev["cos_sin"], True)
    ops.reshape_and_cache(k2.view(S, Hkv, D), dev["v"].view(S, Hkv, D), kc2, vc2, dev["slot"], kv_dtype, kv_scale)
    out2 = torch.full_like(out1, float("nan"))
    if v2:
        es, ml, tmp = bufs()
        ops.paged_attention_v2(out2, es, ml, tmp, q2.view(S, H, D), kc2, vc2, Hkv, scale, dev["bt"], dev["lens"], BS, max_len,
                               None, kv_dtype, kv_scale)
    else:
        ops.paged_attention_v1(out2, q2.view(S, H, D), kc2, vc2, Hkv, scale, dev["bt"], dev["lens"], BS, max_len, None,
                               kv_dtype, kv_scale)
    torch.cuda.synchronize()
    return out1.cpu(), out2.cpu(), (kc1.cpu(), vc1.cpu()), (kc2.cpu(), vc2.cpu()), q2.cpu(), k2.cpu()


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("kv_dtype", ["auto", "fp8"])
@pytest.mark.parametrize("heads", [(32, 8), (8, 1), (8, 8)])
@pytest.mark.parametrize("v2", [False, True])
def test_fused_equals_three_ops(built_extension, dtype, kv_dtype, heads, v2):
    from nm_vllm_b200 import _custom_ops as ops
    H, Hkv = heads
    ctx = [1, 16, 17, 255, 513, 700, 1030] if v2 else [1, 2, 16, 17, 33, 256, 300]
    S = len(ctx)
    kv_scale = 1.0 if kv_dtype == "auto" else 0.0376
    c = make_case(S, H, Hkv, ctx, dtype, kv_dtype, seed=3)
    out1, out2, cache1, cache2, _, _ = run_both(ops, c, S, H, Hkv, max(ctx), kv_dtype, kv_scale, v2)
    assert torch.equal(cache1[0], cache2[0]) and torch.equal(cache1[1], cache2[1]), "cache contents differ"
    assert not torch.isnan(out1.float()).any()
    torch.testing.assert_close(out1.float(), out2.float(), atol=2e-3, rtol=1e-5)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_fused_vs_oracle_chain(built_extension, dtype):
    from nm_vllm_b200 import _custom_ops as ops
    H, Hkv, ctx = 32, 8, [5, 64, 100, 256]
    S = len(ctx)
    c = make_case(S, H, Hkv, ctx, dtype, "auto", seed=11)
    out1, _, cache1, _, _, _ = run_both(ops, c, S, H, Hkv, max(ctx), "auto", 1.0, False)
    qr, kr = ops_ref.rotary_embedding(c["pos"], c["q"], c["k"], D, c["cos_sin"], True)
    kc, vc = c["kc"].clone(), c["vc"].clone()
    ops_ref.reshape_and_cache(kr.reshape(S, Hkv, D), c["v"].reshape(S, Hkv, D), kc, vc, c["slot"])
    assert torch.equal(kc, cache1[0]) and torch.equal(vc, cache1[1]), "cache write differs from the oracle (bit-exact for kv auto)"
    ref = ops_ref.paged_attention(qr.reshape(S, H, D), kc, vc, c["bt"], c["lens"], D ** -0.5)
    # inputs are O(1) here (not the D^-0.5 scale of the reference's attention tests): allow two ulps of the output dtype
    torch.testing.assert_close(out1.float(), ref.float(), atol=2e-3, rtol=2 * (2.0 ** -8 if dtype == torch.bfloat16 else 2.0 ** -11))


def test_fused_padding_slot_skips_write(built_extension):
    from nm_vllm_b200 import _custom_ops as ops
    H, Hkv, ctx = 8, 1, [20, 40]
    c = make_case(2, H, Hkv, ctx, torch.bfloat16, "auto", seed=5)
    c["slot"][1] = -1   # _PAD_SLOT_ID (model_runner.py:60): no cache write for that row
    _, _, cache1, cache2, _, _ = run_both(ops, c, 2, H, Hkv, 40, "auto", 1.0, False)
    assert torch.equal(cache1[0], cache2[0]) and torch.equal(cache1[1], cache2[1])
