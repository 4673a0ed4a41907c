"""Prefill attention over the paged cache (`context_attention_fwd`, csrc/attention_prefill.cu) against the oracle
(`ops_ref.context_attention_ref`: the masked-softmax definition the reference's own test compares its Triton kernel with,
tests/kernels/test_prefix_prefill.py:24-196).  Cases follow that test (GQA ratios 1 / 8 / 64, head sizes 128 / 96 / 24,
sliding windows, block size 32, ragged query / context lengths) plus block sizes 8 / 16, bf16, ALiBi, empty context,
empty query, and the Llama-3-8B 2k-prefill geometry.  Tolerance: atol 1e-2 * max|ref| + rtol 2e-2 (16-bit P.V)."""
import random

import pytest
import torch

from oracle import ops_ref

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _case(seed, bs, H, Hkv, D, BS, dtype, q_lens, c_lens, amp=1.0):
    g = torch.Generator().manual_seed(seed)
    T = sum(q_lens)
    maxb = max(1, max((c + BS - 1) // BS for c in c_lens))
    NB = bs * maxb + 3
    q = (torch.randn((T, H, D), generator=g) * amp).to(dtype)
    k = (torch.randn((T, Hkv, D), generator=g) * amp).to(dtype)
    v = torch.randn((T, Hkv, D), generator=g).to(dtype)
    kc = (torch.randn((NB, Hkv, D // 8, BS, 8), generator=g) * amp).to(dtype)
    vc = torch.randn((NB, Hkv, D, BS), generator=g).to(dtype)
    perm = torch.randperm(NB, generator=g)[:bs * maxb].view(bs, maxb).to(torch.int32)
    start = torch.tensor([sum(q_lens[:i]) for i in range(bs)], dtype=torch.int32)
    seq = torch.tensor([a + b for a, b in zip(q_lens, c_lens)], dtype=torch.int32)
    ctx = torch.tensor(c_lens, dtype=torch.int32)
    return q, k, v, kc, vc, perm, start, seq, ctx


def _misaligned(t):
    """Same values at a 4-byte (not 16-byte) aligned address: takes the direct-from-global kernel instead of the tiled one."""
    buf = torch.empty(t.numel() + 2, dtype=t.dtype, device=DEV)
    out = buf[2:].view(t.shape)
    out.copy_(t)
    return out


def _run(case, max_q, alibi=None, sw=None, direct=False):
    from nm_vllm_b200.attention.prefix_prefill import context_attention_fwd
    q, k, v, kc, vc, bt, start, seq, ctx = case
    o = torch.full_like(q, float("nan")).to(DEV)
    kd, vd = (_misaligned(k), _misaligned(v)) if direct else (k.to(DEV), v.to(DEV))
    context_attention_fwd(q.to(DEV), kd, vd, o, kc.to(DEV), vc.to(DEV), bt.to(DEV), start.to(DEV), seq.to(DEV),
                          ctx.to(DEV), max_q, None if alibi is None else alibi.to(DEV), sw)
    ref = ops_ref.context_attention_ref(q, k, v, kc, vc, bt, start, seq, ctx, alibi, sw or 0)
    got = o.cpu()
    assert not torch.isnan(got).any(), "an output row was not written"
    tol = 1e-2 * float(ref.float().abs().max()) + 2e-2 * ref.float().abs()
    bad = (got.float() - ref.float()).abs() > tol
    assert int(bad.sum()) == 0, "%d of %d outputs out of tolerance (max err %g)" % (
        int(bad.sum()), bad.numel(), float((got.float() - ref.float()).abs().max()))


@pytest.mark.parametrize("direct", [False, True], ids=["tiled", "direct"])
@pytest.mark.parametrize("G", [1, 8, 64])
@pytest.mark.parametrize("D", [128, 96, 24])
@pytest.mark.parametrize("sw", [0, 16, 256])
def test_contexted_kv_attention_like_the_reference_test(built_extension, G, D, sw, direct):
    random.seed(G * 1000 + D + sw)
    bs, H = 4, 64
    q_lens = [random.randint(16, 200) for _ in range(bs)]
    c_lens = [random.randint(16, 300) for _ in range(bs)]
    _run(_case(G + D + sw, bs, H, H // G, D, 32, torch.float16, q_lens, c_lens), max(q_lens), sw=sw, direct=direct)


@pytest.mark.parametrize("direct", [False, True], ids=["tiled", "direct"])
@pytest.mark.parametrize("BS", [8, 16, 32])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_block_sizes_dtypes_and_ragged_edges(built_extension, BS, dtype, direct):
    # context 0 (pure causal prefill), query of 1 token (a decode step), lengths that are not tile multiples
    q_lens = [1, 17, 64, 65, 5]
    c_lens = [33, 0, 15, 129, 1]
    _run(_case(BS, 5, 8, 2, 64, BS, dtype, q_lens, c_lens), max(q_lens), direct=direct)


def test_stale_cache_beyond_the_context_is_never_read_as_a_value(built_extension):
    # NaN in every cache slot past the context (and in unused blocks): outputs stay finite and correct
    q, k, v, kc, vc, bt, start, seq, ctx = _case(21, 2, 8, 2, 128, 16, torch.bfloat16, [30, 70], [37, 5])
    kc2, vc2 = torch.full_like(kc, float("nan")), torch.full_like(vc, float("nan"))
    for b in range(2):
        c = int(ctx[b])
        for p in range(c):
            blk, off = int(bt[b, p // 16]), p % 16
            kc2[blk, :, :, off, :] = kc[blk, :, :, off, :]
            vc2[blk, :, :, off] = vc[blk, :, :, off]
    for direct in (False, True):
        _run((q, k, v, kc2, vc2, bt, start, seq, ctx), 70, direct=direct)


@pytest.mark.parametrize("direct", [False, True], ids=["tiled", "direct"])
def test_alibi(built_extension, direct):
    H = 8
    slopes = torch.tensor([2.0 ** (-(i + 1)) for i in range(H)], dtype=torch.float32)
    _run(_case(3, 3, H, 4, 128, 16, torch.float16, [40, 7, 100], [50, 120, 0]), 100, alibi=slopes, direct=direct)


def test_empty_query_sequence_and_padding_rows(built_extension):
    # a sequence with nothing to prefill (seq_len == ctx_len) leaves its (absent) rows alone; max_input_len may exceed every query
    _run(_case(9, 3, 4, 4, 32, 16, torch.bfloat16, [12, 0, 30], [20, 64, 3]), 128)


def test_llama3_8b_prefill_geometry(built_extension):
    # 32 query / 8 kv heads, d 128, block 16: one 512-token chunk over a 1024-token prefix (chunked prefill of a 2k prompt)
    _run(_case(11, 1, 32, 8, 128, 16, torch.bfloat16, [512], [1024]), 512)


def test_forward_prefix_goes_through_the_prefill_kernel(built_extension):
    from nm_vllm_b200.attention import PagedAttention
    q, k, v, kc, vc, bt, start, seq, ctx = _case(5, 2, 8, 2, 128, 16, torch.bfloat16, [20, 33], [48, 10])
    qsl = torch.cat([start, torch.tensor([53], dtype=torch.int32)])
    out = PagedAttention.forward_prefix(q.to(DEV), k.to(DEV), v.to(DEV), kc.to(DEV), vc.to(DEV), bt.to(DEV), qsl.to(DEV),
                                        seq.to(DEV), ctx.to(DEV), 33, None, None)
    ref = ops_ref.context_attention_ref(q, k, v, kc, vc, bt, start, seq, ctx)
    assert torch.allclose(out.cpu().float(), ref.float(), atol=2e-2, rtol=2e-2)


def test_rejects_unsupported_shapes(built_extension):
    from nm_vllm_b200.attention.prefix_prefill import context_attention_fwd
    q, k, v, kc, vc, bt, start, seq, ctx = _case(1, 1, 2, 2, 256, 16, torch.float16, [4], [4])
    with pytest.raises(RuntimeError):
        context_attention_fwd(q.to(DEV), k.to(DEV), v.to(DEV), torch.empty_like(q).to(DEV), kc.to(DEV), vc.to(DEV), bt.to(DEV),
                              start.to(DEV), seq.to(DEV), ctx.to(DEV), 4)


def test_model_prefill_equals_token_by_token_decode(built_extension):
    """Size-independent property of the whole path: prefilling a prompt in one call (M = prompt length GEMMs, rotary,
    reshape_and_cache, prefill attention) gives the logits of feeding the same tokens one decode step at a time (M = 1
    GEMMs, fused decode attention), and leaves the same KV cache behind.  Also: a chunked prefill (second chunk attends to
    the first through the paged cache) ends in the same logits."""
    from nm_vllm_b200 import decoder
    torch.manual_seed(0)
    T, BS = 40, 16
    m = decoder.SyntheticLlama(decoder.MINI, quant="w4a16", device=DEV, num_blocks=64)
    toks = torch.randint(0, decoder.MINI.vocab, (T,), device=DEV)
    bt = torch.tensor([[5, 9, 2, 7]], dtype=torch.int32)

    def clear():
        for kc, vc in m.kv:
            kc.zero_(); vc.zero_()

    # token by token
    clear()
    for t in range(T):
        p = torch.tensor([t], device=DEV)
        meta = dict(slot_mapping=(bt[0, t // BS].long() * BS + t % BS).view(1).to(DEV), block_tables=bt.to(DEV),
                    seq_lens=torch.tensor([t + 1], dtype=torch.int32, device=DEV), max_seq_len=t + 1)
        ref = m.forward(toks[t:t + 1], p, meta, return_logits=True)
    kv_ref = [(k.clone(), v.clone()) for k, v in m.kv]
    # one prefill call
    clear()
    pos, meta = m.prefill_meta([T], [0], bt)
    got = m.forward(toks, pos, meta, return_logits=True)
    assert got.shape == ref.shape == (1, decoder.MINI.vocab)
    err = float((got.float() - ref.float()).abs().mean() / ref.float().abs().mean())
    assert err < 2e-2, "prefill vs decode logits: mean relative error %g" % err
    for (k, v), (k0, v0) in zip(m.kv, kv_ref):
        assert torch.allclose(k.float(), k0.float(), atol=3e-2, rtol=3e-2) and torch.allclose(v.float(), v0.float(), atol=3e-2, rtol=3e-2)
    # chunked: 24 tokens, then 16 more over the cached 24
    clear()
    pos, meta = m.prefill_meta([24], [0], bt)
    m.forward(toks[:24], pos, meta, return_logits=True)
    pos, meta = m.prefill_meta([T - 24], [24], bt)
    got2 = m.forward(toks[24:], pos, meta, return_logits=True)
    err2 = float((got2.float() - ref.float()).abs().mean() / ref.float().abs().mean())
    assert err2 < 2e-2, "chunked prefill vs decode logits: mean relative error %g" % err2
