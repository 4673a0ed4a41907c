"""GPU parity through the LinearMethod / attention-backend mirrors (what a model built on the reference's layer
framework calls): GPTQ-format checkpoints -> REPACK -> Marlin GEMM, compressed-tensors schemes, FP8 / INT8 W8A8, and the
decode branch of the attention backend, each against the oracle's dense fp32 statement of the same layer."""
import pytest
import torch

from oracle import ops_ref, quant_layouts as ql

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _elementwise_ok(out, ref, rel=3e-2):
    """no single output may be off by more than `rel` of the output scale (a wrong tile cannot hide in a mean)"""
    scale = float(ref.abs().mean()) + 1e-6
    return float((out.float() - ref.float()).abs().max()) < rel * scale * 8


@pytest.mark.parametrize("bits,group,act_order,dtype", [(4, 128, False, torch.bfloat16), (4, 128, False, torch.float16),
                                                        (4, -1, False, torch.float16), (8, 128, False, torch.float16),
                                                        (4, 64, False, torch.bfloat16), (4, 128, True, torch.float16),
                                                        (8, 128, True, torch.bfloat16)])
def test_gptq_marlin_linear_method(built_extension, bits, group, act_order, dtype):
    from nm_vllm_b200.quantization import GPTQMarlinConfig, GPTQMarlinLinearMethod, GPTQMarlinState
    K, N, M = 1024, 768, 24
    g = torch.Generator().manual_seed(bits * 100 + (group if group > 0 else 7))
    w = (torch.randn((K, N), generator=g) * 0.05).to(dtype)
    w_ref, q, s, g_idx, _ = ql.quantize_weights(w, bits, group, act_order, seed=3)
    method = GPTQMarlinLinearMethod(GPTQMarlinConfig(bits, group, act_order, True, False))
    layer = torch.nn.Module()
    with torch.device(DEV):
        method.create_weights(layer, K, [N // 2, N // 2], K, N, dtype)
    layer.qweight.data.copy_(ql.gptq_pack(q, bits))          # the checkpoint's tensors, GPTQ layout
    layer.scales.data.copy_(s.to(dtype))
    layer.g_idx.data.copy_(g_idx if act_order else torch.arange(K, dtype=torch.int32) // (group if group > 0 else K))
    assert layer.marlin_state == GPTQMarlinState.REPACK
    x = (torch.randn((2, M // 2, K), generator=g)).to(dtype)
    bias = torch.randn(N, generator=g).to(dtype)
    out = method.apply(layer, x.to(DEV), bias.to(DEV))
    assert layer.marlin_state == GPTQMarlinState.READY and out.shape == (2, M // 2, N)
    # the repacked tensors are exactly what the oracle's own packer produces from the (group-sorted) integer weights
    # (rows of one group may come in any order: use the order the layer chose)
    qs = q[layer.g_idx_sort_indices.cpu().long(), :].contiguous() if act_order else q
    if act_order:
        assert bool((layer.g_idx.cpu()[1:] >= layer.g_idx.cpu()[:-1]).all()), "g_idx must be sorted after the repack"
    assert torch.equal(layer.qweight.cpu(), ql.marlin_pack(qs, bits, ql.marlin_perm_tables(bits)[0]))
    ref = ops_ref.marlin_gemm_ref(x.reshape(-1, K), w_ref.to(dtype)).float() + bias.float()
    o = out.reshape(-1, N).cpu()
    assert ops_ref.rel_mean_err(o, ref) < 1e-2 and _elementwise_ok(o, ref)
    out2 = method.apply(layer, x.to(DEV), bias.to(DEV))     # READY state: no second repack
    assert torch.equal(out2, out)


@pytest.mark.parametrize("bits,dtype,M", [(4, torch.float16, 5), (4, torch.bfloat16, 70), (8, torch.float16, 33)])
def test_gptq_marlin_act_order_row_parallel_shards(built_extension, bits, dtype, M):
    """desc_act checkpoint under tensor parallelism (gptq_marlin.py:256-263): each rank holds a K shard whose rows belong to
    arbitrary groups, keeps ALL scale rows and runs with is_k_full=False; the sum of the two shards' outputs is the full
    GEMM."""
    from nm_vllm_b200.quantization import GPTQMarlinConfig, GPTQMarlinLinearMethod
    K, N, group, tp = 1024, 256, 128, 2
    g = torch.Generator().manual_seed(77 + bits)
    w = (torch.randn((K, N), generator=g) * 0.05).to(dtype)
    w_ref, q, s, g_idx, _ = ql.quantize_weights(w, bits, group, True, seed=5)
    x = torch.randn((M, K), generator=g).to(dtype)
    packed = ql.gptq_pack(q, bits)
    pf = 32 // bits
    total = torch.zeros((M, N), dtype=torch.float32)
    for r in range(tp):
        k0, k1 = r * K // tp, (r + 1) * K // tp
        method = GPTQMarlinLinearMethod(GPTQMarlinConfig(bits, group, True, True, False))
        layer = torch.nn.Module()
        with torch.device(DEV):
            method.create_weights(layer, K // tp, [N], K, N, dtype)
        assert layer.is_k_full is False and layer.scales.shape[0] == K // group
        layer.qweight.data.copy_(packed[k0 // pf:k1 // pf])
        layer.scales.data.copy_(s.to(dtype))
        layer.g_idx.data.copy_(g_idx[k0:k1])
        total += method.apply(layer, x[:, k0:k1].contiguous().to(DEV)).float().cpu()
    ref = ops_ref.marlin_gemm_ref(x, w_ref.to(dtype)).float()
    assert ops_ref.rel_mean_err(total, ref) < 1e-2 and _elementwise_ok(total, ref)


@pytest.mark.parametrize("bits,group", [(4, 128), (8, None)])
def test_compressed_tensors_wna16(built_extension, bits, group):
    from nm_vllm_b200.quantization import CompressedTensorsWNA16
    K, N, M, dtype = 512, 256, 9, torch.float16
    g = torch.Generator().manual_seed(5)
    w = (torch.randn((K, N), generator=g) * 0.05).to(dtype)
    w_ref, q, s, _, _ = ql.quantize_weights(w, bits, group if group else -1)
    scheme = CompressedTensorsWNA16("group" if group else "channel", bits, group)
    layer = torch.nn.Module()
    with torch.device(DEV):
        scheme.create_weights(layer=layer, input_size=K, output_partition_sizes=[N], input_size_per_partition=K, params_dtype=dtype,
                              weight_loader=None)
    layer.weight_packed.data.copy_(ql.gptq_pack(q, bits).t().contiguous())      # [N, K / pf]: packed along K, transposed
    layer.weight_scale.data.copy_(s.t().contiguous().to(dtype))
    x = torch.randn((M, K), generator=g).to(dtype)
    out = scheme.apply_weights(layer, x.to(DEV)).cpu()
    ref = ops_ref.marlin_gemm_ref(x, w_ref.to(dtype))
    assert ops_ref.rel_mean_err(out, ref) < 1e-2 and _elementwise_ok(out, ref)


def test_compressed_tensors_w4a16_sparse24(built_extension):
    from nm_vllm_b200.quantization import CompressedTensorsW4A16Sparse24
    K, N, M, dtype = 512, 256, 5, torch.float16
    g = torch.Generator().manual_seed(9)
    w = (torch.randn((K, N), generator=g) * 0.05).to(dtype)
    w24_ref, qp, meta, s = ql.marlin_24_quantize(w, 4, 128)
    scheme = CompressedTensorsW4A16Sparse24("group", 4, 128)
    layer = torch.nn.Module()
    with torch.device(DEV):
        scheme.create_weights(layer=layer, input_size=K, output_partition_sizes=[N], input_size_per_partition=K, params_dtype=dtype,
                              weight_loader=None)
    layer.weight_packed.data.copy_(qp)
    layer.meta.data.copy_(meta)
    layer.scale_packed.data.copy_(s.to(dtype))
    x = torch.randn((M, K), generator=g).to(dtype)
    out = scheme.apply_weights(layer, x.to(DEV)).cpu()
    ref = ops_ref.marlin_gemm_ref(x, w24_ref.to(dtype))
    assert ops_ref.rel_mean_err(out, ref) < 1e-2 and _elementwise_ok(out, ref)


@pytest.mark.parametrize("serialized,scheme", [(False, "dynamic"), (True, "dynamic"), (True, "static")])
def test_fp8_linear_method(built_extension, serialized, scheme):
    from nm_vllm_b200.quantization import Fp8Config, Fp8LinearMethod
    from nm_vllm_b200.quantization.fp8 import per_tensor_quantize
    K, widths, M, dtype = 512, [256, 64, 64], 12, torch.bfloat16
    N = sum(widths)
    g = torch.Generator().manual_seed(21)
    w = (torch.randn((N, K), generator=g) * 0.05)
    method = Fp8LinearMethod(Fp8Config(serialized, scheme))
    layer = torch.nn.Module()
    with torch.device(DEV):
        method.create_weights(layer, K, widths, K, N, dtype)
    if serialized:   # a checkpoint that quantised q / k / v separately
        start, w_eff = 0, torch.empty_like(w)
        for i, n in enumerate(widths):
            sc = w[start:start + n].abs().max() / 448.0
            layer.weight.data[start:start + n] = per_tensor_quantize(w[start:start + n], sc).to(DEV)
            layer.weight_scale.data[i] = sc
            w_eff[start:start + n] = layer.weight.data[start:start + n].cpu().float() * sc
            start += n
        if scheme == "static":
            layer.input_scale.data[:] = 4.0 / 448.0
    else:
        layer.weight.data.copy_(w.to(dtype))
        w_eff = w.to(dtype).float()
    method.process_weights_after_loading(layer)
    assert layer.weight.shape == (K, N) and layer.weight.stride(0) == 1 and layer.weight_scale.numel() == 1
    x = torch.randn((M, K), generator=g).to(dtype)
    bias = torch.randn(N, generator=g).to(dtype)
    out = method.apply(layer, x.to(DEV), bias.to(DEV)).cpu().float()
    ref = x.float() @ w_eff.t() + bias.float()
    # fp8 activations + (re)quantised weights: 2^-4 relative per element, averaged over K
    assert float((out - ref).abs().mean() / ref.abs().mean()) < 4e-2


@pytest.mark.parametrize("strategy,static", [("tensor", True), ("channel", False), ("tensor", False)])
def test_compressed_tensors_w8a8(built_extension, strategy, static):
    from nm_vllm_b200.quantization import CompressedTensorsW8A8
    K, widths, M, dtype = 256, [128, 64], 7, torch.float16
    N = sum(widths)
    g = torch.Generator().manual_seed(31)
    wq = torch.randint(-127, 128, (N, K), generator=g, dtype=torch.int8)
    scheme = CompressedTensorsW8A8(strategy, static)
    layer = torch.nn.Module()
    with torch.device(DEV):
        scheme.create_weights(layer=layer, output_partition_sizes=widths, input_size_per_partition=K, params_dtype=dtype, weight_loader=None)
    layer.weight.data.copy_(wq)
    if strategy == "channel":
        ws = torch.rand((N, 1), generator=g) * 0.01 + 0.001
        layer.weight_scale.data.copy_(ws)
    else:
        per = torch.tensor([0.004, 0.009])
        layer.weight_scale.data.copy_(per)
        ws = torch.cat([per[i].repeat(n) for i, n in enumerate(widths)]).reshape(N, 1)
    if static:
        layer.input_scale.data[:] = 0.03
    scheme.process_weights_after_loading(layer)
    x = torch.randn((M, K), generator=g).to(dtype)
    out = scheme.apply_weights(layer, x.to(DEV)).cpu().float()
    if static:
        xq, xs = ops_ref.scaled_int8_quant(x, torch.tensor([0.03]))
    else:
        xq, xs = ops_ref.scaled_int8_quant(x)
    ref = (xq.float() @ wq.float().t()) * xs.float().reshape(-1, 1) * ws.reshape(1, N)
    torch.testing.assert_close(out, ref.to(dtype).float(), rtol=1e-1, atol=1.0)    # the reference's int8 tolerance (test_cutlass.py:111)
    assert float((out - ref).abs().mean() / ref.abs().mean()) < 2e-2


@pytest.mark.parametrize("kv_dtype", ["auto", "fp8"])
def test_paged_decode_backend_forward(built_extension, kv_dtype):
    """PagedDecodeImpl.forward (cache write + forward_decode) and PagedAttention.forward_decode_fused against the oracle."""
    from nm_vllm_b200.attention import PagedAttention, PagedDecodeBackend
    H, Hkv, D, BS, NB, dtype = 32, 8, 128, 16, 64, torch.bfloat16
    ctx = [3, 40, 100, 17]
    S = len(ctx)
    g = torch.Generator().manual_seed(2)
    q = (torch.randn((S, H * D), generator=g) * 0.3).to(dtype)
    k = (torch.randn((S, Hkv * D), generator=g) * 0.3).to(dtype)
    v = (torch.randn((S, Hkv * D), generator=g) * 0.3).to(dtype)
    cache_t = dtype if kv_dtype == "auto" else torch.uint8
    kv = torch.zeros(PagedDecodeBackend.get_kv_cache_shape(NB, BS, Hkv, D), dtype=cache_t)
    kc, vc = PagedAttention.split_kv_cache(kv, Hkv, D)
    if kv_dtype == "auto":
        kc.copy_((torch.randn(kc.shape, generator=g) * 0.3).to(dtype)); vc.copy_((torch.randn(vc.shape, generator=g) * 0.3).to(dtype))
    else:
        kc.copy_(torch.randint(0, 100, kc.shape, generator=g, dtype=torch.uint8)); vc.copy_(torch.randint(0, 100, vc.shape, generator=g, dtype=torch.uint8))
    maxb = (max(ctx) + BS - 1) // BS
    perm = torch.randperm(NB, generator=g)
    bt = torch.zeros((S, maxb), dtype=torch.int32)
    o = 0
    for s_ in range(S):
        nb = (ctx[s_] + BS - 1) // BS
        bt[s_, :nb] = perm[o:o + nb].to(torch.int32)
        o += nb
    lens = torch.tensor(ctx, dtype=torch.int32)
    slot = torch.tensor([int(bt[i, (ctx[i] - 1) // BS]) * BS + (ctx[i] - 1) % BS for i in range(S)], dtype=torch.long)
    kv_scale = 1.0 if kv_dtype == "auto" else 0.05
    md = PagedDecodeBackend.make_metadata(seq_lens_tensor=lens.to(DEV), max_decode_seq_len=max(ctx), block_tables=bt.to(DEV),
                                          num_prefills=0, num_prefill_tokens=0, num_decode_tokens=S, slot_mapping=slot.to(DEV))
    impl = PagedDecodeBackend.get_impl_cls()(H, D, D ** -0.5, Hkv, None, None, kv_dtype)
    kv_d = kv.to(DEV)
    out = impl.forward(q.to(DEV), k.to(DEV), v.to(DEV), kv_d, md, kv_scale).cpu()
    kc_ref, vc_ref = kc.clone(), vc.clone()
    ops_ref.reshape_and_cache(k.view(S, Hkv, D), v.view(S, Hkv, D), kc_ref, vc_ref, slot, kv_dtype, kv_scale)
    kcd, vcd = PagedAttention.split_kv_cache(kv_d, Hkv, D)
    if kv_dtype == "auto":
        assert torch.equal(kcd.cpu(), kc_ref) and torch.equal(vcd.cpu(), vc_ref)
    ref = ops_ref.paged_attention(q.view(S, H, D), kcd.cpu(), vcd.cpu(), bt, lens, D ** -0.5, None, kv_dtype, kv_scale)
    torch.testing.assert_close(out.view(S, H, D).float(), ref.float(), atol=1e-2 if kv_dtype != "auto" else 4e-3, rtol=1.6e-2)
    with pytest.raises(NotImplementedError):
        md.num_prefill_tokens = 1
        impl.forward(q.to(DEV), k.to(DEV), v.to(DEV), kv_d, md, kv_scale)


def test_forward_prefix_matches_causal_reference(built_extension):
    """PagedAttention.forward_prefix: new tokens (already written to the cache) attend causally to the cached prefix + themselves."""
    from nm_vllm_b200.attention import PagedAttention
    H, Hkv, D, BS, NB, dtype = 8, 2, 128, 16, 64, torch.float16
    ctx, qlen = [20, 0, 33], [5, 17, 1]
    S = len(ctx)
    g = torch.Generator().manual_seed(8)
    tot = [c + q for c, q in zip(ctx, qlen)]
    kc = (torch.randn((NB, Hkv, D // 8, BS, 8), generator=g) * 0.3).to(dtype)
    vc = (torch.randn((NB, Hkv, D, BS), generator=g) * 0.3).to(dtype)
    maxb = (max(tot) + BS - 1) // BS
    perm = torch.randperm(NB, generator=g)
    bt = torch.zeros((S, maxb), dtype=torch.int32)
    o = 0
    for s_ in range(S):
        nb = (tot[s_] + BS - 1) // BS
        bt[s_, :nb] = perm[o:o + nb].to(torch.int32)
        o += nb
    T = sum(qlen)
    q = (torch.randn((T, H, D), generator=g) * 0.3).to(dtype)
    start = torch.tensor([0] + list(torch.tensor(qlen).cumsum(0)), dtype=torch.int32)
    out = PagedAttention.forward_prefix(q.to(DEV), None, None, kc.to(DEV), vc.to(DEV), bt.to(DEV), start.to(DEV),
                                        torch.tensor(tot, dtype=torch.int32, device=DEV), torch.tensor(ctx, dtype=torch.int32, device=DEV),
                                        max(qlen), None, None).cpu()
    # oracle: one paged-attention row per query token with its causal length
    rows_bt, rows_len = [], []
    for s_ in range(S):
        for i in range(qlen[s_]):
            rows_bt.append(bt[s_]); rows_len.append(ctx[s_] + i + 1)
    ref = ops_ref.paged_attention(q, kc, vc, torch.stack(rows_bt), torch.tensor(rows_len, dtype=torch.int32), D ** -0.5)
    torch.testing.assert_close(out.float(), ref.float(), atol=2e-3, rtol=2e-3)
