"""GPU parity: paged_attention_v1 / v2 through the torch shim vs the oracle restatement of the reference test's
ref_single_query_cached_kv_attention, and vs outputs of the reference's own CPU operator (golden).  Tolerances are the
reference's (tests/kernels/test_attention.py:276-284): atol 1e-3 (1e-2 with fp8 KV), rtol 1e-5."""
import pytest
import torch

from oracle import fixtures, ops_ref
from tests.util import as_torch, golden, golden_files

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def ops(built_extension):
    from nm_vllm_b200 import _custom_ops
    return _custom_ops


def run_case(ops, c, version, kv_dtype="auto", kv_scale=1.0):
    q = c["query"].to(DEV)
    S, H, D = q.shape
    kc, vc = c["key_cache"], c["value_cache"]
    if kv_dtype != "auto":
        kq = torch.empty(kc.shape[:2] + (D // 16, kc.shape[3], 16), dtype=torch.uint8, device=DEV)
        # re-layout K for x=16 then quantise with the product's own convert_fp8 (what the reference fixtures do)
        k_full = kc.permute(0, 1, 3, 2, 4).reshape(kc.shape[0], kc.shape[1], kc.shape[3], D)  # [NB,H,BS,D]
        k16 = k_full.reshape(kc.shape[0], kc.shape[1], kc.shape[3], D // 16, 16).permute(0, 1, 3, 2, 4).contiguous()
        if kv_dtype == "fp8_e5m2":  # convert_fp8 only takes auto/fp8/fp8_e4m3 (cache_kernels.cu:356-387): quantise on the host
            kq = ops_ref.fp8_quant(k16, kv_scale, kv_dtype).to(DEV)
            vq = ops_ref.fp8_quant(vc, kv_scale, kv_dtype).to(DEV)
        else:
            ops.convert_fp8(kq, k16.to(DEV), kv_scale, kv_dtype)
            vq = torch.empty(vc.shape, dtype=torch.uint8, device=DEV)
            ops.convert_fp8(vq, vc.to(DEV), kv_scale, kv_dtype)
        kc_d, vc_d = kq, vq
        kc_ref, vc_ref = kq.cpu(), vq.cpu()
    else:
        kc_d, vc_d = kc.to(DEV), vc.to(DEV)
        kc_ref, vc_ref = kc, vc
    bt, sl = c["block_tables"].to(DEV), c["seq_lens"].to(DEV)
    alibi = c["alibi_slopes"].to(DEV) if c["alibi_slopes"] is not None else None
    out = torch.full_like(q, float("nan"))
    if version == "v1":
        ops.paged_attention_v1(out, q, kc_d, vc_d, c["num_kv_heads"], c["scale"], bt, sl, c["block_size"],
                               c["max_seq_len"], alibi, kv_dtype, kv_scale)
    else:
        P = (c["max_seq_len"] + 511) // 512
        tmp = torch.empty((S, H, P, D), dtype=q.dtype, device=DEV)
        es = torch.empty((S, H, P), dtype=torch.float32, device=DEV)
        ml = torch.empty_like(es)
        ops.paged_attention_v2(out, es, ml, tmp, q, kc_d, vc_d, c["num_kv_heads"], c["scale"], bt, sl, c["block_size"],
                               c["max_seq_len"], alibi, kv_dtype, kv_scale)
    ref = ops_ref.paged_attention(c["query"], kc_ref, vc_ref, c["block_tables"], c["seq_lens"], c["scale"],
                                  c["alibi_slopes"], kv_dtype, kv_scale)
    return out.cpu(), ref


@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("heads", [(8, 8), (16, 2), (8, 1), (12, 4)])
@pytest.mark.parametrize("head_size", [64, 80, 128, 256])
@pytest.mark.parametrize("block_size", [8, 16, 32])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
def test_paged_attention_shapes(ops, version, heads, head_size, block_size, dtype):
    if dtype == torch.float32 and (head_size in (80, 256) or heads == (12, 4)):
        pytest.skip("fp32 covered on the other shapes")
    H, Hkv = heads
    c = fixtures.attention_case(5, H, Hkv, head_size, block_size, 128, 700, dtype, seed=head_size + block_size)
    out, ref = run_case(ops, c, version)
    assert torch.allclose(out.float(), ref.float(), atol=1e-3, rtol=1e-5), (out.float() - ref.float()).abs().max()


@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("kind,kv_scale", [("fp8", 1.0), ("fp8", 0.0376), ("fp8_e5m2", 1.0)])
@pytest.mark.parametrize("heads", [(8, 2), (8, 1), (4, 4)])
def test_paged_attention_fp8_kv(ops, version, dtype, kind, kv_scale, heads):
    c = fixtures.attention_case(4, heads[0], heads[1], 128, 16, 96, 900, dtype, seed=21)
    out, ref = run_case(ops, c, version, kind, kv_scale)
    assert torch.allclose(out.float(), ref.float(), atol=1e-2, rtol=1e-5), (out.float() - ref.float()).abs().max()


@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("heads", [(8, 8), (16, 2)])
def test_paged_attention_alibi(ops, version, heads):
    c = fixtures.attention_case(6, heads[0], heads[1], 128, 16, 64, 600, torch.bfloat16, seed=9, use_alibi=True)
    out, ref = run_case(ops, c, version)
    assert torch.allclose(out.float(), ref.float(), atol=1e-3, rtol=1e-5), (out.float() - ref.float()).abs().max()


@pytest.mark.parametrize("version", ["v1", "v2"])
def test_paged_attention_ragged_and_empty(ops, version):
    # padded CUDA-graph rows have seq_len 0 and zeroed block tables: must not fault, output 0
    c = fixtures.attention_case(6, 8, 2, 128, 16, 64, 530, torch.bfloat16, seed=3)
    c["seq_lens"][0] = 0
    c["seq_lens"][1] = 1
    c["seq_lens"][2] = 512
    c["seq_lens"][3] = 513
    c["block_tables"][0] = 0
    out, ref = run_case(ops, c, version)
    assert torch.all(out[0] == 0)
    assert torch.allclose(out.float(), ref.float(), atol=1e-3, rtol=1e-5)


@pytest.mark.parametrize("version", ["v1", "v2"])
def test_paged_attention_long_context(ops, version):
    # 8k context (v1's upper bound in paged_attn.py:120) and GQA 8 like Llama-3-70B TP8
    c = fixtures.attention_case(2, 8, 1, 128, 16, 1200, 8192, torch.bfloat16, seed=13, fixed_len=8192)
    out, ref = run_case(ops, c, version)
    assert torch.allclose(out.float(), ref.float(), atol=1e-3, rtol=1e-5), (out.float() - ref.float()).abs().max()


def test_paged_attention_strided_query(ops):
    c = fixtures.attention_case(4, 8, 2, 128, 16, 64, 300, torch.float16, seed=17)
    S, H, D = c["query"].shape
    qkv = torch.zeros((S, (H + 4) * D), dtype=torch.float16, device=DEV)
    qkv[:, :H * D] = c["query"].reshape(S, H * D).to(DEV)
    qv = qkv[:, :H * D].view(S, H, D)   # stride(0) = (H+4)*D
    out = torch.empty((S, H, D), dtype=torch.float16, device=DEV)
    ops.paged_attention_v1(out, qv, c["key_cache"].to(DEV), c["value_cache"].to(DEV), 2, c["scale"],
                           c["block_tables"].to(DEV), c["seq_lens"].to(DEV), 16, c["max_seq_len"], None, "auto", 1.0)
    ref = ops_ref.paged_attention(c["query"], c["key_cache"], c["value_cache"], c["block_tables"], c["seq_lens"], c["scale"])
    assert torch.allclose(out.cpu().float(), ref.float(), atol=1e-3, rtol=1e-5)


@pytest.mark.parametrize("fname", golden_files("attn_"))
@pytest.mark.parametrize("version", ["v1", "v2"])
def test_paged_attention_vs_reference_cpu_op_golden(ops, fname, version):
    g = golden(fname)
    S, H, Hkv, D, BS, NB, max_len = (int(v) for v in g["params"])
    dts = str(g["dtype"])
    c = fixtures.attention_case(S, H, Hkv, D, BS, NB, max_len, getattr(torch, dts), seed=int(g["seed"]),
                                use_alibi=bool(g["alibi"]))
    out, _ = run_case(ops, c, version)
    ref = as_torch(g["out_" + version], dts)
    assert torch.allclose(out.float(), ref.float(), atol=1e-3, rtol=1e-5), (out.float() - ref.float()).abs().max()


def test_unsupported_head_size_raises(ops):
    q = torch.zeros((1, 4, 72), dtype=torch.float16, device=DEV)
    kc = torch.zeros((4, 4, 9, 16, 8), dtype=torch.float16, device=DEV)
    vc = torch.zeros((4, 4, 72, 16), dtype=torch.float16, device=DEV)
    with pytest.raises(RuntimeError, match="Unsupported head size"):
        ops.paged_attention_v1(torch.empty_like(q), q, kc, vc, 4, 1.0, torch.zeros((1, 1), dtype=torch.int32, device=DEV),
                               torch.ones(1, dtype=torch.int32, device=DEV), 16, 1, None, "auto", 1.0)
