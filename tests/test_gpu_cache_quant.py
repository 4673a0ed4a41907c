"""GPU parity: KV-cache write path, fp8 cache conversion, activation quantizers -- CUDA (through the torch shim -> C ABI)
vs the oracle restatements and the golden vectors produced by the reference's own CPU operators."""
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


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("shape", [(7, 2, 128, 16, 16), (42, 8, 64, 32, 40), (3, 1, 80, 8, 9), (64, 8, 128, 16, 256)])
def test_reshape_and_cache_auto_bit_exact(ops, dtype, shape):
    T, H, D, BS, NB = shape
    c = fixtures.cache_write_case(T, H, D, BS, NB, dtype, seed=11)
    kc, vc = c["key_cache"].to(DEV), c["value_cache"].to(DEV)
    qkv = torch.stack([c["key"], c["key"], c["value"]], dim=1).to(DEV)  # key/value as strided slices
    ops.reshape_and_cache(qkv[:, 1], qkv[:, 2], kc, vc, c["slot_mapping"].to(DEV), "auto", 1.0)
    ops_ref.reshape_and_cache(c["key"], c["value"], c["key_cache"], c["value_cache"], c["slot_mapping"])
    assert torch.equal(kc.cpu(), c["key_cache"])
    assert torch.equal(vc.cpu(), c["value_cache"])


@pytest.mark.parametrize("fname", golden_files("cache_"))
def test_reshape_and_cache_golden(ops, fname):
    g = golden(fname)
    T, Hkv, D, BS, NB = (int(v) for v in g["params"])
    dts = str(g["dtype"])
    c = fixtures.cache_write_case(T, Hkv, D, BS, NB, getattr(torch, dts), seed=int(g["seed"]))
    kc, vc = c["key_cache"].to(DEV), c["value_cache"].to(DEV)
    ops.reshape_and_cache(c["key"].to(DEV), c["value"].to(DEV), kc, vc, c["slot_mapping"].to(DEV), "auto", 1.0)
    assert torch.equal(kc.cpu(), as_torch(g["key_cache"], dts))
    assert torch.equal(vc.cpu(), as_torch(g["value_cache"], dts))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("kind", ["fp8", "fp8_e5m2"])
@pytest.mark.parametrize("kv_scale", [1.0, 0.0376])
def test_reshape_and_cache_fp8(ops, dtype, kind, kv_scale):
    T, H, D, BS, NB = 33, 4, 128, 16, 24
    c = fixtures.cache_write_case(T, H, D, BS, NB, dtype, seed=5)
    x = 16
    kc = torch.zeros((NB, H, D // x, BS, x), dtype=torch.uint8, device=DEV)
    vc = torch.zeros((NB, H, D, BS), dtype=torch.uint8, device=DEV)
    ops.reshape_and_cache(c["key"].to(DEV), c["value"].to(DEV), kc, vc, c["slot_mapping"].to(DEV), kind, kv_scale)
    kc_ref = torch.zeros_like(kc, device="cpu")
    vc_ref = torch.zeros_like(vc, device="cpu")
    ops_ref.reshape_and_cache(c["key"], c["value"], kc_ref, vc_ref, c["slot_mapping"], kind, kv_scale)
    # same rounding rule on both sides (RN-even, SATFINITE) -> identical bytes
    assert torch.equal(kc.cpu(), kc_ref)
    assert torch.equal(vc.cpu(), vc_ref)
    if kind == "fp8":
        # the reference test's own check (tests/kernels/test_cache.py:191-202): dequantise and compare loosely
        back = torch.empty(kc.shape, dtype=dtype, device=DEV)
        ops.convert_fp8(back, kc, kv_scale, kind)
        want = ops_ref.fp8_dequant(kc_ref, kv_scale, dtype, kind)
        assert torch.allclose(back.cpu().float(), want.float(), atol=1e-3, rtol=0.1)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
def test_convert_fp8_round_trip(ops, dtype):
    # tests/kernels/test_cache.py:369-394: values in +-224, round trip within atol 1e-3 / rtol 0.1
    g = torch.Generator().manual_seed(3)
    src = ((torch.rand((37, 8, 16, 16, 8), generator=g) * 2 - 1) * 224).to(dtype).to(DEV)
    q = torch.empty(src.shape, dtype=torch.uint8, device=DEV)
    ops.convert_fp8(q, src)
    back = torch.empty_like(src)
    ops.convert_fp8(back, q)
    assert torch.allclose(src.float(), back.float(), atol=1e-3, rtol=0.1)
    assert torch.equal(q.cpu(), ops_ref.fp8_quant(src.cpu(), 1.0))
    assert torch.equal(back.cpu(), ops_ref.fp8_dequant(q.cpu(), 1.0, dtype))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("shape", [(1, 17), (64, 8192), (7, 1024), (33, 3584), (0, 128)])
def test_scaled_fp8_quant(ops, dtype, shape):
    g = torch.Generator().manual_seed(1)
    x = (torch.randn(shape, generator=g) * 3).to(dtype)
    q, s = ops.scaled_fp8_quant(x.to(DEV))
    assert q.dtype == torch.float8_e4m3fn and s.shape == (1,)
    if x.numel():
        q_ref, s_ref = ops_ref.scaled_fp8_quant(x)
        assert torch.allclose(s.cpu(), s_ref, rtol=1e-6)
        # tests/quantization/test_fp8.py:58-94: dequantised values close
        assert torch.allclose(q.cpu().float() * s.cpu(), q_ref.float() * s_ref, atol=float(s_ref) * 0.51 * 32, rtol=0.13)
        mism = (q.cpu().view(torch.uint8) != q_ref.view(torch.uint8)).float().mean().item()
        assert mism < 1e-3, mism  # 1/scale vs division may flip a rare tie
        sc = torch.tensor([0.05], dtype=torch.float32)
        q2, s2 = ops.scaled_fp8_quant(x.to(DEV), sc.to(DEV))
        assert torch.equal(q2.cpu().view(torch.uint8), ops_ref.scaled_fp8_quant(x, sc)[0].view(torch.uint8))
    q3, _ = ops.scaled_fp8_quant(x.to(DEV), batch_dim_padding=17)
    assert q3.shape[0] == max(17, shape[0])


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("shape", [(1, 16), (7, 67), (64, 8192), (83, 5120)])
def test_scaled_int8_quant(ops, dtype, shape):
    g = torch.Generator().manual_seed(2)
    x = (torch.rand(shape, generator=g) * 1000 - 300).to(dtype)
    q, s = ops.scaled_int8_quant(x.to(DEV))
    q_ref, s_ref = ops_ref.scaled_int8_quant(x)
    assert torch.allclose(s.cpu(), s_ref, rtol=1e-6)
    assert (q.cpu().int() - q_ref.int()).abs().max() <= 1      # tests/kernels/test_int8_quant.py:44-46
    assert (q.cpu() != q_ref).float().mean() < 5e-3
    sc = torch.tensor([2.1], dtype=torch.float32)
    q2, _ = ops.scaled_int8_quant(x.to(DEV), sc.to(DEV))
    assert (q2.cpu().int() - ops_ref.scaled_int8_quant(x, sc)[0].int()).abs().max() <= 1
    assert (q2.cpu() != ops_ref.scaled_int8_quant(x, sc)[0]).float().mean() < 5e-3


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
def test_glue_kernels(ops, dtype):
    g = torch.Generator().manual_seed(4)
    T, Hd = 9, 4096
    x = torch.randn((T, Hd), generator=g).to(dtype)
    res = torch.randn((T, Hd), generator=g).to(dtype)
    w = (torch.randn(Hd, generator=g) * 0.1 + 1).to(dtype)
    tol = dict(atol=2e-2, rtol=2e-2) if dtype != torch.float32 else dict(atol=1e-5, rtol=1e-5)
    out = torch.empty((T, Hd), dtype=dtype, device=DEV)
    ops.rms_norm(out, x.to(DEV), w.to(DEV), 1e-5)
    assert torch.allclose(out.cpu().float(), ops_ref.rms_norm(x, w, 1e-5).float(), **tol)
    xi, ri = x.to(DEV).clone(), res.to(DEV).clone()
    ops.fused_add_rms_norm(xi, ri, w.to(DEV), 1e-5)
    want, wres = ops_ref.fused_add_rms_norm(x, res, w, 1e-5)
    assert torch.equal(ri.cpu(), wres)
    assert torch.allclose(xi.cpu().float(), want.float(), **tol)
    gate = torch.randn((T, 2 * 1024), generator=g).to(dtype)
    o2 = torch.empty((T, 1024), dtype=dtype, device=DEV)
    ops.silu_and_mul(o2, gate.to(DEV))
    assert torch.allclose(o2.cpu().float(), ops_ref.silu_and_mul(gate).float(), **tol)
    # rotary, neox + gptj
    Hq, Hk, D = 8, 2, 128
    cache = torch.randn((64, D), generator=g).to(dtype)
    pos = torch.randint(0, 64, (T,), generator=g)
    for neox in (True, False):
        q = torch.randn((T, Hq * D), generator=g).to(dtype)
        k = torch.randn((T, Hk * D), generator=g).to(dtype)
        qd, kd = q.to(DEV).clone(), k.to(DEV).clone()
        ops.rotary_embedding(pos.to(DEV), qd, kd, D, cache.to(DEV), neox)
        wq, wk = ops_ref.rotary_embedding(pos, q, k, D, cache, neox)
        assert torch.allclose(qd.cpu().float(), wq.float(), **tol)
        assert torch.allclose(kd.cpu().float(), wk.float(), **tol)


def _first_max(x):
    """lowest index of the row maximum (the tie rule of torch.argmax's CUDA reduction), computed without argmax"""
    xf = x.float()
    eq = xf == xf.max(dim=-1, keepdim=True).values
    idx = torch.arange(x.shape[1], device=x.device).expand_as(xf)
    return torch.where(eq, idx, x.shape[1]).min(dim=-1).values


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("shape", [(64, 128256), (1, 32000), (7, 1000), (3, 8), (5, 16032), (2, 50257)])
def test_greedy_sample_matches_argmax(ops, dtype, shape):
    g = torch.Generator(device=DEV).manual_seed(3)
    x = torch.randn(shape, generator=g, device=DEV).to(dtype)
    assert torch.equal(ops.greedy_sample(x), _first_max(x))
    # ties: few distinct values -> the lowest index of the maximum must win
    q = torch.randint(0, 4, shape, generator=g, device=DEV).to(dtype)
    assert torch.equal(ops.greedy_sample(q), _first_max(q))
    # strided rows (a slice of a wider logits buffer, 16-byte aligned or not)
    wide = torch.randn((shape[0], shape[1] + 24), generator=g, device=DEV).to(dtype)
    for off in (8, 3):
        assert torch.equal(ops.greedy_sample(wide[:, off:off + shape[1]]), _first_max(wide[:, off:off + shape[1]]))


def test_decode_step_tokens_are_first_max_of_logits(ops, monkeypatch):
    """The engine's sampled ids are the lowest-index maxima of the logits it computed (eager step and graph replay)."""
    from nm_vllm_b200 import decoder
    seen = {}
    orig = ops.greedy_sample

    def capture(logits):
        if not torch.cuda.is_current_stream_capturing():
            seen["logits"] = logits.clone()
        return orig(logits)

    monkeypatch.setattr(decoder.ops, "greedy_sample", capture)
    model = decoder.SyntheticLlama(decoder.TINY, device=DEV, num_blocks=64, max_pos=256)
    eng = decoder.DecodeEngine(model, batch=4, max_ctx=96)
    hi = eng.host_inputs(80, seed=0)
    eng.load_inputs(hi)
    eng._device_step(80)
    torch.cuda.synchronize()
    want = _first_max(seen["logits"])
    assert torch.equal(eng.d_out, want)
    assert torch.equal(eng.step(hi).to(DEV), want)   # CUDA-graph replay of the same step
    # informational: where torch.argmax's tie choice differs, both point at equal logits
    ta = torch.argmax(seen["logits"], dim=-1)
    rows = torch.arange(4, device=DEV)
    assert torch.equal(seen["logits"][rows, ta], seen["logits"][rows, want])
