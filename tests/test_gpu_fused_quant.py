"""GPU parity of the glue kernels fused with the activation quantiser (csrc/glue_quant.cu) against the chain of the two
reference ops on the same inputs -- bit-identical outputs, scales and residuals -- and against the oracle restatements; plus
a decode step with the fused path on vs off."""
import pytest
import torch

from oracle import ops_ref

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _unfused_quant(ops, y, kind, scale):
    if kind == "fp8":
        return ops.scaled_fp8_quant(y, scale)
    if kind == "int8":
        return ops.scaled_int8_quant(y, scale)
    return ops.scaled_int8_quant(y, None)


@pytest.mark.parametrize("kind", ["fp8", "int8", "int8_token"])
@pytest.mark.parametrize("hidden", [512, 4096, 8192])
@pytest.mark.parametrize("with_residual", [False, True])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_rms_norm_quant_equals_two_ops(built_extension, kind, hidden, with_residual, dtype):
    from nm_vllm_b200 import _custom_ops as ops
    T = 37
    g = torch.Generator(device=DEV).manual_seed(hidden + T)
    x = torch.randn((T, hidden), generator=g, device=DEV).to(dtype)
    res = torch.randn((T, hidden), generator=g, device=DEV).to(dtype)
    w = (torch.rand(hidden, generator=g, device=DEV) + 0.5).to(dtype)
    scale = torch.tensor([0.02 if kind == "fp8" else 0.03], dtype=torch.float32, device=DEV)
    # two ops
    x2, r2 = x.clone(), res.clone()
    if with_residual:
        ops.fused_add_rms_norm(x2, r2, w, 1e-5)
        y = x2
    else:
        y = torch.empty_like(x)
        ops.rms_norm(y, x2, w, 1e-5)
    q_ref, s_ref = _unfused_quant(ops, y, kind, scale if kind != "int8_token" else None)
    # fused
    r1 = res.clone()
    q, s = ops.rms_norm_quant(x.clone(), w, 1e-5, kind, scale if kind != "int8_token" else None, r1 if with_residual else None)
    assert torch.equal(q.view(torch.uint8), q_ref.view(torch.uint8))
    assert torch.equal(s.flatten(), s_ref.flatten())
    if with_residual:
        assert torch.equal(r1, r2)
    # oracle chain (CPU restatement): norm, then the quantiser formula
    if with_residual:
        y_o, _ = ops_ref.fused_add_rms_norm(x.cpu(), res.cpu(), w.cpu(), 1e-5)
    else:
        y_o = ops_ref.rms_norm(x.cpu(), w.cpu(), 1e-5)
    if kind == "fp8":
        qo, _ = ops_ref.scaled_fp8_quant(y_o, scale.cpu())
        a, b = q.cpu().float(), qo.float()
        assert float((a - b).abs().max()) <= 0.125 * float(b.abs().max()) + 1e-6    # one e4m3 step where the norm differs by an ulp
        assert float(((a - b).abs() > 0).float().mean()) < 0.02
    else:
        qo, so = ops_ref.scaled_int8_quant(y_o, scale.cpu() if kind == "int8" else None)
        assert int((q.cpu().int() - qo.int()).abs().max()) <= 1                      # the reference's +-1 LSB (test_int8_quant.py:44-46)


@pytest.mark.parametrize("kind", ["fp8", "int8", "int8_token"])
@pytest.mark.parametrize("d", [1024, 14336])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_silu_and_mul_quant_equals_two_ops(built_extension, kind, d, dtype):
    from nm_vllm_b200 import _custom_ops as ops
    T = 19
    g = torch.Generator(device=DEV).manual_seed(d)
    x = torch.randn((T, 2 * d), generator=g, device=DEV).to(dtype)
    scale = torch.tensor([0.02 if kind == "fp8" else 0.03], dtype=torch.float32, device=DEV)
    y = torch.empty((T, d), dtype=dtype, device=DEV)
    ops.silu_and_mul(y, x)
    q_ref, s_ref = _unfused_quant(ops, y, kind, scale if kind != "int8_token" else None)
    q, s = ops.silu_and_mul_quant(x, kind, scale if kind != "int8_token" else None)
    assert torch.equal(q.view(torch.uint8), q_ref.view(torch.uint8))
    assert torch.equal(s.flatten(), s_ref.flatten())


@pytest.mark.parametrize("quant", ["fp8_static", "int8", "fp8"])
def test_decode_step_fused_quant_equals_unfused(built_extension, quant, monkeypatch):
    """the W8A8 decode step with the quantiser fused into norm / silu produces the same logits as the op-by-op chain"""
    from nm_vllm_b200 import decoder
    ops = decoder.ops
    seen = []
    orig = ops.greedy_sample

    def capture(logits):
        seen.append(logits.float().cpu())
        return orig(logits)

    monkeypatch.setattr(ops, "greedy_sample", capture)
    m = decoder.SyntheticLlama(decoder.MINI, quant=quant, device=DEV, num_blocks=64, max_pos=256)
    eng = decoder.DecodeEngine(m, batch=5, max_ctx=96, use_graph=False)
    hi = eng.host_inputs(40, seed=3)
    kv0 = [(k.clone(), v.clone()) for k, v in m.kv]
    for fused in (True, False):
        for (k, v), (k0, v0) in zip(m.kv, kv0):
            k.copy_(k0); v.copy_(v0)
        m.fuse_act_quant = fused
        m.fuse_dynamic_fp8 = True    # quant "fp8": the opt-in dynamic per-tensor fusion
        eng.load_inputs(hi)
        eng._device_step(40)
        torch.cuda.synchronize()
    assert torch.equal(seen[0], seen[1])
    assert not torch.isnan(seen[0]).any()


@pytest.mark.parametrize("hidden,T", [(4096, 64), (8192, 1), (512, 300)])
@pytest.mark.parametrize("with_residual", [False, True])
def test_rms_norm_dynamic_fp8_quant_equals_three_kernels(built_extension, hidden, T, with_residual):
    """kind "fp8_dynamic": norm + the dynamic PER-TENSOR fp8 quantiser (absmax kernel + convert kernel in the reference flow,
    fp8/common.cu:39-72) in ONE launch -- same bytes, same scale; called three times in a row with different data (the
    grid-wide maximum lives in a static device word that every call must leave zero)."""
    from nm_vllm_b200 import _custom_ops as ops
    for rep in range(3):
        g = torch.Generator(device=DEV).manual_seed(hidden + T + rep)
        x = (torch.randn((T, hidden), generator=g, device=DEV) * (1.0 + 3.0 * (2 - rep))).to(torch.bfloat16)   # shrinking maxima
        res = torch.randn((T, hidden), generator=g, device=DEV).to(torch.bfloat16)
        w = (torch.rand(hidden, generator=g, device=DEV) + 0.5).to(torch.bfloat16)
        x2, r2 = x.clone(), res.clone()
        if with_residual:
            ops.fused_add_rms_norm(x2, r2, w, 1e-5)
            y = x2
        else:
            y = torch.empty_like(x)
            ops.rms_norm(y, x2, w, 1e-5)
        q_ref, s_ref = ops.scaled_fp8_quant(y, None)
        r1 = res.clone()
        q, s = ops.rms_norm_quant(x.clone(), w, 1e-5, "fp8_dynamic", None, r1 if with_residual else None)
        assert torch.equal(s.flatten(), s_ref.flatten()), "rep %d: scale %r vs %r" % (rep, s, s_ref)
        assert torch.equal(q.view(torch.uint8), q_ref.view(torch.uint8))
        if with_residual:
            assert torch.equal(r1, r2)


@pytest.mark.parametrize("d,T", [(14336, 64), (1024, 7)])
def test_silu_and_mul_dynamic_fp8_quant_equals_three_kernels(built_extension, d, T):
    from nm_vllm_b200 import _custom_ops as ops
    for rep in range(2):
        g = torch.Generator(device=DEV).manual_seed(d + rep)
        x = (torch.randn((T, 2 * d), generator=g, device=DEV) * (2.0 - rep)).to(torch.bfloat16)
        y = torch.empty((T, d), dtype=torch.bfloat16, device=DEV)
        ops.silu_and_mul(y, x)
        q_ref, s_ref = ops.scaled_fp8_quant(y, None)
        q, s = ops.silu_and_mul_quant(x, "fp8_dynamic", None)
        assert torch.equal(s.flatten(), s_ref.flatten())
        assert torch.equal(q.view(torch.uint8), q_ref.view(torch.uint8))


def test_dynamic_fp8_quant_rejects_too_many_tokens_and_replays_in_a_graph(built_extension):
    from nm_vllm_b200 import _custom_ops as ops
    w = torch.ones(512, dtype=torch.bfloat16, device=DEV)
    with pytest.raises(RuntimeError):
        ops.rms_norm_quant(torch.zeros((ops.FUSED_DYNAMIC_FP8_MAX_TOKENS + 1, 512), dtype=torch.bfloat16, device=DEV), w, 1e-5,
                           "fp8_dynamic")
    x = torch.randn((64, 512), device=DEV).to(torch.bfloat16)
    q0, s0 = ops.rms_norm_quant(x, w, 1e-5, "fp8_dynamic")
    torch.cuda.synchronize()
    st = torch.cuda.Stream()
    st.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(st):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            q1, s1 = ops.rms_norm_quant(x, w, 1e-5, "fp8_dynamic")
        for _ in range(3):
            g.replay()
    torch.cuda.synchronize()
    assert torch.equal(s0, s1) and torch.equal(q0.view(torch.uint8), q1.view(torch.uint8))
