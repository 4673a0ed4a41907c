"""GEMM parity AT THE BENCHMARKED SHAPES: every Llama-3-8B layer GEMM and the Llama-3-70B shards at
TP 2 / 4 / 8, M in {1, 16, 64}, for the W4A16 Marlin kernel (stream-K, cluster split-K and the KS = 56 `down` plan all
occur), the 2:4 kernel and W8A8 FP8 -- random packed bits dequantised by the oracle (`marlin_unpack`, as in
oracle/decode_ref._dense), fp32 matmul, mean relative error < 1e-2 AND an element-wise bound."""
import pytest
import torch

from oracle import ops_ref, quant_layouts as ql

pytestmark = pytest.mark.gpu
DEV = "cuda:0"

L8 = {"qkv": (4096, 6144), "o": (4096, 4096), "gate_up": (4096, 28672), "down": (14336, 4096)}


def l70(tp):
    return {"qkv": (8192, (64 + 16) * 128 // tp if tp <= 8 else 0), "o": (8192 // tp, 8192), "gate_up": (8192, 57344 // tp),
            "down": (28672 // tp, 8192)}


SHAPES = [("L8-" + n, kn) for n, kn in L8.items()] + [("L70tp%d-%s" % (tp, n), kn) for tp in (2, 4, 8) for n, kn in l70(tp).items()]


def _check(out, ref):
    err = ops_ref.rel_mean_err(out, ref)
    assert err < 1e-2, "mean relative error %g" % err
    # element-wise: |out - ref| <= 2 % of the output's rms + 2 % of |ref| (a wrong 128-wide tile would be off by ~100 %)
    rms = float(ref.float().pow(2).mean().sqrt())
    bad = (out.float() - ref.float()).abs() > 0.02 * rms + 0.02 * ref.float().abs()
    assert int(bad.sum()) == 0, "%d outputs beyond the element-wise bound" % int(bad.sum())


@pytest.mark.parametrize("name,kn", SHAPES, ids=[s[0] for s in SHAPES])
def test_w4a16_bench_shapes(built_extension, name, kn):
    from nm_vllm_b200 import _custom_ops as ops
    K, N = kn
    g = torch.Generator().manual_seed(K * 7 + N)
    qw = torch.randint(-2**31, 2**31 - 1, (K // 16, N * 2), generator=g, dtype=torch.int32)
    s = (torch.rand((K // 128, N), generator=g) * 0.002 + 0.003).to(torch.bfloat16)
    _, sp, sps = ql.marlin_perm_tables(4)
    w = (ql.marlin_unpack(qw, K, N, 4).float() - 8.0).view(K // 128, 128, N) * ql.unpermute_scales(s, K, N, 128, sp, sps).float().view(K // 128, 1, N)
    w = w.view(K, N)
    qd, sd = qw.to(DEV), s.to(DEV)
    e = torch.empty(0, dtype=torch.int32, device=DEV)
    ws = torch.zeros((N // 64) * 16, dtype=torch.int32, device=DEV)
    for M in (1, 16, 64):
        a = torch.randn((M, K), generator=g).to(torch.bfloat16)
        out = ops.gptq_marlin_gemm(a.to(DEV), qd, sd, e, e, ws, 4, M, N, K, True).cpu()
        _check(out, a.float() @ w)
    assert int(ws.abs().sum()) == 0


@pytest.mark.parametrize("name,kn", [s for s in SHAPES if s[0].startswith("L8")], ids=[s[0] for s in SHAPES if s[0].startswith("L8")])
def test_sparse_marlin_bench_shapes(built_extension, name, kn):
    from nm_vllm_b200 import _custom_ops as ops
    K, N = kn
    g = torch.Generator().manual_seed(K + N)
    w = (torch.randn((K, N), generator=g) * 0.02).to(torch.float16)
    w24_ref, qp, meta, s = ql.marlin_24_quantize(w, 4, 128)
    qd, md, sd = qp.to(DEV), meta.to(DEV), s.to(torch.float16).to(DEV)
    ws = torch.zeros((N // 128) * 64, dtype=torch.int32, device=DEV)
    for M in (1, 16, 64):
        a = torch.randn((M, K), generator=g).to(torch.float16)
        out = ops.gptq_marlin_24_gemm(a.to(DEV), qd, md, sd, ws, 4, M, N, K).cpu()
        _check(out, a.float() @ w24_ref.float())


@pytest.mark.parametrize("name,kn", SHAPES, ids=[s[0] for s in SHAPES])
def test_w8a8_fp8_bench_shapes(built_extension, name, kn):
    from nm_vllm_b200 import _custom_ops as ops
    K, N = kn
    g = torch.Generator().manual_seed(K * 3 + N)
    wq = (torch.randn((N, K), generator=g) * 40).clamp(-448, 448).to(torch.float8_e4m3fn)
    ws = torch.tensor([0.0005])
    wd = wq.to(DEV).t()
    for M in (1, 16, 64):
        x = torch.randn((M, K), generator=g).to(torch.bfloat16)
        xq, xs = ops.scaled_fp8_quant(x.to(DEV))
        out = ops.cutlass_scaled_mm(xq, wd, xs, ws.to(DEV), torch.bfloat16).cpu()
        ref = ops_ref.scaled_mm(xq.cpu(), wq.t(), xs.cpu(), ws, torch.bfloat16)
        torch.testing.assert_close(out.float(), ref.float(), rtol=1e-2, atol=5e-2)
