"""GPU parity: cutlass_scaled_mm (W8A8 FP8 e4m3 / INT8, tcgen05 kind::f8f6f4 / kind::i8) vs the oracle's
baseline_scaled_mm.  Mirrors tests/kernels/test_cutlass.py of the reference: shapes, per-token / per-channel scales,
bias, output dtypes, the M sweep, strided subsets and CUDA-graph capture.  Tolerances are the reference's
(fp8: rtol 1e-2 / atol 5e-2, test_cutlass.py:79; int8: rtol 1e-1 / atol 1, :111); a scale-free INT8 case is bit-exact."""
import pytest
import torch

from oracle import ops_ref

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def ops(built_extension):
    from nm_vllm_b200 import _custom_ops
    return _custom_ops


def to_fp8(t):
    fi = torch.finfo(torch.float8_e4m3fn)
    return torch.round(t.clamp(min=fi.min, max=fi.max)).to(torch.float8_e4m3fn)


def to_int8(t):
    return torch.round(t.clamp(min=-128, max=127)).to(torch.int8)


def _case(ops, kind, m, n, k, per_token, per_channel, use_bias, out_dtype=torch.bfloat16, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed + m * 7 + n * 3 + k)
    if kind == "fp8":
        a = to_fp8(torch.randn((m, k), generator=g))
        b = to_fp8(torch.randn((n, k), generator=g)).t()
    else:
        a = to_int8(torch.randn((m, k), generator=g) * 5)
        b = to_int8(torch.randn((n, k), generator=g) * 5).t()
    sa = torch.randn((m if per_token else 1, 1), generator=g, dtype=torch.float32)
    sb = torch.randn((1, n if per_channel else 1), generator=g, dtype=torch.float32)
    bias = (torch.rand((n,), generator=g) * 10).to(out_dtype) if use_bias else None
    ref = ops_ref.scaled_mm(a, b, sa, sb, out_dtype, bias)
    bd = b.t().contiguous().to(DEV).t()  # column-major on the device
    out = ops.cutlass_scaled_mm(a.to(DEV), bd, sa.to(DEV), sb.to(DEV), out_dtype, None if bias is None else bias.to(DEV))
    assert out.shape == (m, n) and out.dtype == out_dtype
    if kind == "fp8":
        torch.testing.assert_close(out.cpu().float(), ref.float(), rtol=1e-2, atol=5e-2)
    else:
        torch.testing.assert_close(out.cpu().float(), ref.float(), rtol=1e-1, atol=1.0)
    return out


@pytest.mark.parametrize("kind", ["fp8", "int8"])
@pytest.mark.parametrize("m", [512, 222, 100, 33, 1])
@pytest.mark.parametrize("n", [2048, 256, 1024])
@pytest.mark.parametrize("k", [128, 496, 1024])
@pytest.mark.parametrize("scales", [(True, True), (False, False), (True, False), (False, True)])
def test_scaled_mm(ops, kind, m, n, k, scales):
    _case(ops, kind, m, n, k, scales[0], scales[1], use_bias=(m + n + k) % 2 == 0)


@pytest.mark.parametrize("kind", ["fp8", "int8"])
@pytest.mark.parametrize("out_dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("use_bias", [True, False])
def test_output_dtype(ops, kind, out_dtype, use_bias):
    _case(ops, kind, 512, 512, 512, True, True, use_bias, out_dtype)


@pytest.mark.parametrize("kind", ["fp8", "int8"])
@pytest.mark.parametrize("nk", [32, 64, 96])
def test_m_sweep(ops, kind, nk):
    for m in list(range(1, 20)) + [31, 32, 33, 63, 64, 65, 100, 127]:
        _case(ops, kind, m, nk, nk, True, True, m % 2 == 0)


@pytest.mark.parametrize("mnk", [(1, 1280, 8192), (16, 8192, 1024), (64, 7168, 8192), (48, 8192, 3584), (300, 640, 4096)])
def test_split_k_shapes(ops, mnk):
    """Llama-3-70B TP=8 decode shapes: few tiles, deep K -> the stream-K partial / last-arriver path."""
    m, n, k = mnk
    _case(ops, "fp8", m, n, k, True, True, True)
    _case(ops, "int8", m, n, k, False, True, False)


def test_int8_exact(ops):
    """unit scales, fp16-representable sums: the s32 accumulate + split-K partial sums are exact integers."""
    g = torch.Generator().manual_seed(3)
    m, n, k = 24, 384, 4096
    a = torch.randint(-3, 4, (m, k), generator=g, dtype=torch.int8)
    b = torch.randint(-3, 4, (n, k), generator=g, dtype=torch.int8)
    ref = a.to(torch.int32) @ b.to(torch.int32).t()
    assert int(ref.abs().max()) < 2048
    one = torch.ones((1, 1), dtype=torch.float32, device=DEV)
    out = ops.cutlass_scaled_mm(a.to(DEV), b.to(DEV).t(), one, one, torch.float16)
    assert torch.equal(out.cpu().to(torch.int32), ref)


def test_deterministic(ops):
    g = torch.Generator().manual_seed(5)
    a = to_fp8(torch.randn((8, 8192), generator=g)).to(DEV)
    b = to_fp8(torch.randn((1280, 8192), generator=g)).to(DEV).t()
    sa = torch.rand((8, 1), generator=g).to(DEV)
    sb = torch.rand((1, 1280), generator=g).to(DEV)
    outs = [ops.cutlass_scaled_mm(a, b, sa, sb, torch.bfloat16) for _ in range(5)]
    for o in outs[1:]:
        assert torch.equal(o, outs[0])


def test_subset(ops):
    """strided views of larger tensors (test_cutlass.py:250-272)."""
    g = torch.Generator().manual_seed(1)
    whole_a = to_int8(torch.randn((1024, 1024), generator=g) * 5)
    whole_b = to_int8(torch.randn((1024, 1024), generator=g) * 5).t()
    a, b = whole_a[0:512, 0:512], whole_b[0:512, 0:512]
    sa = torch.randn((1, 1), generator=g) / 10
    sb = torch.randn((1, 1), generator=g) / 10
    ref = ops_ref.scaled_mm(a, b, sa, sb, torch.bfloat16)
    wa, wb = whole_a.to(DEV), whole_b.t().contiguous().to(DEV).t()
    out = ops.cutlass_scaled_mm(wa[0:512, 0:512], wb[0:512, 0:512], sa.to(DEV), sb.to(DEV), torch.bfloat16)
    torch.testing.assert_close(out.cpu().float(), ref.float(), rtol=1e-1, atol=1.0)


@pytest.mark.parametrize("per_token", [True, False])
@pytest.mark.parametrize("per_channel", [True, False])
def test_cuda_graph(ops, per_token, per_channel):
    """test_cutlass.py:289-321"""
    m, n, k = 512, 512, 512
    g = torch.Generator().manual_seed(2)
    a = to_int8(torch.randn((m, k), generator=g)).to(DEV)
    b = to_int8(torch.randn((n, k), generator=g)).to(DEV).t()
    sa = (torch.randn((m if per_token else 1, 1), generator=g) / 10).to(DEV)
    sb = (torch.randn((1, n if per_channel else 1), generator=g) / 10).to(DEV)
    ops.cutlass_scaled_mm(a, b, sa, sb, torch.bfloat16)  # eager call reserves the split-K scratch
    stream = torch.cuda.Stream()
    with torch.cuda.stream(stream):
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            out = ops.cutlass_scaled_mm(a, b, sa, sb, torch.bfloat16)
    out.zero_()
    gr.replay()
    torch.cuda.synchronize()
    ref = torch.mm(sa * a.float(), sb * b.float()).to(torch.bfloat16)
    torch.testing.assert_close(out.float(), ref.float(), rtol=1e-1, atol=1.0)


def test_rejects_bad_layout(ops):
    a = torch.zeros((16, 64), dtype=torch.int8, device=DEV)
    b_rowmajor = torch.zeros((64, 32), dtype=torch.int8, device=DEV)
    one = torch.ones((1, 1), dtype=torch.float32, device=DEV)
    with pytest.raises(RuntimeError):
        ops.cutlass_scaled_mm(a, b_rowmajor, one, one, torch.bfloat16)
