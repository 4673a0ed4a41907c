"""GPU parity: gptq_marlin_repack (bit-exact) and the Marlin-layout W4A16 / W8A16 GEMM (tcgen05) vs the oracle.
Mirrors tests/kernels/test_marlin_gemm.py of the reference: MNK factors x chunks, group sizes, dtypes; tolerance
mean|out-ref|/mean|ref| < 0.04 (reference) and the north_star bound <= 1e-2 relative against the fp32 oracle."""
import pytest
import torch

from oracle import fixtures, ops_ref, quant_layouts as ql
from tests.util import as_torch, golden, golden_files

pytestmark = pytest.mark.gpu
DEV = "cuda:0"

MNK_FACTORS = [(1, 1, 1), (1, 4, 8), (1, 7, 5), (13, 17, 67), (26, 37, 13), (67, 13, 11)]


@pytest.fixture(scope="module")
def ops(built_extension):
    from nm_vllm_b200 import _custom_ops
    return _custom_ops


@pytest.mark.parametrize("bits", [4, 8])
@pytest.mark.parametrize("act_order", [False, True])
@pytest.mark.parametrize("kn", [(128, 64), (256, 192), (1024, 256), (640, 448)])
def test_repack_bit_exact(ops, bits, act_order, kn):
    K, N = kn
    w = fixtures.gemm_weights(K, N, seed=K + N + bits)
    _, q_w, s, g_idx, _ = ql.quantize_weights(w, bits, 64 if act_order else 128, act_order, seed=5)
    gptq = ql.gptq_pack(q_w, bits)
    sort_idx = torch.empty(0, dtype=torch.int32)
    if act_order:
        q_w, g_idx, sort_idx = ql.sort_weights(q_w, g_idx)
    want = ql.marlin_pack(q_w, bits)
    got = ops.gptq_marlin_repack(gptq.to(DEV), sort_idx.to(DEV), K, N, bits)
    assert torch.equal(got.cpu(), want)


@pytest.mark.parametrize("fname", golden_files("marlin_"))
def test_repack_golden(ops, fname):
    g = golden(fname)
    K, N, bits = int(g["K"]), int(g["N"]), int(g["bits"])
    got = ops.gptq_marlin_repack(as_torch(g["gptq"]).to(DEV), as_torch(g["sort_idx"]).to(DEV), K, N, bits)
    assert torch.equal(got.cpu(), as_torch(g["marlin_q"]))


def _gemm_case(ops, M, K, N, bits, group, dtype, seed=0, act_order=False):
    w = fixtures.gemm_weights(K, N, seed=seed, dtype=dtype)
    a = fixtures.gemm_acts(M, K, seed=seed, dtype=dtype)
    w_ref, mq, ms, g_idx, sort_idx, _ = ql.marlin_quantize(w, bits, group, act_order, seed=seed)
    g_idx = g_idx.to(torch.int32) if g_idx.numel() else torch.empty(0, dtype=torch.int32)
    ws = torch.zeros((N // 64) * 16, dtype=torch.int32, device=DEV)
    out = ops.gptq_marlin_gemm(a.to(DEV), mq.to(DEV), ms.to(dtype).to(DEV), g_idx.to(DEV), sort_idx.to(torch.int32).to(DEV), ws, bits,
                               M, N, K, True)
    torch.cuda.synchronize()
    assert torch.all(ws == 0), "workspace must be left zeroed"
    ref32 = ops_ref.marlin_gemm_ref(a, w_ref.to(dtype))
    return out.cpu(), ref32


@pytest.mark.parametrize("mnk", MNK_FACTORS)
@pytest.mark.parametrize("n_chunk", [64, 128, 256])
@pytest.mark.parametrize("group", [-1, 32, 64, 128])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_marlin_gemm_4bit(ops, mnk, n_chunk, group, dtype):
    mf, nf, kf = mnk
    M, N, K = mf, n_chunk * nf, 128 * kf
    out, ref = _gemm_case(ops, M, K, N, 4, group, dtype, seed=M + N + K)
    err = ops_ref.rel_mean_err(out, ref)
    assert err < 0.04, err          # reference's own bound
    assert err < 1e-2, err          # north_star bound vs fp32 dequant-then-matmul


@pytest.mark.parametrize("mnk", [(1, 4, 8), (13, 17, 67), (67, 13, 11)])
@pytest.mark.parametrize("group", [-1, 32, 128])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_marlin_gemm_8bit(ops, mnk, group, dtype):
    mf, nf, kf = mnk
    M, N, K = mf, 128 * nf, 128 * kf
    out, ref = _gemm_case(ops, M, K, N, 8, group, dtype, seed=M + N + K)
    err = ops_ref.rel_mean_err(out, ref)
    assert err < 1e-2, err


@pytest.mark.parametrize("M", [1, 2, 8, 16, 17, 32, 33, 64, 65, 128, 129, 200, 256, 300, 1024])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_marlin_gemm_llama_shapes_all_m(ops, M, dtype):
    # Llama-3-8B o_proj-like shape at every M bucket of the kernel (MPAD 16..256, multi-chunk)
    out, ref = _gemm_case(ops, M, 4096, 4096 if M <= 64 else 1024, 4, 128, dtype, seed=M)
    assert ops_ref.rel_mean_err(out, ref) < 1e-2
    assert torch.isfinite(out.float()).all()


@pytest.mark.parametrize("mkn", [(1, 256, 128), (13, 1024, 256), (67, 2048, 192), (16, 384, 64)])
@pytest.mark.parametrize("group", [32, 64, 128])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_marlin_gemm_act_order(ops, mkn, group, dtype):
    # act-order (desc_act) checkpoints: rows sorted by group at repack time, activation columns gathered by `perm`
    # (tests/kernels/test_marlin_gemm.py ACT_ORDER_OPTS; gptq_marlin.cu:283-339 permute_cols)
    M, K, N = mkn
    out, ref = _gemm_case(ops, M, K, N, 4, group, dtype, seed=M + K, act_order=True)
    err = ops_ref.rel_mean_err(out, ref)
    assert err < 1e-2, err


def test_marlin_gemm_k_not_multiple_of_stage(ops):
    # K = 128 * odd: the last 256-deep pipeline stage is half out of bounds (zero-filled by the tensor maps)
    for K in (128, 384, 640):
        for group in (-1, 128):
            out, ref = _gemm_case(ops, 5, K, 256, 4, group, torch.float16, seed=K)
            assert ops_ref.rel_mean_err(out, ref) < 1e-2


def test_marlin_gemm_deterministic_and_repeatable(ops):
    # split-K partials are summed in a fixed order and the tile counters reset themselves
    M, K, N = 16, 8192, 1024
    w = fixtures.gemm_weights(K, N, seed=1)
    a = fixtures.gemm_acts(M, K, seed=1).to(DEV)
    _, mq, ms, g_idx, sort_idx, _ = ql.marlin_quantize(w, 4, 128)
    ws = torch.zeros((N // 64) * 16, dtype=torch.int32, device=DEV)
    args = (mq.to(DEV), ms.to(DEV), g_idx.to(DEV), sort_idx.to(DEV), ws, 4, M, N, K, True)
    first = ops.gptq_marlin_gemm(a, *args)
    for _ in range(5):
        assert torch.equal(ops.gptq_marlin_gemm(a, *args), first)


def test_marlin_gemm_linearity(ops):
    # size-independent property at a BASELINE-sized layer: f(a1 + a2) ~= f(a1) + f(a2), f(0) == 0
    M, K, N = 8, 4096, 6144
    g = torch.Generator().manual_seed(0)
    mq = torch.randint(-2**31, 2**31 - 1, (K // 16, N * 2), generator=g, dtype=torch.int32).to(DEV)
    ms = (torch.rand((K // 128, N), generator=g) * 0.01 + 0.005).half().to(DEV)
    e = torch.empty(0, dtype=torch.int32, device=DEV)
    ws = torch.zeros((N // 64) * 16, dtype=torch.int32, device=DEV)
    a1 = torch.randn((M, K), generator=g).half().to(DEV)
    a2 = torch.randn((M, K), generator=g).half().to(DEV)
    f = lambda a: ops.gptq_marlin_gemm(a, mq, ms, e, e, ws, 4, M, N, K, True).float()
    assert torch.all(f(torch.zeros_like(a1)) == 0)
    lhs, rhs = f(a1 + a2), f(a1) + f(a2)
    assert (lhs - rhs).abs().mean() / rhs.abs().mean() < 5e-3


def test_marlin_gemm_alias(ops):
    M, K, N = 5, 512, 256
    w = fixtures.gemm_weights(K, N, seed=2)
    a = fixtures.gemm_acts(M, K, seed=2)
    w_ref, mq, ms, _, _, _ = ql.marlin_quantize(w, 4, 128)
    ws = torch.zeros((N // 64) * 16, dtype=torch.int32, device=DEV)
    out = ops.marlin_gemm(a.to(DEV), mq.to(DEV), ms.to(DEV), ws, M, N, K)
    assert ops_ref.rel_mean_err(out.cpu(), ops_ref.marlin_gemm_ref(a, w_ref)) < 1e-2


def test_cuda_graph_survives_scratch_regrow(ops):
    """a graph captured while the split-K scratch was small must keep replaying correctly after a later call made
    the scratch grow (the old buffers are retired, never freed)."""
    import ctypes

    from nm_vllm_b200 import capi
    M, K, N = 16, 4096, 4096          # split over the SMs: the kernel uses partial slots and tile counters
    g = torch.Generator().manual_seed(5)
    qw = torch.randint(-2**31, 2**31 - 1, (K // 16, N * 2), generator=g, dtype=torch.int32).to(DEV)
    s = (torch.rand((K // 128, N), generator=g) * 0.002 + 0.003).to(torch.bfloat16).to(DEV)
    a = torch.randn((M, K), generator=g).to(torch.bfloat16).to(DEV)
    e = torch.empty(0, dtype=torch.int32, device=DEV)
    ws = torch.zeros((N // 64) * 16, dtype=torch.int32, device=DEV)
    ref = ops.gptq_marlin_gemm(a, qw, s, e, e, ws, 4, M, N, K, True).clone()
    st = torch.cuda.Stream()
    st.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(st):
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=st):
            out = ops.gptq_marlin_gemm(a, qw, s, e, e, ws, 4, M, N, K, True)
    torch.cuda.synchronize()
    gr.replay()
    torch.cuda.synchronize()
    assert torch.equal(out, ref)
    # force the scratch to grow well beyond its first size, then use it from another kernel
    assert capi.lib().b200_gemm_scratch_reserve(ctypes.c_int64(1024), ctypes.c_int64(65536)) == 0, capi.last_error()
    big = torch.randn((256, K), generator=g).to(torch.bfloat16).to(DEV)
    ops.gptq_marlin_gemm(big, qw, s, e, e, ws, 4, 256, N, K, True)
    torch.cuda.synchronize()
    out.zero_()
    gr.replay()
    torch.cuda.synchronize()
    assert torch.equal(out, ref), "graph replay after a scratch regrow differs"
