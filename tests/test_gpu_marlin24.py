"""GPU parity: gptq_marlin_24_gemm (Sparse-Marlin 2:4, mma.sp fed with the compressed operand) vs the oracle.
Mirrors test_marlin_24_gemm of the reference (tests/kernels/test_marlin_gemm.py:182-227): MNK factors x 128/256 chunks,
group sizes {-1, 128}, fp16; tolerance mean|out-ref|/mean|ref| < 0.04 (reference) and <= 1e-2 vs the fp32 oracle.
A unit-scale / integer-activation case is exact, which pins the kernel's reading of b_meta and of the tile permutation."""
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


def _run(ops, a, q, meta, s, bits, M, N, K):
    ws = torch.zeros((N // 128) * 64, dtype=torch.int32, device=DEV)
    out = ops.gptq_marlin_24_gemm(a.to(DEV), q.to(DEV), meta.to(DEV), s.to(DEV), ws, bits, M, N, K)
    torch.cuda.synchronize()
    assert torch.all(ws == 0), "workspace must be left zeroed"
    return out.cpu()


def _case(ops, M, K, N, bits, group, seed=0):
    w = fixtures.gemm_weights(K, N, seed=seed, dtype=torch.float16)
    a = fixtures.gemm_acts(M, K, seed=seed, dtype=torch.float16)
    w_ref, q, meta, s = ql.marlin_24_quantize(w, bits, group)
    out = _run(ops, a, q, meta, s.to(torch.float16), bits, M, N, K)
    return out, ops_ref.marlin_gemm_ref(a, w_ref.to(torch.float16))


@pytest.mark.parametrize("mnk", MNK_FACTORS)
@pytest.mark.parametrize("chunk", [(128, 128), (256, 128), (128, 256)])
@pytest.mark.parametrize("group", [-1, 128])
def test_marlin24_gemm_4bit(ops, mnk, chunk, group):
    mf, nf, kf = mnk
    M, N, K = mf, chunk[1] * nf, chunk[0] * kf
    out, ref = _case(ops, M, K, N, 4, group, seed=M + N + K)
    err = ops_ref.rel_mean_err(out, ref)
    assert err < 0.04, err
    assert err < 1e-2, err


@pytest.mark.parametrize("mnk", [(1, 4, 8), (13, 17, 67), (67, 13, 11)])
@pytest.mark.parametrize("group", [-1, 128])
def test_marlin24_gemm_8bit(ops, mnk, group):
    mf, nf, kf = mnk
    M, N, K = mf, 128 * nf, 128 * kf
    out, ref = _case(ops, M, K, N, 8, group, seed=M + N + K)
    assert ops_ref.rel_mean_err(out, ref) < 1e-2


@pytest.mark.parametrize("M", [1, 2, 8, 16, 17, 32, 33, 64, 65, 128, 200])
def test_marlin24_batch_sweep(ops, M):
    out, ref = _case(ops, M, 4096, 1024, 4, 128, seed=M)
    assert ops_ref.rel_mean_err(out, ref) < 1e-2


@pytest.mark.parametrize("bits", [4, 8])
@pytest.mark.parametrize("group", [-1, 128])
def test_marlin24_exact_integers(ops, bits, group):
    """weights that quantise exactly (integers, scale 1) x small-integer activations: every product and partial sum is an
    integer below 2048, so any mistake in the metadata / permutation / split-K reduction shows up as a non-zero diff."""
    M, K, N = 8, 512, 256
    g = torch.Generator().manual_seed(7 + bits)
    half = (1 << bits) // 2
    w = torch.randint(-(half - 1), half, (K, N), generator=g).to(torch.float16)
    # pin max|w| per scale group to half-1... scale = max|w| * 2 / (2^bits - 1)-ish is not 1: build the tensors directly
    w24, _ = ql.inject_24(w)
    q = (w24 + half).to(torch.int32)
    comp, meta = ql.compress_quantized_24_weight(q, K, N, bits)
    perm, sp, sps = ql.marlin24_perm_tables(bits)
    qp = ql.marlin_pack(comp, bits, perm)
    ngroups = 1 if group == -1 else K // group
    s = torch.ones((ngroups, N), dtype=torch.float16)
    a = torch.randint(-2, 3, (M, K), generator=g).to(torch.float16)
    if bits == 8:
        a = (a != 0).to(torch.float16) * (torch.randint(0, 2, (M, K), generator=g) * 2 - 1).to(torch.float16) * \
            (torch.rand((M, K), generator=g) < 0.05).to(torch.float16)
    ref = a.double() @ w24.double()
    assert float(ref.abs().max()) < 2048
    out = _run(ops, a, qp, meta, s, bits, M, N, K)
    assert torch.equal(out.double(), ref)


@pytest.mark.parametrize("fname", golden_files("marlin24_"))
def test_marlin24_golden(ops, fname):
    g = golden(fname)
    K, N, bits = int(g["K"]), int(g["N"]), int(g["bits"])
    a = fixtures.gemm_acts(16, K, seed=1, dtype=torch.float16)
    out = _run(ops, a, as_torch(g["q_comp"]), as_torch(g["meta"]), as_torch(g["s24"]), bits, 16, N, K)
    ref = ops_ref.marlin_gemm_ref(a, as_torch(g["w24_ref"]))
    assert ops_ref.rel_mean_err(out, ref) < 1e-2


def test_marlin24_deterministic(ops):
    M, K, N = 16, 4096, 512
    w = fixtures.gemm_weights(K, N, seed=3, dtype=torch.float16)
    a = fixtures.gemm_acts(M, K, seed=3, dtype=torch.float16)
    _, q, meta, s = ql.marlin_24_quantize(w, 4, 128)
    outs = [_run(ops, a, q, meta, s.to(torch.float16), 4, M, N, K) for _ in range(4)]
    for o in outs[1:]:
        assert torch.equal(o, outs[0])


def test_marlin24_rejects(ops):
    a = torch.zeros((1, 128), dtype=torch.bfloat16, device=DEV)
    q = torch.zeros((4, 256), dtype=torch.int32, device=DEV)
    meta = torch.zeros((4, 256), dtype=torch.int16, device=DEV)
    s = torch.ones((1, 128), dtype=torch.bfloat16, device=DEV)
    ws = torch.zeros(64, dtype=torch.int32, device=DEV)
    with pytest.raises(RuntimeError):
        ops.gptq_marlin_24_gemm(a, q, meta, s, ws, 4, 1, 128, 128)   # bf16 activations are rejected like the reference
