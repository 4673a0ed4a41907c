"""GPU: the allocation-free C-ABI entry point b200_gptq_marlin_24_gemm_static (K-split scratch in a file-scope device array)
gives the same result as b200_gptq_marlin_24_gemm, on two calls in a row, and rejects bad arguments before any launch.
Collected last on purpose (file name): it shares one static scratch array between calls, so it is kept away from the rest
of the suite.  It has no counterpart in the reference's tests."""
import ctypes

import pytest
import torch

from oracle import fixtures, ops_ref, quant_layouts as ql

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
B200_F16 = 0   # enum b200_dtype in include/b200_ops.h


@pytest.fixture(scope="module")
def lib(built_extension):
    from nm_vllm_b200 import capi
    return capi.lib()


def _call(lib, name, out, a, q, meta, s, ws, bits, M, N, K, groups):
    fn = getattr(lib, name)
    fn.restype = ctypes.c_int
    fn.argtypes = [ctypes.c_void_p] * 6 + [ctypes.c_int64, ctypes.c_int, ctypes.c_int64, ctypes.c_int64, ctypes.c_int64,
                                           ctypes.c_int64, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
    st = torch.cuda.current_stream(DEV).cuda_stream
    return fn(out.data_ptr(), a.data_ptr(), q.data_ptr(), meta.data_ptr(), s.data_ptr(), ws.data_ptr(), ws.numel(), bits,
              M, N, K, a.stride(0), groups, B200_F16, st)


def _problem(M, K, N, bits, group, seed):
    w = fixtures.gemm_weights(K, N, seed=seed, dtype=torch.float16)
    a = fixtures.gemm_acts(M, K, seed=seed, dtype=torch.float16)
    w_ref, q, meta, s = ql.marlin_24_quantize(w, bits, group)
    ref = ops_ref.marlin_gemm_ref(a, w_ref.to(torch.float16))
    return a.to(DEV), q.to(DEV).contiguous(), meta.to(DEV).contiguous(), s.to(torch.float16).to(DEV).contiguous(), ref


# (M, K, N, bits, group): K splits > 1 in the first four (scratch slots + counters in use), a single split in the last
CASES = [(16, 4096, 1024, 4, 128), (64, 4096, 1024, 4, 128), (1, 8192, 512, 8, -1), (33, 2048, 2048, 4, -1),
         (200, 512, 28672, 4, 128)]


@pytest.mark.parametrize("case", CASES)
def test_static_scratch_entry_matches_allocating_entry(lib, case):
    M, K, N, bits, group = case
    a, q, meta, s, ref = _problem(M, K, N, bits, group, seed=M + K + N)
    groups = s.shape[0]
    ws = torch.zeros((N // 128) * 64, dtype=torch.int32, device=DEV)
    base = torch.empty((M, N), dtype=torch.float16, device=DEV)
    rc = _call(lib, "b200_gptq_marlin_24_gemm", base, a, q, meta, s, ws, bits, M, N, K, groups)
    assert rc == 0, lib.b200_last_error()
    torch.cuda.synchronize()
    assert ops_ref.rel_mean_err(base.cpu(), ref) < 1e-2
    # two calls in a row through the allocation-free entry point; the second one finds the first one's scratch contents
    for i in range(2):
        out = torch.full((M, N), float("nan"), dtype=torch.float16, device=DEV)
        rc = _call(lib, "b200_gptq_marlin_24_gemm_static", out, a, q, meta, s, ws, bits, M, N, K, groups)
        assert rc == 0, lib.b200_last_error()
        torch.cuda.synchronize()
        # same kernel, same deterministic K-split order: bit-identical, and within the suite's tolerance of the oracle
        assert torch.equal(out, base), (i, float((out.float() - base.float()).abs().max()))
        assert ops_ref.rel_mean_err(out.cpu(), ref) < 1e-2
    assert torch.all(ws == 0)


def test_static_scratch_entry_back_to_back_different_shapes(lib):
    """calls of different shapes share the array: each must clear what it needs"""
    outs = []
    for (M, K, N, bits, group) in CASES[:4] + CASES[:2]:
        a, q, meta, s, ref = _problem(M, K, N, bits, group, seed=7)
        ws = torch.zeros((N // 128) * 64, dtype=torch.int32, device=DEV)
        out = torch.empty((M, N), dtype=torch.float16, device=DEV)
        rc = _call(lib, "b200_gptq_marlin_24_gemm_static", out, a, q, meta, s, ws, bits, M, N, K, s.shape[0])
        assert rc == 0, lib.b200_last_error()
        outs.append((out, ref))
    torch.cuda.synchronize()
    for out, ref in outs:
        assert ops_ref.rel_mean_err(out.cpu(), ref) < 1e-2


def test_static_scratch_entry_rejects_bad_arguments(lib):
    M, K, N = 4, 256, 256
    a = torch.zeros((M, K), dtype=torch.float16, device=DEV)
    q = torch.zeros((K // 32, N * 2), dtype=torch.int32, device=DEV)
    meta = torch.zeros((K // 32, 2 * N), dtype=torch.int16, device=DEV)
    s = torch.ones((1, N), dtype=torch.float16, device=DEV)
    ws = torch.zeros((N // 128) * 64, dtype=torch.int32, device=DEV)
    out = torch.empty((M, N), dtype=torch.float16, device=DEV)
    assert _call(lib, "b200_gptq_marlin_24_gemm_static", out, a, q, meta, s, ws, 3, M, N, K, 1) == -1
    assert b"num_bits" in lib.b200_last_error()
    assert _call(lib, "b200_gptq_marlin_24_gemm_static", out, a, q, meta, s, ws, 4, M, N + 64, K, 1) == -1
    assert b"size_n" in lib.b200_last_error()
