"""CPU (no GPU needed): the C-ABI library loads, exports every symbol include/b200_ops.h declares, the torch shim
registers the reference's schemas, and precondition failures are reported before any launch."""
import ctypes

import pytest
import torch


def test_library_exports_every_declared_symbol(built_extension):
    from nm_vllm_b200 import capi
    lib = capi.lib()
    names = capi.declared_symbols()
    assert len(names) >= 20
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    assert b"sm_100a" in lib.b200_build_info()


def test_torch_schemas_match_reference(built_extension):
    # schema strings of /root/reference/csrc/torch_bindings.cpp:24-48, 146-150, 174-182, 194-204, 221-243
    s = str(torch.ops._b200_C.paged_attention_v1.default._schema)
    assert "Tensor(a0!) out" in s or "Tensor($0! -> ) out" in s or "out" in s
    for frag in ["int num_kv_heads", "float scale", "Tensor block_tables", "Tensor seq_lens", "int block_size",
                 "int max_seq_len", "Tensor? alibi_slopes", "str kv_cache_dtype", "float kv_scale", "int tp_rank",
                 "int blocksparse_local_blocks", "int blocksparse_vert_stride", "int blocksparse_block_size",
                 "int blocksparse_head_sliding_step"]:
        assert frag in s, frag
    s2 = str(torch.ops._b200_C.paged_attention_v2.default._schema)
    for frag in ["exp_sums", "max_logits", "tmp_out"]:
        assert frag in s2
    assert "Tensor? bias" in str(torch.ops._b200_C.cutlass_scaled_mm.default._schema)
    assert str(torch.ops._b200_C.gptq_marlin_gemm.default._schema).count("Tensor") >= 7  # 6 inputs + return
    assert "-> Tensor" in str(torch.ops._b200_C.gptq_marlin_repack.default._schema)
    assert "str kv_cache_dtype" in str(torch.ops._b200_C_cache_ops.reshape_and_cache.default._schema)
    assert "float scale, str kv_cache_dtype" in str(torch.ops._b200_C_cache_ops.convert_fp8.default._schema)
    for name in ["static_scaled_fp8_quant", "dynamic_scaled_fp8_quant", "static_scaled_int8_quant",
                 "dynamic_scaled_int8_quant", "marlin_gemm", "gptq_marlin_24_gemm", "cutlass_scaled_mm_supports_fp8"]:
        assert hasattr(torch.ops._b200_C, name)


def _reference_schema_strings():
    """every op the reference registers with an explicit schema string (ops.def("name(args) -> ret")), by library"""
    import os
    import re
    path = os.path.join(os.environ.get("B200_REFERENCE_ROOT", "/root/reference"), "csrc", "torch_bindings.cpp")
    if not os.path.exists(path):
        return None
    txt = open(path).read()
    out = {}
    for m in re.finditer(r"(\w+)\.def\(\s*((?:\"(?:[^\"\\\\]|\\\\.)*\"\s*)+)\)", txt):
        lib, lit = m.group(1), "".join(re.findall(r"\"((?:[^\"\\\\]|\\\\.)*)\"", m.group(2)))
        lit = " ".join(lit.split())
        if "(" in lit:
            out.setdefault(lib, {})[lit.split("(")[0].strip()] = lit
    return out


def test_torch_schemas_equal_reference_strings(built_extension):
    """Every schema string this repo registers for a hot-path op parses to EXACTLY the schema the reference's
    csrc/torch_bindings.cpp declares (argument names, types, mutability annotations, order, return)."""
    ref = _reference_schema_strings()
    if ref is None:
        pytest.skip("reference tree not present")
    libs = {"ops": torch.ops._b200_C, "cache_ops": torch.ops._b200_C_cache_ops, "custom_ar": torch.ops._b200_C_custom_ar}
    checked = 0
    for lib, mine in libs.items():
        for name, lit in ref.get(lib, {}).items():
            if not hasattr(mine, name):
                continue
            want = torch._C.parse_schema("x::" + lit)
            got = getattr(mine, name).default._schema
            assert str(got).split("::", 1)[1] == str(want).split("::", 1)[1], (name, str(got), str(want))
            checked += 1
    assert checked >= 12, checked
    for name in ["init_custom_ar", "should_custom_ar", "all_reduce_reg", "all_reduce_unreg", "dispose", "meta_size",
                 "register_buffer", "get_graph_buffer_ipc_meta", "register_graph_buffers"]:
        assert hasattr(torch.ops._b200_C_custom_ar, name), name


def test_custom_ops_mirror_signatures(built_extension):
    import inspect

    from nm_vllm_b200 import _custom_ops as ops
    sig = inspect.signature(ops.paged_attention_v1)
    assert list(sig.parameters)[:13] == ["out", "query", "key_cache", "value_cache", "num_kv_heads", "scale",
                                         "block_tables", "seq_lens", "block_size", "max_seq_len", "alibi_slopes",
                                         "kv_cache_dtype", "kv_scale"]
    assert sig.parameters["blocksparse_block_size"].default == 64 and sig.parameters["tp_rank"].default == 0
    assert list(inspect.signature(ops.gptq_marlin_gemm).parameters) == [
        "a", "b_q_weight", "b_scales", "g_idx", "perm", "workspace", "num_bits", "size_m", "size_n", "size_k",
        "is_k_full"]
    assert list(inspect.signature(ops.cutlass_scaled_mm).parameters) == ["a", "b", "scale_a", "scale_b", "out_dtype",
                                                                         "bias"]
    assert list(inspect.signature(ops.reshape_and_cache).parameters) == [
        "key", "value", "key_cache", "value_cache", "slot_mapping", "kv_cache_dtype", "kv_scale"]


def test_preconditions_rejected_before_launch(built_extension):
    from nm_vllm_b200 import capi
    lib = capi.lib()
    # num_bits must be 4 or 8 (gptq_marlin_repack.cu:280-304)
    rc = lib.b200_gptq_marlin_repack(None, None, None, ctypes.c_int64(128), ctypes.c_int64(64), 3, None)
    assert rc == -1 and "num_bits must be 4 or 8" in capi.last_error()
    rc = lib.b200_gptq_marlin_repack(None, None, None, ctypes.c_int64(120), ctypes.c_int64(64), 4, None)
    assert rc == -1 and "not divisible by tile_k_size" in capi.last_error()
    # unsupported head size / block size (attention_kernels.cu:763-765, 800-802)
    rc = lib.b200_paged_attention_v1(None, None, None, None, 1, 8, 8, 72, ctypes.c_float(1.0), None, None, 1, 16, 16,
                                     None, ctypes.c_int64(0), ctypes.c_int64(0), ctypes.c_int64(0), 0, 0,
                                     ctypes.c_float(1.0), None)
    assert rc == -1 and "Unsupported head size" in capi.last_error()
    rc = lib.b200_paged_attention_v1(None, None, None, None, 1, 8, 8, 64, ctypes.c_float(1.0), None, None, 1, 24, 16,
                                     None, ctypes.c_int64(0), ctypes.c_int64(0), ctypes.c_int64(0), 0, 0,
                                     ctypes.c_float(1.0), None)
    assert rc == -1 and "Unsupported block size" in capi.last_error()
    assert lib.b200_scaled_mm_supports_fp8(100) == 1 and lib.b200_scaled_mm_supports_fp8(80) == 0


def test_shim_rejects_bad_shapes_on_cpu(built_extension):
    # TORCH_CHECKs of the shim fire before any device work (gptq_marlin.cu:1741-1763 restated)
    a = torch.zeros(2, 128, dtype=torch.float16)
    with pytest.raises(RuntimeError, match="num_bits must be 4 or 8"):
        torch.ops._b200_C.gptq_marlin_gemm(a, a, a, a, a, a, 5, 2, 64, 128, True)
    with pytest.raises(RuntimeError, match="Shape mismatch: a.size\\(0\\)"):
        torch.ops._b200_C.gptq_marlin_gemm(a, a, a, a, a, a, 4, 3, 64, 128, True)


def test_product_path_never_imports_oracle():
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    bad = []
    for dp, _, fns in os.walk(os.path.join(root, "nm_vllm_b200")):
        for fn in fns:
            if fn.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                txt = open(os.path.join(dp, fn), errors="ignore").read()
                if "import oracle" in txt or "from oracle" in txt or "oracle/" in txt:
                    bad.append(os.path.join(dp, fn))
    assert not bad, bad


def test_all_reduce_host_checks_without_gpu():
    """Argument validation of the NVLink all-reduce entry points happens before any CUDA call."""
    import ctypes
    from nm_vllm_b200 import capi
    lib = capi.lib()
    n = ctypes.c_int64()
    assert lib.b200_ar_region_bytes(2, ctypes.c_int64(1 << 20), ctypes.byref(n)) == 0
    two = n.value
    assert lib.b200_ar_region_bytes(8, ctypes.c_int64(1 << 20), ctypes.byref(n)) == 0
    assert n.value > two > 2 * 2 * (1 << 20)          # flags + LL cells + two parities of W slots
    assert lib.b200_ar_region_bytes(9, ctypes.c_int64(1 << 20), ctypes.byref(n)) == -1   # B200_EINVAL
    assert lib.b200_ar_region_bytes(2, ctypes.c_int64(0), ctypes.byref(n)) == -1
    ctx = ctypes.c_int64()
    regions = (ctypes.c_void_p * 3)(1, 2, 3)
    assert lib.b200_ar_init(0, 3, regions, ctypes.c_int64(1 << 20), ctypes.byref(ctx)) == -1   # world must be 2/4/6/8
    assert "world must be" in capi.last_error()
    assert lib.b200_ar_all_reduce(ctypes.c_int64(12345), None, None, ctypes.c_int64(8), 1, None) == -1
    assert lib.b200_ar_fuse_next_gemm(ctypes.c_int64(12345), ctypes.c_int64(4), ctypes.c_int64(128)) == -1
    assert lib.b200_ar_dispose(ctypes.c_int64(0)) == -1
