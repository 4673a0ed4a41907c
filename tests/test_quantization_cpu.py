"""CPU tests of the drop-in Python surface: scale permutation against the golden tables and
against the oracle, parameter names / shapes / dtypes / loader attributes of every LinearMethod mirror against the
REFERENCE's own classes (loaded by file path with its framework imports stubbed; skipped where /root/reference is
absent), config parsing, PagedAttention shapes, and `install()` rebinding."""
import importlib.util
import os
import sys
import types

import pytest
import torch

from nm_vllm_b200 import quantization as Q
from nm_vllm_b200.attention import PagedAttention, PagedDecodeImpl, PagedDecodeMetadata
from nm_vllm_b200.quantization import base, layouts
from oracle import quant_layouts as ql

REF = os.environ.get("B200_REFERENCE_ROOT", "/root/reference")
QDIR = os.path.join(REF, "vllm/model_executor/layers/quantization")


def test_scale_perm_tables_match_oracle_tables():
    _, sp, sps = ql.marlin_perm_tables(4)
    g, s = layouts.get_scale_perms(4)
    assert list(sp) == g and list(sps) == s


@pytest.mark.parametrize("group,k,n", [(128, 512, 256), (64, 256, 128), (-1, 256, 192), (256, 256, 64)])
def test_marlin_permute_scales_roundtrip_with_oracle(group, k, n):
    rows = 1 if (group == -1 or group >= k) else k // group
    s = torch.arange(rows * n, dtype=torch.float32).reshape(rows, n).to(torch.float16)
    p = layouts.marlin_permute_scales(s, k, n, group, 4)
    _, sp, sps = ql.marlin_perm_tables(4)
    back = ql.unpermute_scales(p, k, n, group if group != -1 else k, sp, sps)
    assert torch.equal(back, s)


def _load_reference_quant():
    """The reference's gptq_marlin / gptq_marlin_24 / marlin / fp8 / compressed-tensors modules with vllm's framework
    replaced by this repo's interface mirrors (they only need set_weight_attrs, the base classes and `ops`)."""
    saved = {k: v for k, v in sys.modules.items() if k == "vllm" or k.startswith("vllm.")}
    for k in list(saved):
        del sys.modules[k]

    def mod(name, **attrs):
        m = types.ModuleType(name)
        m.__path__ = []
        m.__dict__.update(attrs)
        sys.modules[name] = m
        return m

    class _Plat:
        @staticmethod
        def get_device_capability():
            return (10, 0)

    class _Log:
        def __getattr__(self, n):
            return lambda *a, **k: None

    class _Stub(torch.nn.Module):
        pass

    mod("vllm", _custom_ops=types.SimpleNamespace(cutlass_scaled_mm_supports_fp8=lambda c: True))
    sys.modules["vllm._custom_ops"] = sys.modules["vllm"]._custom_ops
    mod("vllm.logger", init_logger=lambda n: _Log())
    mod("vllm.platforms", current_platform=_Plat())
    mod("vllm.utils", print_warning_once=lambda *a, **k: None)
    mod("vllm.model_executor")
    mod("vllm.model_executor.utils", set_weight_attrs=base.set_weight_attrs)
    mod("vllm.model_executor.layers")
    mod("vllm.model_executor.layers.linear", LinearBase=base.LinearBase, LinearMethodBase=base.LinearMethodBase,
        set_weight_attrs=base.set_weight_attrs)
    mod("vllm.model_executor.layers.vocab_parallel_embedding", ParallelLMHead=_Stub)
    mod("vllm.model_executor.layers.fused_moe", FusedMoE=_Stub, FusedMoEMethodBase=object, fused_moe=None)
    mod("vllm.model_executor.layers.quantization")
    mod("vllm.model_executor.layers.quantization.base_config", QuantizationConfig=base.QuantizationConfig,
        QuantizeMethodBase=base.QuantizeMethodBase)
    mod("vllm.model_executor.layers.quantization.utils")
    mod("vllm.model_executor.layers.quantization.utils.marlin_utils", pack_fp8_to_int32=None)
    mod("vllm.model_executor.layers.quantization.compressed_tensors")
    out = {}

    def load(name, rel):
        full = "vllm.model_executor.layers.quantization." + name
        spec = importlib.util.spec_from_file_location(full, os.path.join(QDIR, rel))
        m = importlib.util.module_from_spec(spec)
        sys.modules[full] = m
        spec.loader.exec_module(m)
        out[name] = m
        return m

    try:
        load("gptq_marlin", "gptq_marlin.py")
        load("gptq_marlin_24", "gptq_marlin_24.py")
        load("marlin", "marlin.py")
        load("fp8", "fp8.py")
        load("compressed_tensors.utils", "compressed_tensors/utils.py")
        mod("vllm.model_executor.layers.quantization.compressed_tensors.schemes")
        cs = load("compressed_tensors.schemes.compressed_tensors_scheme", "compressed_tensors/schemes/compressed_tensors_scheme.py")
        sys.modules["vllm.model_executor.layers.quantization.compressed_tensors.schemes"].CompressedTensorsScheme = cs.CompressedTensorsScheme
        load("compressed_tensors.schemes.compressed_tensors_wNa16", "compressed_tensors/schemes/compressed_tensors_wNa16.py")
        load("compressed_tensors.schemes.compressed_tensors_w4a16_24", "compressed_tensors/schemes/compressed_tensors_w4a16_24.py")
        load("compressed_tensors.schemes.compressed_tensors_w8a8", "compressed_tensors/schemes/compressed_tensors_w8a8.py")
    finally:
        for k in [k for k in sys.modules if k == "vllm" or k.startswith("vllm.")]:
            del sys.modules[k]
        sys.modules.update(saved)
    return out


def _describe(layer):
    """name -> (shape, dtype, loader attributes) of everything create_weights hung on the layer"""
    keys = ("input_dim", "output_dim", "packed_dim", "pack_factor", "marlin_tile_size", "needs_scalar_to_array", "ignore_warning")
    d = {}
    for name, p in layer.named_parameters():
        d[name] = (tuple(p.shape), str(p.dtype), {k: getattr(p, k) for k in keys if hasattr(p, k)})
    for name in ("workspace", "g_idx_sort_indices"):
        t = layer.__dict__.get(name)
        if isinstance(t, torch.Tensor):
            d[name] = (tuple(t.shape), str(t.dtype), {})
    for name in ("input_size_per_partition", "output_size_per_partition", "input_size", "is_k_full", "group_size", "logical_widths",
                 "process_after_load", "orig_dtype"):
        if name in layer.__dict__:
            d["attr:" + name] = layer.__dict__[name]
    if "marlin_state" in layer.__dict__:
        d["attr:marlin_state"] = layer.__dict__["marlin_state"].name
    return d


needs_ref = pytest.mark.skipif(not os.path.isdir(QDIR), reason="reference tree not present")


@pytest.fixture(scope="module")
def refq():
    # the reference allocates the pre-packed Marlin parameters with device="cuda": route those to the CPU for this test
    real_empty, real_zeros = torch.empty, torch.zeros

    def strip(fn):
        def f(*a, **k):
            if k.get("device") == "cuda":
                k["device"] = "cpu"
            return fn(*a, **k)
        return f

    torch.empty, torch.zeros = strip(real_empty), strip(real_zeros)
    try:
        yield _load_reference_quant()
    finally:
        torch.empty, torch.zeros = real_empty, real_zeros


@needs_ref
@pytest.mark.parametrize("bits,group,desc,k_part,k_full,dtype", [
    (4, 128, False, 4096, 4096, torch.bfloat16), (4, 128, False, 2048, 4096, torch.float16), (4, -1, False, 1024, 1024, torch.float16),
    (8, 64, False, 512, 512, torch.float16), (4, 128, True, 2048, 4096, torch.float16), (4, 128, True, 4096, 4096, torch.bfloat16)])
def test_gptq_marlin_create_weights_matches_reference(refq, bits, group, desc, k_part, k_full, dtype):
    ref = refq["gptq_marlin"]
    a, b = torch.nn.Module(), torch.nn.Module()
    ref.GPTQMarlinLinearMethod(ref.GPTQMarlinConfig(bits, group, desc, True, False)).create_weights(a, k_part, [256, 128], k_full, 384, dtype)
    Q.GPTQMarlinLinearMethod(Q.GPTQMarlinConfig(bits, group, desc, True, False)).create_weights(b, k_part, [256, 128], k_full, 384, dtype)
    assert _describe(a) == _describe(b)


@needs_ref
def test_marlin_family_create_weights_matches_reference(refq):
    for (rm, rc, rl, mc, ml, cargs) in (
            ("gptq_marlin_24", "GPTQMarlin24Config", "GPTQMarlin24LinearMethod", Q.GPTQMarlin24Config, Q.GPTQMarlin24LinearMethod, (4, 128)),
            ("gptq_marlin_24", "GPTQMarlin24Config", "GPTQMarlin24LinearMethod", Q.GPTQMarlin24Config, Q.GPTQMarlin24LinearMethod, (8, -1)),
            ("marlin", "MarlinConfig", "MarlinLinearMethod", Q.MarlinConfig, Q.MarlinLinearMethod, (128, False))):
        ref = refq[rm]
        a, b = torch.nn.Module(), torch.nn.Module()
        getattr(ref, rl)(getattr(ref, rc)(*cargs)).create_weights(a, 1024, [512, 256], 1024, 768, torch.float16)
        ml(mc(*cargs)).create_weights(b, 1024, [512, 256], 1024, 768, torch.float16)
        assert _describe(a) == _describe(b)


@needs_ref
@pytest.mark.parametrize("serialized,scheme", [(False, "dynamic"), (True, "dynamic"), (True, "static")])
def test_fp8_create_weights_matches_reference(refq, serialized, scheme):
    ref = refq["fp8"]
    a, b = torch.nn.Module(), torch.nn.Module()
    ref.Fp8LinearMethod(ref.Fp8Config(serialized, scheme)).create_weights(a, 512, [256, 64, 64], 512, 384, torch.bfloat16)
    Q.Fp8LinearMethod(Q.Fp8Config(serialized, scheme)).create_weights(b, 512, [256, 64, 64], 512, 384, torch.bfloat16)
    da, db = _describe(a), _describe(b)
    assert da == db
    if serialized:
        assert torch.equal(a.weight_scale, b.weight_scale)   # both initialised to finfo(e4m3).min


@needs_ref
def test_compressed_tensors_schemes_match_reference(refq):
    from nm_vllm_b200.quantization import compressed_tensors as ct
    pairs = [
        (refq["compressed_tensors.schemes.compressed_tensors_wNa16"].CompressedTensorsWNA16, ct.CompressedTensorsWNA16,
         dict(strategy="group", num_bits=4, group_size=128)),
        (refq["compressed_tensors.schemes.compressed_tensors_wNa16"].CompressedTensorsWNA16, ct.CompressedTensorsWNA16,
         dict(strategy="channel", num_bits=8, group_size=None)),
        (refq["compressed_tensors.schemes.compressed_tensors_w4a16_24"].CompressedTensorsW4A16Sparse24, ct.CompressedTensorsW4A16Sparse24,
         dict(strategy="group", num_bits=4, group_size=128)),
        (refq["compressed_tensors.schemes.compressed_tensors_w8a8"].CompressedTensorsW8A8, ct.CompressedTensorsW8A8,
         dict(strategy="tensor", is_static_input_scheme=True)),
        (refq["compressed_tensors.schemes.compressed_tensors_w8a8"].CompressedTensorsW8A8, ct.CompressedTensorsW8A8,
         dict(strategy="channel", is_static_input_scheme=False)),
    ]
    for rcls, mcls, kw in pairs:
        a, b = torch.nn.Module(), torch.nn.Module()
        args = dict(input_size=1024, output_partition_sizes=[256, 256], input_size_per_partition=512, params_dtype=torch.float16,
                    weight_loader=None, output_size=512)
        rcls(**kw).create_weights(layer=a, **args)
        mcls(**kw).create_weights(layer=b, **args)
        assert _describe(a) == _describe(b), (rcls.__name__, kw)


def test_compressed_tensors_config_picks_schemes():
    from nm_vllm_b200.quantization import compressed_tensors as ct

    def cfg(fmt, weights, acts=None):
        return ct.CompressedTensorsConfig.from_config({"format": fmt, "ignore": ["lm_head"], "config_groups": {
            "group_0": {"targets": ["Linear"], "weights": weights, "input_activations": acts}}})

    lin = base.LinearBase(64, 64)
    w4 = dict(num_bits=4, strategy="group", group_size=128, symmetric=True, dynamic=False, type="int")
    assert isinstance(cfg("pack-quantized", w4).get_scheme(lin), ct.CompressedTensorsWNA16)
    assert isinstance(cfg("marlin-24", w4).get_scheme(lin), ct.CompressedTensorsW4A16Sparse24)
    w8 = dict(num_bits=8, strategy="tensor", symmetric=True, dynamic=False, type="int")
    a_static = dict(num_bits=8, strategy="tensor", symmetric=True, dynamic=False, type="int")
    a_dyn = dict(num_bits=8, strategy="token", symmetric=True, dynamic=True, type="int")
    s = cfg("int-quantized", w8, a_static).get_scheme(lin)
    assert isinstance(s, ct.CompressedTensorsW8A8) and s.is_static_input_scheme
    s = cfg("int-quantized", w8, a_dyn).get_scheme(lin)
    assert isinstance(s, ct.CompressedTensorsW8A8) and not s.is_static_input_scheme
    with pytest.raises(NotImplementedError):
        cfg("int-quantized", dict(num_bits=8, strategy="block", symmetric=True, dynamic=False), a_static).get_scheme(lin)
    assert isinstance(cfg("pack-quantized", w4).get_quant_method(lin), ct.CompressedTensorsLinearMethod)


def test_gptq_marlin_config_validation_and_parsing():
    c = Q.GPTQMarlinConfig.from_config({"bits": 4, "group_size": 128, "desc_act": True, "sym": True})
    assert (c.weight_bits, c.group_size, c.desc_act, c.pack_factor) == (4, 128, True, 8)
    assert Q.GPTQMarlinConfig(4, -1, True, True, False).desc_act is False
    for bad in ((3, 128, False, True), (4, 100, False, True), (4, 128, False, False)):
        with pytest.raises(ValueError):
            Q.GPTQMarlinConfig(*bad, False)
    assert Q.GPTQMarlinConfig.is_marlin_compatible({"bits": 4, "group_size": 128, "sym": True, "desc_act": False})
    assert not Q.GPTQMarlinConfig.is_marlin_compatible({"bits": 4, "group_size": 128, "sym": False, "desc_act": False})
    assert Q.GPTQMarlinConfig.override_quantization_method({"bits": 4, "group_size": 128, "sym": True, "desc_act": False}, None) == "gptq_marlin"
    assert Q.get_quantization_config("fp8") is Q.Fp8Config
    m = Q.GPTQMarlinLinearMethod(c)
    with pytest.raises(ValueError):
        m.create_weights(torch.nn.Module(), 4096, [100], 4096, 100, torch.float16)      # N % 64
    with pytest.raises(ValueError):
        m.create_weights(torch.nn.Module(), 4000, [128], 4000, 128, torch.float16)      # K % 128
    with pytest.raises(ValueError):
        m.create_weights(torch.nn.Module(), 4096, [128], 4096, 128, torch.float32)


def test_fp8_kv_cache_method_sets_python_float():
    from nm_vllm_b200.attention import Attention
    cc = types.SimpleNamespace(cache_dtype="fp8", sliding_window=None)
    att = Attention(8, 128, 0.088, num_kv_heads=2, cache_config=cc, quant_config=Q.Fp8Config(True, "static"))
    assert isinstance(att.quant_method, Q.Fp8KVCacheMethod) and float(att.kv_scale) == 1.0
    att.kv_scale.data.fill_(0.0376)
    att.quant_method.process_weights_after_loading(att)
    assert isinstance(att._kv_scale, float) and abs(att._kv_scale - 0.0376) < 1e-6 and not hasattr(att, "kv_scale")


def test_paged_attention_shapes_and_views():
    assert PagedAttention.get_kv_cache_shape(10, 16, 8, 128) == (2, 10, 16 * 8 * 128)
    assert PagedAttention.get_supported_head_sizes() == [64, 80, 96, 112, 128, 192, 256]
    kv = torch.zeros(PagedAttention.get_kv_cache_shape(10, 16, 8, 128), dtype=torch.bfloat16)
    k, v = PagedAttention.split_kv_cache(kv, 8, 128)
    assert k.shape == (10, 8, 16, 16, 8) and v.shape == (10, 8, 128, 16)
    kv8 = torch.zeros(PagedAttention.get_kv_cache_shape(10, 16, 8, 128), dtype=torch.uint8)
    k, v = PagedAttention.split_kv_cache(kv8, 8, 128)
    assert k.shape == (10, 8, 8, 16, 16) and v.shape == (10, 8, 128, 16)
    with pytest.raises(ValueError):
        PagedDecodeImpl(8, 100, 0.1)
    md = PagedDecodeMetadata(seq_lens_tensor=torch.ones(2, dtype=torch.int32), max_decode_seq_len=1,
                             block_tables=torch.zeros((2, 1), dtype=torch.int32), num_prefills=0, num_prefill_tokens=0,
                             num_decode_tokens=2, slot_mapping=torch.zeros(2, dtype=torch.long))
    assert md.decode_metadata is md and md.prefill_metadata is None
    assert set(md.asdict_zerocopy()) >= {"slot_mapping", "block_tables", "seq_lens_tensor", "max_decode_seq_len"}


def test_install_rebinds_vllm_custom_ops(monkeypatch):
    """install() puts this repo's operators onto an importable `vllm._custom_ops` (what the reference's LinearMethod /
    PagedAttention classes call): exercised on a stand-in module so that an installed vllm is not modified."""
    from nm_vllm_b200 import _custom_ops as ours
    fake_pkg, fake_ops = types.ModuleType("vllm"), types.ModuleType("vllm._custom_ops")
    fake_pkg.__path__ = []
    for n in ("gptq_marlin_gemm", "paged_attention_v1", "reshape_and_cache", "cutlass_scaled_mm", "scaled_fp8_quant"):
        setattr(fake_ops, n, lambda *a, **k: "reference")
    fake_pkg._custom_ops = fake_ops
    monkeypatch.setitem(sys.modules, "vllm", fake_pkg)
    monkeypatch.setitem(sys.modules, "vllm._custom_ops", fake_ops)
    done = ours.install()
    for n in ("gptq_marlin_gemm", "paged_attention_v1", "reshape_and_cache", "cutlass_scaled_mm", "scaled_fp8_quant"):
        assert getattr(fake_ops, n) is getattr(ours, n)
        assert n in done


def test_sparse24_helpers_bit_exact_against_reference_golden():
    """product-side 2:4 compression / metadata reorder / Marlin-24 packing against vectors the REFERENCE's own utilities
    produced (oracle/make_golden.py)"""
    from nm_vllm_b200.quantization import sparse24
    from tests.util import as_torch, golden, golden_files
    files = golden_files("marlin24_")
    assert files
    for f in files:
        g = golden(f)
        K, N, bits, group = int(g["K"]), int(g["N"]), int(g["bits"]), int(g["group"])
        q24 = as_torch(g["q_w_24"])
        comp, meta = sparse24.compress_quantized_24_weight(q24, K, N, bits)
        assert torch.equal(comp, as_torch(g["comp_plain"]).to(torch.int32)), f
        assert torch.equal(meta, as_torch(g["meta"])), f
        assert torch.equal(sparse24.marlin_24_pack(comp, bits), as_torch(g["q_comp"])), f
    tabs = golden("perm_tables.npz")
    for bits in (4, 8):
        perm, sp, sps = sparse24.get_perms_24(bits)
        assert perm.tolist() == as_torch(tabs["perm24_%d" % bits]).tolist()
        assert sp == as_torch(tabs["scale_perm24_%d" % bits]).tolist() and sps == as_torch(tabs["scale_perm_single24_%d" % bits]).tolist()


def test_sparse24_quantize_matches_oracle():
    from nm_vllm_b200.quantization import sparse24
    w = (torch.randn((256, 128), generator=torch.Generator().manual_seed(4)) * 0.05).to(torch.float16)
    a = sparse24.marlin_24_quantize(w, 4, 128)
    b = ql.marlin_24_quantize(w, 4, 128)
    for x, y in zip(a, b):
        assert torch.equal(x, y.to(x.dtype))
    mask = sparse24.mask_2_of_4(w.t().contiguous())
    assert bool((mask.reshape(-1, 4).sum(1) == 2).all())
