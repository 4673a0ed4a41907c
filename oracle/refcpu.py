"""ORACLE helper: loader for oracle/_ref/_refcpu_C.so -- the reference's own CPU operators (csrc/cpu/*.cpp compiled
unmodified by oracle/Makefile).  Gives torch.ops._refcpu_C.paged_attention_v1/v2 and
torch.ops._refcpu_C_cache_ops.reshape_and_cache on CPU tensors (limits: block_size 16, kv_scale 1.0 / no fp8,
dtypes fp32 / bf16 -- csrc/cpu/attention.cpp:241,430; csrc/cpu/cache.cpp:111)."""
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_ref", "_refcpu_C.so")
_state = {"loaded": False, "why": None}


def _isa_ok() -> bool:
    isa = os.path.join(_HERE, "_ref", "isa.txt")
    if not os.path.exists(isa):
        return True
    need = [f[2:] for f in open(isa).read().split() if f.startswith("-m")]
    try:
        flags = set()
        for line in open("/proc/cpuinfo"):
            if line.startswith("flags"):
                flags = set(line.split(":", 1)[1].split())
                break
    except OSError:
        return False
    alias = {"avx512bf16": "avx512_bf16"}
    return all(alias.get(n, n) in flags for n in need)


def available() -> bool:
    if _state["loaded"]:
        return True
    if not os.path.exists(_SO):
        _state["why"] = "oracle/_ref/_refcpu_C.so not built"
        return False
    if not _isa_ok():
        _state["why"] = "host CPU lacks the ISA the reference ops were compiled for"
        return False
    try:
        import torch
        torch.ops.load_library(_SO)
        _state["loaded"] = True
        return True
    except Exception as e:  # pragma: no cover
        _state["why"] = repr(e)
        return False


def why_unavailable():
    return _state["why"]


def paged_attention_v1(out, query, key_cache, value_cache, num_kv_heads, scale, block_tables, seq_lens, block_size,
                       max_seq_len, alibi_slopes=None):
    import torch
    assert available(), why_unavailable()
    torch.ops._refcpu_C.paged_attention_v1(out, query, key_cache, value_cache, num_kv_heads, scale, block_tables,
                                           seq_lens, block_size, max_seq_len, alibi_slopes, "auto", 1.0, 0, 0, 0, 64, 0)


def paged_attention_v2(out, exp_sums, max_logits, tmp_out, query, key_cache, value_cache, num_kv_heads, scale,
                       block_tables, seq_lens, block_size, max_seq_len, alibi_slopes=None):
    import torch
    assert available(), why_unavailable()
    torch.ops._refcpu_C.paged_attention_v2(out, exp_sums, max_logits, tmp_out, query, key_cache, value_cache,
                                           num_kv_heads, scale, block_tables, seq_lens, block_size, max_seq_len,
                                           alibi_slopes, "auto", 1.0, 0, 0, 0, 64, 0)


def reshape_and_cache(key, value, key_cache, value_cache, slot_mapping):
    import torch
    assert available(), why_unavailable()
    torch.ops._refcpu_C_cache_ops.reshape_and_cache(key, value, key_cache, value_cache, slot_mapping, "auto", 1.0)
