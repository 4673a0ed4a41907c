"""nm_vllm_b200 -- B200-native (sm_100a) decode hot path behind nm-vllm's operator API.

Only what the path needs lives here:
  csrc/                 hand-written CUDA kernels + the C ABI (include/b200_ops.h) + the torch operator shim
  _custom_ops.py        mirror of vllm/_custom_ops.py for the hot-path operators (same names / argument order / returns)
  quantization/         QuantizationConfig / LinearMethodBase mirrors: gptq_marlin, marlin, gptq_marlin_24, fp8 (+ the fp8
                        kv-scale method), compressed-tensors (WNA16, W4A16 2:4, W8A8) + Marlin scale permutation helpers
  attention/            PagedAttention (+ metadata; forward_decode, forward_prefix), prefix_prefill.context_attention_fwd, the
                        decode half of the AttentionBackend / Impl / Metadata contract, the Attention layer that owns kv_scale
  custom_all_reduce.py  host side of the NVLink all-reduce kernels (mirror of the reference's CustomAllreduce)
  tp.py                 tensor-parallel sharding rules of the linear layers (column / row parallel, Marlin shard validity)
  decoder.py            synthetic Llama decode step and prefill built from the pieces above (bench / smoke / e2e harness)
  capi.py               ctypes view of the C ABI

There is no CPU fallback: calling an operator without the built extension raises.
"""
import os

__all__ = ["load_extension", "extension_paths", "is_built"]

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_OPS = os.path.join(_HERE, "libb200ops.so")
_LIB_TORCH = os.path.join(_HERE, "_b200_C.so")
_loaded = False


def extension_paths():
    return _LIB_OPS, _LIB_TORCH


def is_built() -> bool:
    return os.path.exists(_LIB_OPS) and os.path.exists(_LIB_TORCH)


def load_extension():
    """Load libb200ops.so + _b200_C.so (registers torch.ops._b200_C*).  Raises if the extension is not built."""
    global _loaded
    if _loaded:
        return
    if not is_built():
        raise RuntimeError(
            "nm_vllm_b200: the sm_100a extension is not built (expected %s and %s). "
            "Run `python -m nm_vllm_b200.build` (or __graft_entry__.build()). There is no CPU fallback." %
            (_LIB_OPS, _LIB_TORCH))
    import ctypes

    import torch

    # B200_LIB_OPS: tuning builds of the same library (tools/build_variant.py), same soname
    ctypes.CDLL(os.environ.get("B200_LIB_OPS", _LIB_OPS), mode=ctypes.RTLD_GLOBAL)
    torch.ops.load_library(_LIB_TORCH)
    _loaded = True
