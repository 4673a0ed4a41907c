"""ORACLE helper: import the REFERENCE's own Python quant/pack helpers by file path (authoring container only --
/root/reference does not exist on the GPU box, so nothing at test/bench run time may call this; it is used by
oracle/make_golden.py to produce tests/golden/*.npz and by tests that skip when the reference is absent)."""
import importlib.util
import os
import sys
import types

REF = os.environ.get("B200_REFERENCE_ROOT", "/root/reference")
_UTILS = os.path.join(REF, "vllm/model_executor/layers/quantization/utils")


def available() -> bool:
    return os.path.isdir(_UTILS)


def load():
    """Returns a namespace with quant_utils, format_24, marlin_perms, marlin_24_perms, marlin_utils (reference code)."""
    if not available():
        raise RuntimeError("reference tree not present")
    saved = {k: v for k, v in sys.modules.items() if k == "vllm" or k.startswith("vllm.")}
    for k in list(saved):
        del sys.modules[k]
    try:
        for name in ["vllm", "vllm.model_executor", "vllm.model_executor.layers",
                     "vllm.model_executor.layers.quantization", "vllm.model_executor.layers.quantization.utils",
                     "vllm.platforms"]:
            m = types.ModuleType(name)
            m.__path__ = []  # mark as package
            sys.modules[name] = m

        class _Plat:
            @staticmethod
            def get_device_capability():
                return (10, 0)

        sys.modules["vllm.platforms"].current_platform = _Plat()
        ns = types.SimpleNamespace()
        for name in ["quant_utils", "format_24", "marlin_perms", "marlin_24_perms", "marlin_utils"]:
            full = "vllm.model_executor.layers.quantization.utils." + name
            spec = importlib.util.spec_from_file_location(full, os.path.join(_UTILS, name + ".py"))
            mod = importlib.util.module_from_spec(spec)
            sys.modules[full] = mod
            spec.loader.exec_module(mod)
            setattr(ns, name, mod)

        # inject_24 hard-codes .cuda() (marlin_utils.py:112); same arithmetic without the device move
        def inject_24_cpu(w, size_k, size_n):
            mask = ns.format_24.mask_creator(w.t()).t().bool()
            return (mask * w).contiguous(), mask.contiguous()

        ns.marlin_utils.inject_24 = inject_24_cpu
        return ns
    finally:
        for k in [k for k in sys.modules if k == "vllm" or k.startswith("vllm.")]:
            del sys.modules[k]
        sys.modules.update(saved)
