"""ctypes view of the C ABI (include/b200_ops.h).  Used by tests (symbol/export checks, direct C-ABI calls) and by
INTEGRATION.md as the example of a binding that needs no torch headers."""
import ctypes
import os
import re

from . import extension_paths

_lib = None

# every extern "C" symbol include/b200_ops.h declares
def declared_symbols():
    hdr = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "b200_ops.h")
    txt = open(hdr).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", txt)))


def lib():
    global _lib
    if _lib is None:
        path = extension_paths()[0]
        if not os.path.exists(path):
            raise RuntimeError("libb200ops.so is not built: run `python -m nm_vllm_b200.build`")
        _lib = ctypes.CDLL(path, mode=ctypes.RTLD_GLOBAL)
        _lib.b200_last_error.restype = ctypes.c_char_p
        _lib.b200_build_info.restype = ctypes.c_char_p
    return _lib


def last_error() -> str:
    return lib().b200_last_error().decode()
