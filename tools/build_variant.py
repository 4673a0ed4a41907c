"""Tuning helper: build a variant of libb200ops.so with extra -D flags for the decode GEMM translation units.
usage: python tools/build_variant.py NAME -DB200_DEC_NSHF_FMA=0 ...   ->  build/variants/NAME/libb200ops.so
Run with B200_LIB_OPS=build/variants/NAME/libb200ops.so.

PITFALL (cost a session of wrong conclusions): _b200_C.so has a NEEDED entry for libb200ops.so.  Without a SONAME on the
variant the loader does not recognise the pre-loaded variant as that library and maps the in-tree copy as well; two fatbins
then register kernels of the same name and WHICH one a launch runs differs per instantiation (some ablations silently ran
the production kernel and "changed nothing").  Both link lines now set -soname=libb200ops.so; tools/variant_check.py prints a
checksum to confirm that a variant really runs (an ablated kernel gives a different one)."""
import os, subprocess, sys
from concurrent.futures import ThreadPoolExecutor
sys.path.insert(0, ".")
from nm_vllm_b200 import build as B
name, flags = sys.argv[1], sys.argv[2:]
out = os.path.join(B.ROOT, "build", "variants", name)
os.makedirs(out, exist_ok=True)
B.build_ops()
srcs = B.cuda_sources()
var = [s for s in srcs if os.path.basename(s).startswith(("marlin_decode_inst", "marlin_gemm_inst")) or os.path.basename(s) in ("marlin_gemm.cu", "scaled_mm.cu")]
def cc(s):
    o = os.path.join(out, os.path.basename(s)[:-3] + ".o")
    B._run([B.NVCC] + B.NVCC_FLAGS + flags + ["-c", s, "-o", o], log=o + ".log")
    return o
with ThreadPoolExecutor(8) as ex:
    vobjs = list(ex.map(cc, var))
objs = [os.path.join(B.OBJ, os.path.basename(s)[:-3] + ".o") for s in srcs if s not in var] + vobjs
B._run([B.NVCC, "-shared", "-o", os.path.join(out, "libb200ops.so")] + objs + ["-lcudart", "-Xlinker", "-rpath=/usr/local/cuda/lib64",
                                                                          "-Xlinker", "-soname=libb200ops.so"])
print("built", out)
