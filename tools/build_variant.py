"""Tuning helper: build a variant of libb200ops.so with extra -D flags for the decode GEMM translation units.
usage: python tools/build_variant.py NAME -DB200_DEC_NSHF_FMA=0 ...   ->  build/variants/NAME/libb200ops.so
Run with B200_LIB_OPS=build/variants/NAME/libb200ops.so."""
import os, subprocess, sys
from concurrent.futures import ThreadPoolExecutor
sys.path.insert(0, ".")
from nm_vllm_b200 import build as B
name, flags = sys.argv[1], sys.argv[2:]
out = os.path.join(B.ROOT, "build", "variants", name)
os.makedirs(out, exist_ok=True)
B.build_ops()
srcs = B.cuda_sources()
var = [s for s in srcs if os.path.basename(s).startswith(("marlin_decode_inst", "marlin_gemm_inst")) or os.path.basename(s) == "marlin_gemm.cu"]
def cc(s):
    o = os.path.join(out, os.path.basename(s)[:-3] + ".o")
    B._run([B.NVCC] + B.NVCC_FLAGS + flags + ["-c", s, "-o", o], log=o + ".log")
    return o
with ThreadPoolExecutor(8) as ex:
    vobjs = list(ex.map(cc, var))
objs = [os.path.join(B.OBJ, os.path.basename(s)[:-3] + ".o") for s in srcs if s not in var] + vobjs
B._run([B.NVCC, "-shared", "-o", os.path.join(out, "libb200ops.so")] + objs + ["-lcudart", "-Xlinker", "-rpath=/usr/local/cuda/lib64"])
print("built", out)
