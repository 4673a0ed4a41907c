"""In-tree build of the sm_100a extension (no JIT cache: the .so files must travel with the repo snapshot).

  libb200ops.so  -- all CUDA kernels + the C ABI of include/b200_ops.h   (nvcc, sm_100a only)
  _b200_C.so     -- torch operator registrations forwarding to the C ABI (g++ against torch headers)

`python -m nm_vllm_b200.build` or `__graft_entry__.build()`.  Incremental on mtimes; `force=True` rebuilds.
"""
import os
import subprocess
import sys
import sysconfig
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(ROOT, "build", "obj")
LIB_OPS = os.path.join(HERE, "libb200ops.so")
LIB_TORCH = os.path.join(HERE, "_b200_C.so")

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "-Xptxas", "-v",
]


def _newer(src_list, target):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in src_list)


def _run(cmd, log=None):
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if log is not None:
        with open(log, "w") as f:
            f.write(" ".join(cmd) + "\n" + p.stdout)
    if p.returncode != 0:
        sys.stderr.write(p.stdout)
        raise RuntimeError("build step failed: " + " ".join(cmd[:6]) + " ...")
    return p.stdout


def cuda_sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))


def build_ops(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    headers.append(os.path.join(ROOT, "include", "b200_ops.h"))
    srcs = cuda_sources()
    jobs = []
    for s in srcs:
        o = os.path.join(OBJ, os.path.basename(s)[:-3] + ".o")
        if force or _newer([s] + headers, o):
            jobs.append((s, o))

    def compile_one(job):
        s, o = job
        _run([NVCC] + NVCC_FLAGS + ["-c", s, "-o", o], log=o + ".log")
        return s

    if jobs:
        with ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            for s in ex.map(compile_one, jobs):
                if verbose:
                    print("nvcc", os.path.basename(s))
    objs = [os.path.join(OBJ, os.path.basename(s)[:-3] + ".o") for s in srcs]
    if jobs or force or _newer(objs, LIB_OPS):
        _run([NVCC, "-shared", "-o", LIB_OPS] + objs + ["-lcudart", "-Xlinker", "-rpath=/usr/local/cuda/lib64",
                                                          "-Xlinker", "-soname=libb200ops.so"])
    return LIB_OPS


def build_torch(force=False, verbose=False):
    import torch
    from torch.utils import cpp_extension as ce

    src = os.path.join(CSRC, "torch_bindings.cpp")
    hdr = os.path.join(ROOT, "include", "b200_ops.h")
    if not (force or _newer([src, hdr], LIB_TORCH)):
        return LIB_TORCH
    inc = []
    for p in ce.include_paths(device_type="cuda") if "device_type" in ce.include_paths.__code__.co_varnames else ce.include_paths(True):
        inc += ["-I", p]
    inc += ["-I", sysconfig.get_paths()["include"], "-I", "/usr/local/cuda/include"]
    tlib = os.path.join(os.path.dirname(torch.__file__), "lib")
    abi = int(torch._C._GLIBCXX_USE_CXX11_ABI)
    cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-w", f"-D_GLIBCXX_USE_CXX11_ABI={abi}",
           "-DTORCH_EXTENSION_NAME=_b200_C", src, "-o", LIB_TORCH] + inc + [
           "-L", tlib, "-ltorch", "-ltorch_cpu", "-lc10", "-ltorch_cuda", "-lc10_cuda",
           "-L", HERE, "-lb200ops", "-Wl,-rpath,$ORIGIN", f"-Wl,-rpath,{tlib}"]
    if verbose:
        print("g++ torch_bindings.cpp")
    _run(cmd)
    return LIB_TORCH


def build_all(force=False, verbose=False):
    build_ops(force, verbose)
    build_torch(force, verbose)
    return LIB_OPS, LIB_TORCH


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, verbose=True)
    print("built", LIB_OPS, LIB_TORCH)
