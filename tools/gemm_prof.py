"""Bring-up helper: per-role cycle counters of the Marlin GEMM (B200_GEMM_PROF=1), eager launches."""
import os, sys
os.environ["B200_GEMM_PROF"] = "1"
import torch
sys.path.insert(0, ".")
from nm_vllm_b200 import _custom_ops as ops
DEV = "cuda:0"
SHAPES = {"qkv": (4096, 6144), "o": (4096, 4096), "gate_up": (4096, 28672), "down": (14336, 4096)}
for name, (K, N) in SHAPES.items():
    g = torch.Generator(device=DEV).manual_seed(0)
    w = torch.randint(-2**31, 2**31 - 1, (K // 16, N * 2), generator=g, dtype=torch.int32, device=DEV)
    s = (torch.rand((K // 128, N), generator=g, device=DEV) * 0.01 + 0.005).to(torch.bfloat16)
    e = torch.empty(0, dtype=torch.int32, device=DEV)
    wk = torch.zeros((N // 64) * 16, dtype=torch.int32, device=DEV)
    for M in (1, 16, 64):
        a = torch.randn((M, K), device=DEV, dtype=torch.bfloat16)
        for i in range(2):
            print(name, "M", M, "rep", i, file=sys.stderr, flush=True)
            ops.gptq_marlin_gemm(a, w, s, e, e, wk, 4, M, N, K, True)
            torch.cuda.synchronize()
