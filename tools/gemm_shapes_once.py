"""One launch of the W4A16 GEMM per (Llama-3-8B layer shape, M in {1,16,64}) after a warm-up launch each (for ncu metric runs)."""
import sys
import torch
sys.path.insert(0, ".")
from nm_vllm_b200 import _custom_ops as ops
DEV = "cuda:0"
SHAPES = {"qkv": (4096, 6144), "o": (4096, 4096), "gate_up": (4096, 28672), "down": (14336, 4096)}
g = torch.Generator(device=DEV).manual_seed(0)
for name, (K, N) in SHAPES.items():
    ws = [torch.randint(-2**31, 2**31 - 1, (K // 16, N * 2), generator=g, dtype=torch.int32, device=DEV) for _ in range(2)]
    s = (torch.rand((K // 128, N), generator=g, device=DEV) * 0.01 + 0.005).to(torch.bfloat16)
    e = torch.empty(0, dtype=torch.int32, device=DEV)
    wk = torch.zeros((N // 64) * 16, dtype=torch.int32, device=DEV)
    for M in (1, 16, 64):
        a = torch.randn((M, K), device=DEV, dtype=torch.bfloat16)
        for i in range(2):   # launch 0 = warm-up, launch 1 = the one to read (other weights: cold L2)
            ops.gptq_marlin_gemm(a, ws[i], s, e, e, wk, 4, M, N, K, True)
        torch.cuda.synchronize()
print("ok")
