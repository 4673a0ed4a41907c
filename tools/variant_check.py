"""Does a B200_LIB_OPS variant really run?  One GEMM, prints a checksum of the output (an ablated kernel gives garbage / zeros)."""
import os, sys
import torch
sys.path.insert(0, ".")
from nm_vllm_b200 import _custom_ops as ops
DEV = "cuda:0"
K, N, M = 4096, 28672, 16
g = torch.Generator(device=DEV).manual_seed(0)
w = torch.randint(-2**31, 2**31 - 1, (K // 16, N * 2), generator=g, dtype=torch.int32, device=DEV)
s = (torch.rand((K // 128, N), generator=g, device=DEV) * 0.01 + 0.005).to(torch.bfloat16)
e = torch.empty(0, dtype=torch.int32, device=DEV)
wk = torch.zeros((N // 64) * 16, dtype=torch.int32, device=DEV)
a = torch.randn((M, K), generator=g, device=DEV, dtype=torch.bfloat16)
out = ops.gptq_marlin_gemm(a, w, s, e, e, wk, 4, M, N, K, True)
torch.cuda.synchronize()
print("lib", os.environ.get("B200_LIB_OPS"), "exp", os.environ.get("B200_GEMM_EXP"), "checksum", float(out.float().abs().sum()), "nan", bool(torch.isnan(out.float()).any()))
