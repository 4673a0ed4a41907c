"""Bring-up helper: a few launches of one Sparse-Marlin GEMM shape (for ncu captures).  usage: K N M"""
import sys
import torch
sys.path.insert(0, ".")
from nm_vllm_b200 import _custom_ops as ops
DEV = "cuda:0"
K, N, M = (int(x) for x in sys.argv[1:4])
g = torch.Generator(device=DEV).manual_seed(0)
LEGAL = torch.tensor([0x4, 0x8, 0x9, 0xC, 0xD, 0xE], dtype=torch.int32, device=DEV)
qs, ms = [], []
for _ in range(3):
    qs.append(torch.randint(-2**31, 2**31 - 1, (K // 32, N * 2), generator=g, dtype=torch.int32, device=DEV))
    quads = LEGAL[torch.randint(0, 6, (K // 32, 2 * N, 4), generator=g, device=DEV)]
    ms.append((quads[..., 0] | (quads[..., 1] << 4) | (quads[..., 2] << 8) | (quads[..., 3] << 12)).to(torch.int16))
s = (torch.rand((K // 128, N), generator=g, device=DEV) * 0.01 + 0.005).to(torch.float16)
ws = torch.zeros((N // 128) * 64, dtype=torch.int32, device=DEV)
a = torch.randn((M, K), device=DEV, dtype=torch.float16)
for i in range(4):
    out = ops.gptq_marlin_24_gemm(a, qs[i % 3], ms[i % 3], s, ws, 4, M, N, K)
torch.cuda.synchronize()
print("ok", float(out.float().abs().mean()))
