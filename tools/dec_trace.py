"""Bring-up helper: event trace of one CTA of the decode GEMM (needs a -DB200_DEC_TRACE variant build).  usage: K N M [cta]"""
import os, sys
os.environ.setdefault("B200_DEC_TRACE", sys.argv[4] if len(sys.argv) > 4 else "5")
import torch
sys.path.insert(0, ".")
from nm_vllm_b200 import _custom_ops as ops
DEV = "cuda:0"
K, N, M = (int(x) for x in sys.argv[1:4])
g = torch.Generator(device=DEV).manual_seed(0)
ws = [torch.randint(-2**31, 2**31 - 1, (K // 16, N * 2), generator=g, dtype=torch.int32, device=DEV) for _ in range(3)]
s = (torch.rand((K // 128, N), generator=g, device=DEV) * 0.01 + 0.005).to(torch.bfloat16)
e = torch.empty(0, dtype=torch.int32, device=DEV)
wk = torch.zeros((N // 64) * 16, dtype=torch.int32, device=DEV)
a = torch.randn((M, K), device=DEV, dtype=torch.bfloat16)
for i in range(3):
    print("rep", i, file=sys.stderr, flush=True)
    ops.gptq_marlin_gemm(a, ws[i], s, e, e, wk, 4, M, N, K, True)
    torch.cuda.synchronize()
