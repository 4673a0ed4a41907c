"""Bring-up helper (not a test): runs the Marlin GEMM on a few shapes and prints the error vs the fp32 oracle."""
import sys
import time

import torch

sys.path.insert(0, ".")
from nm_vllm_b200 import _custom_ops as ops  # noqa: E402
from oracle import fixtures, ops_ref, quant_layouts as ql  # noqa: E402

DEV = "cuda:0"
cases = [(1, 128, 64, 4, 128, torch.float16), (1, 128, 128, 4, 128, torch.float16), (16, 256, 128, 4, 128, torch.float16),
         (16, 1024, 256, 4, 128, torch.bfloat16), (3, 512, 192, 4, -1, torch.float16), (33, 1024, 512, 4, 64, torch.float16),
         (64, 4096, 4096, 4, 128, torch.float16), (8, 4096, 6144, 4, 128, torch.bfloat16), (200, 1024, 256, 4, 128, torch.float16),
         (5, 512, 256, 8, 128, torch.float16), (5, 512, 256, 8, 128, torch.bfloat16), (300, 512, 256, 4, 32, torch.bfloat16)]
if len(sys.argv) > 1:
    cases = cases[:int(sys.argv[1])]
for (M, K, N, bits, group, dt) in cases:
    w = fixtures.gemm_weights(K, N, seed=1, dtype=dt)
    a = fixtures.gemm_acts(M, K, seed=1, dtype=dt)
    w_ref, mq, ms, g_idx, sort_idx, _ = ql.marlin_quantize(w, bits, group)
    ws = torch.zeros((N // 64) * 16, dtype=torch.int32, device=DEV)
    t0 = time.time()
    out = ops.gptq_marlin_gemm(a.to(DEV), mq.to(DEV), ms.to(dt).to(DEV), g_idx.to(DEV), sort_idx.to(DEV), ws, bits, M, N, K, True)
    torch.cuda.synchronize()
    ref = ops_ref.marlin_gemm_ref(a, w_ref.to(dt))
    err = ops_ref.rel_mean_err(out.cpu(), ref)
    print(f"M={M} K={K} N={N} bits={bits} group={group} {dt}: rel_mean_err={err:.5f} maxabs={float((out.cpu().float()-ref).abs().max()):.4f} "
          f"refmax={float(ref.abs().max()):.2f} ({time.time()-t0:.2f}s)", flush=True)
    if err > 0.05:
        o = out.cpu().float()
        print("  out[0,:8] ", o[0, :8].tolist())
        print("  ref[0,:8] ", ref[0, :8].tolist())
        print("  out[0,64:72]", o[0, 64:72].tolist() if N > 64 else None)
        print("  ref[0,64:72]", ref[0, 64:72].tolist() if N > 64 else None)
print("done")
