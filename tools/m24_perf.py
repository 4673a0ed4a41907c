"""Kernel micro-benchmark: gptq_marlin_24_gemm (2:4 W4A16 g128) on the Llama-3-8B layer shapes, cold L2 (rotating buffers)."""
import json
import sys

import torch

sys.path.insert(0, ".")
from nm_vllm_b200 import _custom_ops as ops  # noqa: E402

DEV = "cuda:0"
SHAPES = {"qkv": (4096, 6144), "o": (4096, 4096), "gate_up": (4096, 28672), "down": (14336, 4096)}
MS = [1, 16, 64] if len(sys.argv) < 2 else [int(x) for x in sys.argv[1].split(",")]
if len(sys.argv) > 2:
    SHAPES = {k: v for k, v in SHAPES.items() if k in sys.argv[2].split(",")}
peak = 6524.6
try:
    peak = json.load(open("MEASURED_PEAKS.json"))["hbm_gbs"]
except Exception:
    pass
LEGAL = torch.tensor([0x4, 0x8, 0x9, 0xC, 0xD, 0xE], dtype=torch.int32, device=DEV)  # format_24.py:96-102
res = {}
for name, (K, N) in SHAPES.items():
    wbytes = K * N // 4 + K * N // 8
    ncopy = max(3, int(400e6 // wbytes) + 1)
    g = torch.Generator(device=DEV).manual_seed(0)
    qs, ms = [], []
    for _ in range(ncopy):
        qs.append(torch.randint(-2**31, 2**31 - 1, (K // 32, N * 2), generator=g, dtype=torch.int32, device=DEV))
        quads = LEGAL[torch.randint(0, 6, (K // 32, 2 * N, 4), generator=g, device=DEV)]
        m16 = quads[..., 0] | (quads[..., 1] << 4) | (quads[..., 2] << 8) | (quads[..., 3] << 12)
        ms.append(m16.to(torch.int16))
    s = (torch.rand((K // 128, N), generator=g, device=DEV) * 0.01 + 0.005).half()
    wk = torch.zeros((N // 128) * 64, dtype=torch.int32, device=DEV)
    for M in MS:
        a = torch.randn((M, K), device=DEV, dtype=torch.float16)
        for i in range(3):
            ops.gptq_marlin_24_gemm(a, qs[i % ncopy], ms[i % ncopy], s, wk, 4, M, N, K)
        torch.cuda.synchronize()
        reps = ncopy * 2
        st = torch.cuda.Stream()
        with torch.cuda.stream(st):
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=st):
                for i in range(reps):
                    ops.gptq_marlin_24_gemm(a, qs[i % ncopy], ms[i % ncopy], s, wk, 4, M, N, K)
            for _ in range(5):
                gr.replay()
            torch.cuda.synchronize()
            evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(10)]
            for e0, e1 in evs:
                e0.record()
                gr.replay()
                e1.record()
            torch.cuda.synchronize()
        ts = sorted(e0.elapsed_time(e1) / reps for e0, e1 in evs)
        med = ts[len(ts) // 2] * 1e-3
        alg = K * N / 4 + K * N / 8 + (K / 128) * N * 2 + 2 * M * (K + N)
        gbs = alg / med / 1e9
        res[f"{name}_M{M}"] = dict(us=med * 1e6, gbs=gbs, frac=gbs / peak)
        print(f"2:4 {name:8s} M={M:3d} K={K} N={N}: {med*1e6:8.2f} us  {gbs:7.1f} GB/s  {gbs/peak*100:5.1f}% of measured {peak:.0f}" + (f"  {2*M*N*K/med/1e12:7.1f} TFLOP/s" if M >= 256 else ""), flush=True)
    del qs, ms
json.dump(res, open("gpurun_out/m24_perf.json", "w"), indent=1)
