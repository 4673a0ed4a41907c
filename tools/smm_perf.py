"""Kernel micro-benchmark: cutlass_scaled_mm (W8A8 fp8 / int8) time, achieved HBM GB/s (decode M) or TFLOP/s (prefill M)
on Llama-3-70B TP=8 / TP=1 shapes, cold L2 by rotating weight buffers."""
import json
import sys

import torch

sys.path.insert(0, ".")
from nm_vllm_b200 import _custom_ops as ops  # noqa: E402

DEV = "cuda:0"
SHAPES = {"qkv_tp8": (8192, 1280), "o_tp8": (1024, 8192), "gate_up_tp8": (8192, 7168), "down_tp8": (3584, 8192),
          "qkv_tp1": (8192, 10240), "down_tp1": (28672, 8192), "tiny": (128, 128), "one_wave": (128, 148 * 128)}
if len(sys.argv) > 3:
    SHAPES = {k: v for k, v in SHAPES.items() if k in sys.argv[3].split(",")}
MS = [1, 16, 64, 2048] if len(sys.argv) < 2 else [int(x) for x in sys.argv[1].split(",")]
kinds = ["fp8", "int8"] if len(sys.argv) < 3 else sys.argv[2].split(",")
peaks = {"hbm_gbs": 6524.6}
try:
    peaks = json.load(open("MEASURED_PEAKS.json"))
except Exception:
    pass
res = {}
for kind in kinds:
    for name, (K, N) in SHAPES.items():
        ncopy = max(3, int(400e6 // (K * N)) + 1)
        if kind == "fp8":
            ws_ = [torch.randn((N, K), device=DEV).to(torch.float8_e4m3fn) for _ in range(ncopy)]
        else:
            ws_ = [torch.randint(-128, 128, (N, K), device=DEV, dtype=torch.int8) for _ in range(ncopy)]
        sb = torch.rand((1, N), device=DEV) * 0.01
        for M in MS:
            if M > 64 and "tp1" in name and K > 10000:
                pass
            a = torch.randn((M, K), device=DEV)
            a = a.to(torch.float8_e4m3fn) if kind == "fp8" else (a * 20).round().clamp(-128, 127).to(torch.int8)
            sa = torch.rand((M, 1), device=DEV) * 0.01
            for i in range(3):
                ops.cutlass_scaled_mm(a, ws_[i % ncopy].t(), sa, sb, torch.bfloat16)
            torch.cuda.synchronize()
            reps = ncopy * 2
            st = torch.cuda.Stream()
            with torch.cuda.stream(st):
                gr = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gr, stream=st):
                    for i in range(reps):
                        ops.cutlass_scaled_mm(a, ws_[i % ncopy].t(), sa, sb, torch.bfloat16)
                for _ in range(5):
                    gr.replay()
                torch.cuda.synchronize()
                evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(10)]
                for s, t in evs:
                    s.record()
                    gr.replay()
                    t.record()
                torch.cuda.synchronize()
            ts = sorted(s.elapsed_time(t) / reps for s, t in evs)
            med = ts[len(ts) // 2] * 1e-3
            alg = K * N + M * K + 2 * M * N + 4 * (M + N)
            gbs = alg / med / 1e9
            tf = 2.0 * M * N * K / med / 1e12
            res[f"{kind}_{name}_M{M}"] = dict(us=med * 1e6, gbs=gbs, tflops=tf)
            print(f"{kind:4s} {name:12s} M={M:5d} K={K:5d} N={N:5d}: {med*1e6:9.2f} us  {gbs:7.1f} GB/s "
                  f"({gbs/peaks['hbm_gbs']*100:5.1f}% hbm)  {tf:7.1f} TFLOP/s", flush=True)
        del ws_
json.dump(res, open("gpurun_out/smm_perf.json", "w"), indent=1)
