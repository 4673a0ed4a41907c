"""Kernel micro-benchmark: paged_attention_v1/v2 decode, achieved HBM GB/s (KV counted once per kv head), cache > L2."""
import json
import sys

import torch

sys.path.insert(0, ".")
from nm_vllm_b200 import _custom_ops as ops  # noqa: E402

DEV = "cuda:0"
peak = 6524.6
try:
    peak = json.load(open("MEASURED_PEAKS.json"))["hbm_gbs"]
except Exception:
    pass
CASES = [(64, 256, 32, 8), (64, 2048, 32, 8), (8, 8192, 32, 8), (64, 8192, 8, 1), (1, 32768, 32, 8), (64, 32768, 8, 1)]
if len(sys.argv) > 1:
    CASES = [tuple(int(x) for x in c.split("x")) for c in sys.argv[1].split(",")]
D, BS = 128, 16
KVD = sys.argv[2] if len(sys.argv) > 2 else "auto"
EB = 2 if KVD == "auto" else 1
dt = torch.bfloat16
res = {}
for S, ctx, H, Hkv in CASES:
    nb_seq = (ctx + BS - 1) // BS
    NB = max(S * nb_seq, int(1.0e9 // (Hkv * D * BS * 2 * EB)))   # >= 1 GB of cache so that L2 cannot hold it
    NB = min(NB, int(24e9 // (Hkv * D * BS * 2 * EB)))
    if KVD == "auto":
        kc = torch.empty((NB, Hkv, D // 8, BS, 8), dtype=dt, device=DEV).uniform_(-0.09, 0.09)
        vc = torch.empty((NB, Hkv, D, BS), dtype=dt, device=DEV).uniform_(-0.09, 0.09)
    else:
        kc = torch.randint(0, 120, (NB, Hkv, D // 16, BS, 16), dtype=torch.uint8, device=DEV)
        vc = torch.randint(0, 120, (NB, Hkv, D, BS), dtype=torch.uint8, device=DEV)
    q = torch.empty((S, H, D), dtype=dt, device=DEV).uniform_(-0.09, 0.09)
    bt = torch.randperm(NB, device=DEV)[: S * nb_seq].to(torch.int32).reshape(S, nb_seq) if NB >= S * nb_seq else \
        torch.randint(0, NB, (S, nb_seq), dtype=torch.int32, device=DEV)
    sl = torch.full((S,), ctx, dtype=torch.int32, device=DEV)
    out = torch.empty_like(q)
    P = (ctx + 511) // 512
    use_v1 = ctx <= 8192 and (P == 1 or S * H > 512)   # paged_attn.py:120-121
    tmp = torch.empty((S, H, P, D), dtype=dt, device=DEV)
    es = torch.empty((S, H, P), dtype=torch.float32, device=DEV)
    ml = torch.empty_like(es)
    scale = D ** -0.5

    def run():
        if use_v1:
            ops.paged_attention_v1(out, q, kc, vc, Hkv, scale, bt, sl, BS, ctx, None, KVD, 0.0376)
        else:
            ops.paged_attention_v2(out, es, ml, tmp, q, kc, vc, Hkv, scale, bt, sl, BS, ctx, None, KVD, 0.0376)
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    # REPS launches replayed from one CUDA graph: removes the host launch latency that dominates a single ~20 us launch
    REPS = 20
    st = torch.cuda.Stream()
    st.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(st):
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=st):
            for _ in range(REPS):
                run()
        gr.replay()
        torch.cuda.synchronize()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(10)]
        for e0, e1 in evs:
            e0.record(st); gr.replay(); e1.record(st)
        torch.cuda.synchronize()
    ts = sorted(e0.elapsed_time(e1) / REPS for e0, e1 in evs)
    med = ts[len(ts) // 2] * 1e-3
    alg = S * ctx * Hkv * D * 2 * EB + 2 * S * H * D * 2 + 4 * S * nb_seq
    gbs = alg / med / 1e9
    name = f"S{S}_ctx{ctx}_H{H}_Hkv{Hkv}_{KVD}"
    res[name] = dict(us=med * 1e6, gbs=gbs, frac=gbs / peak, v=1 if use_v1 else 2)
    print(f"attn {name:26s} v{1 if use_v1 else 2}: {med*1e6:9.1f} us  {gbs:7.1f} GB/s  {gbs/peak*100:5.1f}% of measured {peak:.0f}", flush=True)
    del kc, vc
json.dump(res, open("gpurun_out/attn_perf.json", "w"), indent=1)
