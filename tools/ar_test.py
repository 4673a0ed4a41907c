"""One-shot NVLink all-reduce: parity + timing, run under torchrun (world 2/4/8).
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/ar_test.py
Checks (reference test: tests/distributed/test_custom_all_reduce.py:23-80): result == fp32 sum of the ranks' inputs rounded
once (exact), identical bits on every rank, eager and CUDA-graph replay with changing inputs; prints us per call vs NCCL."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nm_vllm_b200 import load_extension  # noqa: E402
from nm_vllm_b200.custom_all_reduce import CustomAllreduce  # noqa: E402

load_extension()
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
ar = CustomAllreduce(None, dev)
fails = []


def inputs(shape, dtype, seed):
    """every rank can rebuild every rank's input (same generator seeds) -> local oracle"""
    xs = []
    for r in range(world):
        g = torch.Generator(device=dev).manual_seed(1000 * seed + r)
        xs.append(torch.randn(shape, generator=g, device=dev, dtype=torch.float32).to(dtype))
    return xs


def oracle(xs):
    acc = torch.zeros_like(xs[0], dtype=torch.float32)
    for x in xs:          # rank order, fp32 accumulation, one rounding
        acc += x.float()
    return acc.to(xs[0].dtype)


for dtype in (torch.bfloat16, torch.float16, torch.float32):
    for shape in ((1, 4096), (64, 4096), (64, 8192), (7, 1024), (3, 8), (257, 4096)):
        xs = inputs(shape, dtype, seed=shape[0] + shape[1])
        out = ar.all_reduce(xs[rank].clone())
        if not torch.equal(out, oracle(xs)):
            fails.append(("eager", str(dtype), shape))
        t = xs[rank].clone()
        ar.all_reduce_(t)
        if not torch.equal(t, oracle(xs)):
            fails.append(("inplace", str(dtype), shape))
        gathered = [torch.empty_like(out) for _ in range(world)]
        dist.all_gather(gathered, out)
        if any(not torch.equal(g, out) for g in gathered):
            fails.append(("ranks differ", str(dtype), shape))

# CUDA graph: 6 chained in-place calls per replay on a buffer refilled before each replay
buf = torch.zeros((64, 4096), dtype=torch.bfloat16, device=dev)
st = torch.cuda.Stream()
with torch.cuda.stream(st):
    g = torch.cuda.CUDAGraph()
    for _ in range(2):
        ar.all_reduce_(buf)
    torch.cuda.synchronize()
    with torch.cuda.graph(g, stream=st):
        for _ in range(6):
            ar.all_reduce_(buf)
            buf.mul_(1.0 / world)
    for it in range(50):
        xs = inputs((64, 4096), torch.bfloat16, seed=500 + it)
        buf.copy_(xs[rank])
        g.replay()
        st.synchronize()
        ref = [x.clone() for x in xs]
        for _ in range(6):
            s = oracle(ref)
            ref = [(s * (1.0 / world)).to(s.dtype) for _ in range(world)]
        if not torch.equal(buf, ref[0]):
            fails.append(("graph", it))
            break

# timing: graph of 64 calls, custom vs NCCL
res = {}
for name, fn in (("custom", lambda t: ar.all_reduce_(t)), ("nccl", lambda t: dist.all_reduce(t))):
    for shape in ((1, 4096), (16, 4096), (64, 4096), (64, 8192)):
        t = torch.zeros(shape, dtype=torch.bfloat16, device=dev)
        with torch.cuda.stream(st):
            for _ in range(3):
                fn(t)
            torch.cuda.synchronize()
            gg = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gg, stream=st):
                for _ in range(64):
                    fn(t)
            for _ in range(3):
                gg.replay()
            dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            for _ in range(10):
                gg.replay()
            e1.record(st)
            torch.cuda.synchronize()
        res["%s_%dx%d" % (name, *shape)] = e0.elapsed_time(e1) * 1e3 / 640

# fused GEMM + all-reduce (marlin_gemm_kernel FUSE): must equal GEMM -> all_reduce bit for bit, eager and replayed
from nm_vllm_b200 import _custom_ops as ops  # noqa: E402
fused_res = {}
e_i32 = torch.empty(0, dtype=torch.int32, device=dev)
for dtype in (torch.bfloat16, torch.float16):
    for (M, K, N) in ((1, 512, 4096), (16, 2048, 4096), (64, 7168, 4096), (64, 1792, 4096), (33, 4096, 8192),
                      (128, 1024, 1024), (16, 1024, 12288), (5, 4096, 128)):
        g = torch.Generator(device=dev).manual_seed(77 + rank)
        w = torch.randint(-2**31, 2**31 - 1, (K // 16, N * 2), generator=g, dtype=torch.int32, device=dev)
        sc = (torch.rand((K // 128, N), generator=g, device=dev) * 0.01 + 0.005).to(dtype)
        x = torch.randn((M, K), generator=g, device=dev).to(dtype)
        wk = torch.zeros((N // 64) * 16, dtype=torch.int32, device=dev)
        ref = ar.all_reduce(ops.gptq_marlin_gemm(x, w, sc, e_i32, e_i32, wk, 4, M, N, K, True))
        if not ar.fuse_next_gemm(M, N):
            fused_res["%s_%dx%dx%d" % (str(dtype)[6:], M, K, N)] = "not fused"
            continue
        for rep in range(3):
            if rep:
                assert ar.fuse_next_gemm(M, N)
            out = ops.gptq_marlin_gemm(x, w, sc, e_i32, e_i32, wk, 4, M, N, K, True)
            if not torch.equal(out, ref):
                fails.append(("fused", str(dtype), (M, K, N), rep, float((out.float() - ref.float()).abs().max())))
                break
        fused_res["%s_%dx%dx%d" % (str(dtype)[6:], M, K, N)] = "ok"
# timing: graph of 32 x (GEMM + all-reduce) vs 32 fused GEMMs, down-proj shard shape of this world size
Kd = max(128, (14336 // world) // 128 * 128)
g = torch.Generator(device=dev).manual_seed(5 + rank)
ws_ = [torch.randint(-2**31, 2**31 - 1, (Kd // 16, 4096 * 2), generator=g, dtype=torch.int32, device=dev) for _ in range(8)]
sc = (torch.rand((Kd // 128, 4096), generator=g, device=dev) * 0.01 + 0.005).to(torch.bfloat16)
wk = torch.zeros((4096 // 64) * 16, dtype=torch.int32, device=dev)
for M in (1, 16, 64):
    x = torch.randn((M, Kd), generator=g, device=dev).to(torch.bfloat16)
    for name in ("gemm_then_ar", "fused"):
        def one(i):
            if name == "fused" and ar.fuse_next_gemm(M, 4096):
                return ops.gptq_marlin_gemm(x, ws_[i % 8], sc, e_i32, e_i32, wk, 4, M, 4096, Kd, True)
            return ar.all_reduce_(ops.gptq_marlin_gemm(x, ws_[i % 8], sc, e_i32, e_i32, wk, 4, M, 4096, Kd, True))
        with torch.cuda.stream(st):
            for i in range(3):
                one(i)
            torch.cuda.synchronize()
            gg = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gg, stream=st):
                for i in range(32):
                    one(i)
            for _ in range(3):
                gg.replay()
            dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            for _ in range(10):
                gg.replay()
            e1.record(st)
            torch.cuda.synchronize()
        res["%s_M%d_K%d" % (name, M, Kd)] = e0.elapsed_time(e1) * 1e3 / 320
nf = torch.tensor([len(fails)], device=dev)
dist.all_reduce(nf)
if rank == 0:
    print(json.dumps({"world": world, "fails_total": int(nf), "fails_rank0": [str(f) for f in fails[:5]], "fused": fused_res, "us_per_call": res}), flush=True)
dist.barrier()
torch.cuda.synchronize()
sys.stdout.flush()
os._exit(0 if int(nf) == 0 else 1)
