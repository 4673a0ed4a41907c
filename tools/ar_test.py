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
nf = torch.tensor([len(fails)], device=dev)
dist.all_reduce(nf)
if rank == 0:
    print(json.dumps({"world": world, "fails_total": int(nf), "fails_rank0": fails[:5], "us_per_call": res}), flush=True)
dist.barrier()
torch.cuda.synchronize()
sys.stdout.flush()
os._exit(0 if int(nf) == 0 else 1)
