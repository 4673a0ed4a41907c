"""The reference's custom all-reduce OP surface (torch.ops._b200_C_custom_ar.*, same call sequence as
vllm/distributed/device_communicators/custom_all_reduce.py:104-262) on world-size ranks, against the fp32 sum through NCCL.
Run under torchrun (tests/test_gpu_allreduce.py does, on >= 2 GPUs)."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nm_vllm_b200 import _custom_ops as ops  # noqa: E402

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
dev = torch.device("cuda", torch.cuda.current_device())
dist.init_process_group("nccl", device_id=dev)
max_size = 1024 * 1024
meta = torch.zeros(ops.meta_size() + max_size, dtype=torch.uint8, device=dev)
# IPC handle + offset of the tensor's allocation (what the reference's _get_ipc_meta obtains through torch's storage
# sharing; taken from the CUDA runtime / driver directly here so that the test does not depend on torch's handle format)
import ctypes
rt = ctypes.CDLL("libcudart.so.12") if os.path.exists("/usr/local/cuda/lib64/libcudart.so.12") else ctypes.CDLL("libcudart.so")
drv = ctypes.CDLL("libcuda.so.1")
base, size = ctypes.c_uint64(), ctypes.c_size_t()
assert drv.cuMemGetAddressRange_v2(ctypes.byref(base), ctypes.byref(size), ctypes.c_uint64(meta.data_ptr())) == 0
hbuf = ctypes.create_string_buffer(64)
assert rt.cudaIpcGetMemHandle(hbuf, ctypes.c_void_p(base.value)) == 0
handle, offset = hbuf.raw, meta.data_ptr() - base.value
gathered = [None] * world
dist.all_gather_object(gathered, (bytes(handle), int(offset)))
handles, offsets = [g[0] for g in gathered], [g[1] for g in gathered]
rank_data = torch.empty(8 * 1024 * 1024, dtype=torch.uint8, device=dev)
fa = ops.init_custom_ar(meta, rank_data, handles, offsets, rank, True)
dist.barrier()
fails = 0
for dtype in (torch.bfloat16, torch.float16, torch.float32):
    for n in (16, 4096, 64 * 4096, 64 * 8192):
        g = torch.Generator(device=dev).manual_seed(100 * rank + n)
        x = torch.randn(n, generator=g, device=dev).to(dtype)
        if not ops.should_custom_ar(x, max_size, world, True):
            continue
        ref = x.float().clone()
        dist.all_reduce(ref)
        out = torch.empty_like(x)
        ops.all_reduce_reg(fa, x, out)
        out2 = torch.empty_like(x)
        ops.all_reduce_unreg(fa, x, torch.empty(0, device=dev), out2)
        torch.cuda.synchronize()
        tol = ref.abs() * {torch.bfloat16: 2.0 ** -7, torch.float16: 2.0 ** -10, torch.float32: 1e-6}[dtype] + 1e-5
        if int(((out.float() - ref).abs() > tol).sum()) or not torch.equal(out, out2):
            fails += 1
t = torch.tensor([fails], device=dev)
dist.all_reduce(t)
dist.barrier()
ops.dispose(fa)
if rank == 0:
    print(json.dumps({"world": world, "fails_total": int(t)}), flush=True)
torch.cuda.synchronize()
os._exit(0)
