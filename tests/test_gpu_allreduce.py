"""One-shot NVLink all-reduce (csrc/all_reduce.cu) against the fp32-sum oracle on 2 GPUs; mirrors the reference's
tests/distributed/test_custom_all_reduce.py (eager + CUDA-graph, several sizes / dtypes).  Needs >= 2 GPUs: skipped on
a single-GPU box (the check itself lives in tools/ar_test.py so that `gpurun --gpus 2` can run it directly)."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_custom_all_reduce_world2():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29541", os.path.join(ROOT, "tools", "ar_test.py")]
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600, cwd=ROOT)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    line = [l for l in p.stdout.splitlines() if l.startswith("{")][-1]
    assert json.loads(line)["fails_total"] == 0


@pytest.mark.gpu
def test_custom_ar_op_surface_world2():
    """torch.ops._b200_C_custom_ar.* driven the way the reference's CustomAllreduce drives its ops (tools/ar_ops_test.py)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29543", os.path.join(ROOT, "tools", "ar_ops_test.py")]
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600, cwd=ROOT)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    line = [l for l in p.stdout.splitlines() if l.startswith("{")][-1]
    assert json.loads(line)["fails_total"] == 0
