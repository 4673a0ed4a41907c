#!/bin/bash
# N-GPU pass (N = $1): multi-GPU tests, then the driver-style TP=N bench (parity self-check + 70B extras)
N=$1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
if [ "$N" = "2" ]; then timeout 600 python -m pytest tests/test_gpu_allreduce.py -m gpu -x -q > gpurun_out/pytest_ar.log 2>&1; echo "pytest allreduce rc=$?"; tail -3 gpurun_out/pytest_ar.log; fi
timeout 800 $TR --master-port 29534 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_tp$N.log 2>&1; echo "bench rc=$?"; grep '^{' gpurun_out/bench_tp$N.log | tail -1 > gpurun_out/bench_tp$N.json; python -c "
import json;d=json.load(open('gpurun_out/bench_tp$N.json'));print(d['ms_per_step'], d['value'], d.get('parity_check'), d['config'].get('allreduce'), json.dumps(d.get('extra_models_at_this_tp'))[:900], sum(d['gpu_launches_per_step'].values()))"; tail -2 gpurun_out/bench_tp$N.log | cut -c1-300
true
