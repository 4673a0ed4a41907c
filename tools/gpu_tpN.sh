#!/bin/bash
# N-GPU pass (N = $1): one-shot all-reduce parity/timing at world N, then the TP=N bench
N=$1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 240 $TR --master-port 29533 tools/ar_test.py > gpurun_out/ar_test_w$N.log 2>&1; echo "ar_test rc=$?"; grep '^{' gpurun_out/ar_test_w$N.log | tail -1; tail -2 gpurun_out/ar_test_w$N.log | cut -c1-300
timeout 300 $TR --master-port 29534 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_tp$N.log 2>&1; echo "bench rc=$?"; grep '^{' gpurun_out/bench_tp$N.log | cut -c1-600; tail -2 gpurun_out/bench_tp$N.log | cut -c1-300
true
