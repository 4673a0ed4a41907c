#!/bin/bash
# 2-GPU pass: all-reduce + fused GEMM/all-reduce parity and timing, then the TP=2 bench fused / unfused / NCCL
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29533 tools/ar_test.py > gpurun_out/ar_test_w$N.log 2>&1; echo "ar_test rc=$?"; grep '^{' gpurun_out/ar_test_w$N.log | tail -1; tail -3 gpurun_out/ar_test_w$N.log | cut -c1-300
timeout 240 $TR --master-port 29534 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_tp${N}_fused.log 2>&1; echo "bench fused rc=$?"; grep '^{' gpurun_out/bench_tp${N}_fused.log | cut -c1-600; tail -2 gpurun_out/bench_tp${N}_fused.log | cut -c1-300
timeout 240 $TR --master-port 29535 bench.py --gpus $N --steps 20 --warmup 5 --no-fuse > gpurun_out/bench_tp${N}_custom.log 2>&1; echo "bench unfused rc=$?"; grep '^{' gpurun_out/bench_tp${N}_custom.log | cut -c1-300
true
