#!/bin/bash
# 2-GPU pass: one-shot all-reduce parity/timing, then the TP=2 bench with the custom kernel and with NCCL
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 240 $TR --master-port 29533 tools/ar_test.py > gpurun_out/ar_test.log 2>&1; echo "ar_test rc=$?"; grep '^{' gpurun_out/ar_test.log | tail -1; tail -3 gpurun_out/ar_test.log | cut -c1-300
timeout 240 $TR --master-port 29534 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench_tp2_custom.log 2>&1; echo "bench custom rc=$?"; grep '^{' gpurun_out/bench_tp2_custom.log | cut -c1-700; tail -2 gpurun_out/bench_tp2_custom.log | cut -c1-300
timeout 240 $TR --master-port 29535 bench.py --gpus 2 --steps 20 --warmup 5 --allreduce nccl > gpurun_out/bench_tp2_nccl.log 2>&1; echo "bench nccl rc=$?"; grep '^{' gpurun_out/bench_tp2_nccl.log | cut -c1-400
true
