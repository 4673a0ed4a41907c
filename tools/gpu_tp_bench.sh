#!/bin/bash
# TP=N bench variants: fused two-shot (default), fused one-shot, unfused
N=$1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 240 $TR --master-port 29534 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_tp${N}_fused.log 2>&1; echo "bench fused rc=$?"; grep '^{' gpurun_out/bench_tp${N}_fused.log | cut -c1-520; tail -1 gpurun_out/bench_tp${N}_fused.log | cut -c1-200
B200_AR_TWO_SHOT_BYTES=100000000 B200_AR_FUSE_MAX_BYTES=100000000 timeout 240 $TR --master-port 29535 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_tp${N}_fused_oneshot.log 2>&1; echo "bench fused one-shot rc=$?"; grep '^{' gpurun_out/bench_tp${N}_fused_oneshot.log | cut -c1-230
timeout 240 $TR --master-port 29536 bench.py --gpus $N --steps 20 --warmup 5 --no-fuse > gpurun_out/bench_tp${N}_custom.log 2>&1; echo "bench unfused rc=$?"; grep '^{' gpurun_out/bench_tp${N}_custom.log | cut -c1-230
true
