#!/bin/bash
# decode-step pass: bench line + ncu launch list of ONE decode step (eager, profiler-bracketed)
mkdir -p gpurun_out
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench.log
timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --profile-step > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --profile-step > gpurun_out/ncu_launch.log 2>&1
echo "ncu launches rc=$?"
