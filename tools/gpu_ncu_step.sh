#!/bin/bash
# ncu launch list of one decode step (device time per launch, cold-cache and serialised: compare SHARES)
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-sweep --profile-step > gpurun_out/plain_step.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches_step.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-sweep --profile-step > gpurun_out/ncu_step.log 2>&1
echo rc=$?; tail -2 gpurun_out/ncu_step.log; python tools/launch_summary.py gpurun_out/launches_step.csv | tail -30
