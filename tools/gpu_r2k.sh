#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fused_quant.py -m gpu -x -q > gpurun_out/pytest_fq.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_fq.log
for fq in 1 0; do
B200_FUSE_ACT_QUANT=$fq timeout 600 python bench.py --model llama-3-8b --quant fp8 --kv fp8 --steps 20 --warmup 5 --no-cpu-baseline --no-sweep > gpurun_out/bench_fp8_dyn_$fq.json 2>gpurun_out/bench_fp8_dyn_$fq.err; python -c "
import json;d=json.load(open('gpurun_out/bench_fp8_dyn_$fq.json'));print('fuse=$fq', d['ms_per_step'], sum(d['gpu_launches_per_step'].values()), d['gpu_launches_per_step'])"
done
