#!/bin/bash
mkdir -p gpurun_out
{
echo "=== warm-up (discard)"; timeout 200 python tools/gemm_perf.py 64 bf16 o 2>&1 | grep "M="
echo "=== current build"; timeout 200 python tools/gemm_perf.py 1,16,64 bf16 2>&1 | grep "M="
} > gpurun_out/gemm_now.txt 2>&1
cat gpurun_out/gemm_now.txt
timeout 900 python -m pytest tests/test_gpu_marlin.py tests/test_gpu_gemm_silu.py tests/test_gpu_bench_shapes.py -m gpu -x -q > gpurun_out/pytest_marlin.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_marlin.log
timeout 600 python bench.py --steps 30 --warmup 5 > gpurun_out/bench_now.json 2> gpurun_out/bench_now.err; echo "bench rc=$?"; python -c "import json;d=json.load(open('gpurun_out/bench_now.json'));print(d['ms_per_step'], d['e2e']['ms_per_step'], sum(d['gpu_launches_per_step'].values()), d['tok_s_by_batch'], d['roofline']['frac'], d['roofline']['by_batch'], d.get('prefill_2k'))"
