#!/bin/bash
# prefill attention tests, silu tests, A/B of the silu fusion in the decode step
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_prefix_prefill.py tests/test_gpu_gemm_silu.py tests/test_gpu_linear_methods.py -m gpu -x -q > gpurun_out/pytest_prefill.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_prefill.log
timeout 600 python bench.py --steps 30 --warmup 5 > gpurun_out/bench_silu_on.json 2> gpurun_out/bench_silu_on.err; echo "bench rc=$?"; python -c "import json;d=json.load(open('gpurun_out/bench_silu_on.json'));print('fused', d['ms_per_step'], d['e2e']['ms_per_step'], sum(d['gpu_launches_per_step'].values()), d['tok_s_by_batch'])"
B200_NO_FUSE_SILU=1 timeout 600 python bench.py --steps 30 --warmup 5 > gpurun_out/bench_silu_off.json 2> gpurun_out/bench_silu_off.err; echo "bench rc=$?"; python -c "import json;d=json.load(open('gpurun_out/bench_silu_off.json'));print('unfused', d['ms_per_step'], d['e2e']['ms_per_step'], sum(d['gpu_launches_per_step'].values()), d['tok_s_by_batch'])"
true
