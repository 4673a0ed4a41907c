#!/bin/bash
# silu epilogue (balanced) A/B at every batch + silu tests + model prefill test + 2k prefill numbers
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_gemm_silu.py tests/test_gpu_prefix_prefill.py -m gpu -x -q > gpurun_out/pytest_silu2.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_silu2.log
B200_FUSE_SILU_MAX_M=64 timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_silu_on.json 2> gpurun_out/bench_silu_on.err; echo "bench rc=$?"; python -c "import json;d=json.load(open('gpurun_out/bench_silu_on.json'));print('fused', d['ms_per_step'], d['e2e']['ms_per_step'], sum(d['gpu_launches_per_step'].values()), d['tok_s_by_batch'], d.get('prefill_2k'))"
B200_NO_FUSE_SILU=1 timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_silu_off.json 2> gpurun_out/bench_silu_off.err; echo "bench rc=$?"; python -c "import json;d=json.load(open('gpurun_out/bench_silu_off.json'));print('unfused', d['ms_per_step'], d['e2e']['ms_per_step'], sum(d['gpu_launches_per_step'].values()), d['tok_s_by_batch'])"
timeout 600 python tools/prefill_perf.py 2048 w4a16,w4a16_24 > gpurun_out/prefill_perf.txt 2> gpurun_out/prefill_perf.err; echo "prefill rc=$?"; cat gpurun_out/prefill_perf.txt; tail -3 gpurun_out/prefill_perf.err
true
