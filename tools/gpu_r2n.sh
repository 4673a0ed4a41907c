#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/attn_perf.py > gpurun_out/attn_perf_bf16.txt 2>&1; tail -6 gpurun_out/attn_perf_bf16.txt
for i in 1 2; do timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_pf.json 2>/dev/null; python -c "import json;d=json.load(open('gpurun_out/bench_pf.json'));print(d['ms_per_step'], d['e2e']['ms_per_step'], d['tok_s_by_batch'])"; done
