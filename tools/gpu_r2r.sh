#!/bin/bash
mkdir -p gpurun_out
for pf in 2 0 1 2 0 1; do
B200_ATTN_PREFETCH=$pf timeout 600 python bench.py --steps 40 --warmup 5 --no-cpu-baseline > gpurun_out/bench_pf$pf.json 2>/dev/null; python -c "import json;d=json.load(open('gpurun_out/bench_pf$pf.json'));print('prefetch=$pf', round(d['ms_per_step'],4), round(d['e2e']['ms_per_step'],4), {k:round(v) for k,v in d['tok_s_by_batch'].items()})"
done
