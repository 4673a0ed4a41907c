#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_decode_e2e.py tests/test_gpu_fused_attention.py -m gpu -x -q -k "glue or norm or decode" 2>&1 | tail -2
timeout 600 python -m pytest tests -m gpu -x -q -k "glue_kernels or rms_norm" 2>&1 | tail -2
timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_norm.json 2>/dev/null; python -c "import json;d=json.load(open('gpurun_out/bench_norm.json'));print(d['ms_per_step'], d['e2e']['ms_per_step'], d['tok_s_by_batch'])"
