#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fused_attention.py tests/test_gpu_attention.py tests/test_gpu_decode_e2e.py -m gpu -x -q 2>&1 | tail -4
for w in 4 3 2; do echo "== B200_ATTN_WARPS=$w"; B200_ATTN_WARPS=$w timeout 300 python tools/attn_perf.py 64x256x32x8,64x2048x32x8,32x256x32x8,16x1024x32x8 2>&1 | tail -4; done
echo "== auto"; timeout 300 python tools/attn_perf.py 2>&1 | tail -6
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-sweep 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], d['e2e']['ms_per_step'])"
