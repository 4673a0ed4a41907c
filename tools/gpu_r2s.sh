#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_attention.py tests/test_gpu_fused_attention.py tests/test_gpu_decode_e2e.py -m gpu -x -q 2>&1 | tail -2
timeout 300 python tools/attn_perf.py > gpurun_out/attn_perf_bf16.txt 2>&1; grep "^attn" gpurun_out/attn_perf_bf16.txt
timeout 300 python tools/attn_perf.py 64x256x32x8,64x2048x32x8,8x8192x32x8,64x8192x8x1,1x32768x32x8,64x32768x8x1 fp8 > gpurun_out/attn_perf_fp8.txt 2>&1; grep "^attn" gpurun_out/attn_perf_fp8.txt
timeout 600 python bench.py > gpurun_out/bench.log 2>gpurun_out/bench.err; tail -1 gpurun_out/bench.log | python -c "
import json,sys;d=json.loads(sys.stdin.read());print(d['ms_per_step'],d['e2e']['ms_per_step'],d['roofline']['frac'],d['tok_s_by_batch'],d['prefill_2k'])"
