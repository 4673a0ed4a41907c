#!/bin/bash
# fused attention + e2e parity, then the default bench line
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fused_attention.py tests/test_gpu_decode_e2e.py tests/test_gpu_attention.py -m gpu -x -q > gpurun_out/pytest_fused.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_fused.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_n1.err; cat gpurun_out/bench_n1.json
