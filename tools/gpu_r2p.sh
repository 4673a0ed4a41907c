#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_scaled_mm.py -m gpu -x -q 2>&1 | tail -2
timeout 400 python tools/smm_perf.py > gpurun_out/smm_perf.txt 2>&1; grep "^fp8" gpurun_out/smm_perf.txt | head -24
