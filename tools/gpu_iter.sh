#!/bin/bash
# iteration pass: marlin parity tests, per-shape GEMM numbers, role counters
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_marlin.py -m gpu -x -q > gpurun_out/pytest_marlin.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_marlin.log
timeout 300 python tools/gemm_perf.py 1,16,32,64 bf16 > gpurun_out/gemm_perf.log 2>&1; echo "gemm_perf rc=$?"; cat gpurun_out/gemm_perf.log | tail -18
timeout 300 python tools/gemm_prof.py > gpurun_out/gemm_prof.log 2>&1; echo "prof rc=$?"; grep "gemm prof" gpurun_out/gemm_prof.log | awk 'NR%2==0' | tail -6
