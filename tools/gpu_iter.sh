#!/bin/bash
# iteration pass: marlin parity tests, per-shape GEMM numbers, role counters
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_marlin.py -m gpu -x -q > gpurun_out/pytest_marlin.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_marlin.log
timeout 300 python tools/gemm_perf.py 1,16,64 bf16 > gpurun_out/gemm_perf.log 2>&1; echo "gemm_perf rc=$?"; cat gpurun_out/gemm_perf.log | tail -14
true
