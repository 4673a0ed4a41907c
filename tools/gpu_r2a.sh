#!/bin/bash
# round-2 iteration pass for the decode GEMM: quick smoke, marlin parity tests, per-shape numbers (variants)
mkdir -p gpurun_out
B200_DEC_VERBOSE=1 timeout 120 python tools/gemm_check.py > gpurun_out/gemm_check.log 2>&1; echo "gemm_check rc=$?"; tail -16 gpurun_out/gemm_check.log
timeout 900 python -m pytest tests/test_gpu_marlin.py -m gpu -x -q > gpurun_out/pytest_marlin.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_marlin.log
timeout 300 python tools/gemm_perf.py 1,16,64 bf16 > gpurun_out/gemm_perf_new.log 2>&1; echo "gemm_perf rc=$?"; tail -12 gpurun_out/gemm_perf_new.log
for v in "$@"; do
  echo "== $v"; env $v timeout 300 python tools/gemm_perf.py 1,16,64 bf16 2>&1 | tail -12
done
true
