#!/bin/bash
# baseline pass: whole GPU parity suite, per-shape GEMM numbers, role counters, bench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python tools/gemm_perf.py 1,16,64 bf16 > gpurun_out/gemm_perf.log 2>&1; echo "gemm_perf rc=$?"; tail -14 gpurun_out/gemm_perf.log
timeout 300 python tools/gemm_prof.py > gpurun_out/gemm_prof.log 2>&1; echo "prof rc=$?"
timeout 600 python bench.py > gpurun_out/bench.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench.log
