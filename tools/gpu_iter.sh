#!/bin/bash
# iteration pass: marlin parity tests, per-shape GEMM numbers, bench line
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_marlin.py -m gpu -x -q > gpurun_out/pytest_marlin.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_marlin.log
timeout 300 python tools/gemm_perf.py 1,16,64 bf16 > gpurun_out/gemm_perf.log 2>&1; echo "gemm_perf rc=$?"; cat gpurun_out/gemm_perf.log | tail -14
B200_NO_PDL=1 timeout 300 python tools/gemm_perf.py 1,64 bf16 > gpurun_out/gemm_perf_nopdl.log 2>&1; echo "nopdl:"; cat gpurun_out/gemm_perf_nopdl.log | tail -9
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_iter.json 2> gpurun_out/bench_iter.err; echo "bench rc=$?"; cat gpurun_out/bench_iter.json; tail -3 gpurun_out/bench_iter.err
