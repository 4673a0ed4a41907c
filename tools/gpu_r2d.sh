#!/bin/bash
# silu-in-epilogue pass: new tests, Marlin parity (the weight / scale TMA boxes changed for every call), GEMM numbers, bench
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_gemm_silu.py -m gpu -x -q > gpurun_out/pytest_silu.log 2>&1; echo "pytest silu rc=$?"; tail -5 gpurun_out/pytest_silu.log
timeout 1200 python -m pytest tests/test_gpu_marlin.py tests/test_gpu_bench_shapes.py tests/test_gpu_decode_e2e.py tests/test_gpu_linear_methods.py -m gpu -x -q > gpurun_out/pytest_marlin.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_marlin.log
timeout 300 python tools/gemm_perf.py 1,16,64 bf16 > gpurun_out/gemm_perf.log 2>&1; echo "gemm_perf rc=$?"; cat gpurun_out/gemm_perf.log | tail -14
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_silu.json 2> gpurun_out/bench_silu.err; echo "bench rc=$?"; cat gpurun_out/bench_silu.json
true
