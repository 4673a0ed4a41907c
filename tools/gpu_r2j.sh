#!/bin/bash
mkdir -p gpurun_out
{
echo "=== warm-up (discard)"; timeout 200 python tools/gemm_perf.py 64 bf16 o 2>&1 | grep "M="
echo "=== current build"; timeout 300 python tools/gemm_perf.py 16,64,2048 bf16 2>&1 | grep "M="
} > gpurun_out/gemm_now.txt 2>&1
cat gpurun_out/gemm_now.txt
timeout 600 python tools/prefill_perf.py 2048 w4a16 > gpurun_out/prefill_perf.txt 2> gpurun_out/prefill_perf.err; echo "prefill rc=$?"; cat gpurun_out/prefill_perf.txt; tail -3 gpurun_out/prefill_perf.err
timeout 600 python -m pytest tests/test_gpu_marlin.py -m gpu -x -q -k "128 or 200 or 256 or prefill or large" > gpurun_out/pytest_m128.log 2>&1; tail -3 gpurun_out/pytest_m128.log
