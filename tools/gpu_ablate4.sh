#!/bin/bash
mkdir -p gpurun_out
{
echo "=== warm-up (discard)"; timeout 200 python tools/gemm_perf.py 64 bf16 o 2>&1 | grep "M="
echo "=== production"; timeout 200 python tools/gemm_perf.py 1,16,64 bf16 2>&1 | grep "M="
echo "=== ablate B200_GEMM_EXP=8 (no units: launch skeleton only)"
B200_LIB_OPS=build/variants/ablate/libb200ops.so B200_GEMM_EXP=8 timeout 200 python tools/gemm_perf.py 1,16,64 bf16 2>&1 | grep "M="
} > gpurun_out/gemm_ablate4.txt 2>&1
cat gpurun_out/gemm_ablate4.txt
