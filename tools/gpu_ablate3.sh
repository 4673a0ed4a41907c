#!/bin/bash
mkdir -p gpurun_out
{
echo "=== warm-up (discard)"; timeout 200 python tools/gemm_perf.py 64 bf16 o 2>&1 | grep "M="
for e in 0 7 9; do
  echo "=== ablate B200_GEMM_EXP=$e"
  B200_LIB_OPS=build/variants/ablate/libb200ops.so B200_GEMM_EXP=$e timeout 200 python tools/gemm_perf.py 16,64 bf16 2>&1 | grep "M="
done
} > gpurun_out/gemm_ablate3.txt 2>&1
cat gpurun_out/gemm_ablate3.txt
