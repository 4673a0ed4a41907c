#!/bin/bash
# ablations of the W4A16 GEMM (build/variants/ablate): which stage bounds the unit time
mkdir -p gpurun_out
export B200_LIB_OPS=build/variants/ablate/libb200ops.so
for e in 0 1 2 3 4 5 6; do
  echo "=== B200_GEMM_EXP=$e"
  B200_GEMM_EXP=$e timeout 200 python tools/gemm_perf.py 16,64 bf16 2>&1 | grep "M=" 
done > gpurun_out/gemm_ablate.txt 2>&1
cat gpurun_out/gemm_ablate.txt
