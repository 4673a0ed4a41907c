#!/bin/bash
mkdir -p gpurun_out
{
echo "=== warm-up (discard)"; timeout 200 python tools/gemm_perf.py 64 bf16 o 2>&1 | grep "M="
echo "=== baseline"; timeout 200 python tools/gemm_perf.py 16,64 bf16 2>&1 | grep "M="
echo "=== variant spin"; B200_LIB_OPS=build/variants/spin/libb200ops.so timeout 200 python tools/gemm_perf.py 16,64 bf16 2>&1 | grep "M="
for e in 1 2 3 4 5 6; do
  echo "=== ablate B200_GEMM_EXP=$e"
  B200_LIB_OPS=build/variants/ablate/libb200ops.so B200_GEMM_EXP=$e timeout 200 python tools/gemm_perf.py 16,64 bf16 gate_up,down 2>&1 | grep "M="
done
} > gpurun_out/gemm_ablate2.txt 2>&1
cat gpurun_out/gemm_ablate2.txt
