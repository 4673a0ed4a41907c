#!/bin/bash
mkdir -p gpurun_out
{
echo "=== warm-up (discard)"; timeout 200 python tools/gemm_perf.py 64 bf16 o 2>&1 | grep "M="
echo "=== baseline"; timeout 200 python tools/gemm_perf.py 1,16,64 bf16 2>&1 | grep "M="
for v in "$@"; do
echo "=== variant $v"; B200_LIB_OPS=build/variants/$v/libb200ops.so timeout 200 python tools/gemm_perf.py 1,16,64 bf16 2>&1 | grep "M="
done
} > gpurun_out/gemm_variants.txt 2>&1
cat gpurun_out/gemm_variants.txt
