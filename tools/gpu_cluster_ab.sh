#!/bin/bash
mkdir -p gpurun_out
{
echo "=== warm-up (discard)"; timeout 200 python tools/gemm_perf.py 64 bf16 o 2>&1 | grep "M="
echo "=== cluster split-K where tiles*2 <= SMs (default)"; timeout 200 python tools/gemm_perf.py 1,16,64 bf16 qkv,o,down 2>&1 | grep "M="
echo "=== B200_GEMM_NO_CLUSTER=1 (stream-K everywhere)"; B200_GEMM_NO_CLUSTER=1 timeout 200 python tools/gemm_perf.py 1,16,64 bf16 qkv,o,down 2>&1 | grep "M="
} > gpurun_out/gemm_cluster_ab.txt 2>&1
cat gpurun_out/gemm_cluster_ab.txt
