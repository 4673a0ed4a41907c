#!/bin/bash
mkdir -p gpurun_out
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o /tmp/tma_bw tools/tma_bw.cu -lcuda 2>&1 | tail -3
timeout 300 /tmp/tma_bw > gpurun_out/tma_bw.log 2>&1
for d in 0 4 5 1; do
  echo "== DBG=$d" >> gpurun_out/gemm_dbg.log
  B200_GEMM_DBG=$d timeout 200 python tools/gemm_perf.py 1,64 bf16 qkv,gate_up >> gpurun_out/gemm_dbg.log 2>&1
done
cat gpurun_out/gemm_dbg.log
grep -E "nst=(16|24)" gpurun_out/tma_bw.log
