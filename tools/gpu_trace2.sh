#!/bin/bash
mkdir -p gpurun_out
export B200_LIB_OPS=build/variants/prof/libb200ops.so
for cfg in "4096 4096 16" "4096 28672 16"; do
  echo "=== $cfg"; timeout 120 python tools/gemm_trace.py $cfg 5 > /tmp/tr.log 2>&1; python tools/trace_fmt.py /tmp/tr.log | head -120; grep "gemm prof" /tmp/tr.log | tail -1
done > gpurun_out/gemm_trace2.txt 2>&1
cat gpurun_out/gemm_trace2.txt | head -150
