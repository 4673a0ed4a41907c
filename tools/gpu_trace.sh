#!/bin/bash
mkdir -p gpurun_out
export B200_LIB_OPS=build/variants/trace/libb200ops.so
for cfg in "2 16" "2 64"; do
  set -- $cfg
  echo "=== cps=$1 M=$2"; B200_DEC_CTAS_PER_SM=$1 timeout 120 python tools/dec_trace.py 4096 28672 $2 5 2>&1 | grep -A80 "rep 2" | grep "dec trace" | head -${3:-24}
done
