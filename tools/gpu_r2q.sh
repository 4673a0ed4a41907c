#!/bin/bash
mkdir -p gpurun_out
{
echo "=== warm-up"; timeout 200 python tools/smm_perf.py 64 fp8 down_tp8 2>&1 | grep "^fp8"
echo "=== stepped (current)"; timeout 300 python tools/smm_perf.py 1,16,64 fp8 qkv_tp8,gate_up_tp8,down_tp8,qkv_tp1,down_tp1 2>&1 | grep "^fp8"
echo "=== divisions (variant smmdiv)"; B200_LIB_OPS=build/variants/smmdiv/libb200ops.so timeout 300 python tools/smm_perf.py 1,16,64 fp8 qkv_tp8,gate_up_tp8,down_tp8,qkv_tp1,down_tp1 2>&1 | grep "^fp8"
echo "=== stepped (current) again"; timeout 300 python tools/smm_perf.py 1,16,64 fp8 qkv_tp8,gate_up_tp8,down_tp8,qkv_tp1,down_tp1 2>&1 | grep "^fp8"
} > gpurun_out/smm_ab.txt 2>&1
cat gpurun_out/smm_ab.txt
