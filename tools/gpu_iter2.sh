#!/bin/bash
# iteration pass: role counters + one full ncu capture of the gate_up GEMM
mkdir -p gpurun_out
timeout 300 python tools/gemm_prof.py > gpurun_out/gemm_prof.log 2>&1; echo "prof rc=$?"; grep "gemm prof" gpurun_out/gemm_prof.log | tail -12
timeout 300 python tools/gemm_one.py 4096 28672 16 6 > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:marlin_gemm_kernel -s 3 -c 2 -o gpurun_out/prof_gemm2 python tools/gemm_one.py 4096 28672 16 6 > gpurun_out/ncu_full2.log 2>&1
echo "ncu rc=$?"
