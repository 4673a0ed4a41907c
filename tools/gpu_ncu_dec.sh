#!/bin/bash
# one ncu --set full capture (with source counters) of the decode GEMM on gate_up M=$1
mkdir -p gpurun_out
M=${1:-16}
python tools/gemm_one.py 4096 28672 $M 4 > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:marlin_decode3 -s 2 -c 1 -f -o gpurun_out/dec3_m$M python tools/gemm_one.py 4096 28672 $M 4 > gpurun_out/ncu.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/ncu.log
