#!/bin/bash
mkdir -p gpurun_out
python tools/gemm_shapes_once.py > gpurun_out/plain_g.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,dram__throughput.avg.pct_of_peak_sustained_elapsed,dram__bytes_read.sum,sm__inst_executed.sum,smsp__warps_eligible.avg.per_cycle_active --clock-control none -k regex:marlin_gemm_kernel --csv --log-file gpurun_out/r2_ncu_marlin_gemm.csv python tools/gemm_shapes_once.py > gpurun_out/ncu_g.log 2>&1
echo "gemm rc=$?"
python tools/attn_perf.py 64x2048x32x8,64x8192x8x1 fp8 > gpurun_out/plain_a.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:paged_attn_gqa -s 6 -c 1 -f -o gpurun_out/attn_fp8 python tools/attn_perf.py 64x2048x32x8 fp8 > gpurun_out/ncu_a.log 2>&1
echo "attn rc=$?"; tail -2 gpurun_out/ncu_a.log
