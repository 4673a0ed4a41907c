#!/bin/bash
# round-end pass: whole GPU parity suite, smoke, bench (+ reference arm), ncu launch list of one decode step, ncu traffic of the GEMM
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench.log 2>gpurun_out/bench.err; echo "bench rc=$?"; tail -1 gpurun_out/bench.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.log 2>&1; echo "bench ref rc=$?"; tail -1 gpurun_out/bench_ref.log
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-sweep --profile-step > gpurun_out/plain_step.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches_step.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-sweep --profile-step > gpurun_out/ncu_step.log 2>&1
echo "ncu launches rc=$?"; python tools/launch_summary.py gpurun_out/launches_step.csv | tail -12
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,dram__throughput.avg.pct_of_peak_sustained_elapsed --clock-control none --profile-from-start off -k regex:marlin_gemm_kernel -s 8 -c 8 --csv --log-file gpurun_out/gemm_traffic.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-sweep --profile-step > gpurun_out/ncu_traffic.log 2>&1
echo "ncu traffic rc=$?"; tail -20 gpurun_out/gemm_traffic.csv | cut -c1-220
