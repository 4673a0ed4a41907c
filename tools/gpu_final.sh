#!/bin/bash
# round-end pass: whole GPU parity suite, smoke, per-shape GEMM numbers, bench (+ reference arm), ncu launch list of one
# decode step and one full capture of the dominant kernel
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 300 python tools/gemm_perf.py 1,16,64 bf16 > gpurun_out/gemm_perf.log 2>&1; echo "gemm_perf rc=$?"
timeout 600 python bench.py > gpurun_out/bench.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.log 2>&1; echo "bench ref rc=$?"; tail -1 gpurun_out/bench_ref.log
timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --profile-step > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --profile-step > gpurun_out/ncu_launch.log 2>&1
echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:marlin_gemm_kernel -s 6 -c 4 -f -o gpurun_out/prof_gemm_step python bench.py --steps 1 --warmup 3 --no-cpu-baseline --profile-step > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
