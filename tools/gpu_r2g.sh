#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_prefix_prefill.py -m gpu -x -q > gpurun_out/pytest_prefill2.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_prefill2.log
timeout 600 python tools/prefill_perf.py 2048 w4a16 > gpurun_out/prefill_perf.txt 2> gpurun_out/prefill_perf.err; echo "prefill rc=$?"; cat gpurun_out/prefill_perf.txt; tail -3 gpurun_out/prefill_perf.err
true
