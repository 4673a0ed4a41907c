#!/bin/bash
# fused attention + e2e parity, then the default bench line
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_decode_e2e.py tests/test_gpu_fused_attention.py tests/test_gpu_attention.py tests/test_gpu_cache_quant.py -m gpu -x -q > gpurun_out/pytest_fused.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_fused.log | cut -c1-300
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_n1.err; python - <<'PY'
import json
d=json.load(open("gpurun_out/bench_n1.json"))
print({k:d[k] for k in ("value","ms_per_step","gpu_launches_per_step","tok_s_by_batch")}, d["e2e"], d["roofline"]["frac"], d["roofline"]["gemm_ms_per_step"], d["roofline"]["by_batch"])
PY
