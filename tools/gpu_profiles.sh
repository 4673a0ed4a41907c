#!/bin/bash
# round-2 evidence pass (kept under profiles/): kernel micro-benchmarks, 2:4 e2e, prefill GEMMs, config 1 on the host cores
mkdir -p gpurun_out
timeout 600 python tools/attn_perf.py > gpurun_out/attn_perf_bf16.txt 2>&1; cp gpurun_out/attn_perf.json gpurun_out/attn_perf_bf16.json; tail -7 gpurun_out/attn_perf_bf16.txt
timeout 600 python tools/attn_perf.py 64x256x32x8,64x2048x32x8,8x8192x32x8,64x8192x8x1,1x32768x32x8,64x32768x8x1 fp8 > gpurun_out/attn_perf_fp8.txt 2>&1; cp gpurun_out/attn_perf.json gpurun_out/attn_perf_fp8.json; tail -7 gpurun_out/attn_perf_fp8.txt
timeout 600 python tools/m24_perf.py 1,16,64,2048 > gpurun_out/m24_perf.txt 2>&1; tail -17 gpurun_out/m24_perf.txt
timeout 600 python tools/smm_perf.py 1,16,64,2048 fp8,int8 > gpurun_out/smm_perf.txt 2>&1; tail -40 gpurun_out/smm_perf.txt
timeout 600 python tools/gemm_perf.py 1,16,64,2048 bf16 > gpurun_out/gemm_perf_w4a16.txt 2>&1; tail -17 gpurun_out/gemm_perf_w4a16.txt
timeout 600 python bench.py --quant w4a16_24 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n1_w4a16_24.json 2> gpurun_out/bench_24.err; tail -2 gpurun_out/bench_24.err; cut -c1-900 gpurun_out/bench_n1_w4a16_24.json
timeout 900 python -m oracle.cpu_opt125m > gpurun_out/config1_opt125m_cpu.json 2>&1; cat gpurun_out/config1_opt125m_cpu.json
timeout 900 python -m oracle.cpu_opt125m --w4a16 > gpurun_out/config1_opt125m_cpu_w4a16.json 2>&1; cat gpurun_out/config1_opt125m_cpu_w4a16.json
