#!/bin/bash
mkdir -p gpurun_out
python tools/variant_check.py
for e in 0 6 2; do B200_LIB_OPS=build/variants/ablate/libb200ops.so B200_GEMM_EXP=$e python tools/variant_check.py; done
B200_LIB_OPS=build/variants/ablate/libb200ops.so python -c "
import sys; sys.path.insert(0,'.')
import nm_vllm_b200; nm_vllm_b200.load_extension()
print([l.split()[-1] for l in open('/proc/self/maps') if 'b200ops' in l and 'r-xp' in l])"
timeout 900 python -m pytest tests/test_gpu_marlin.py tests/test_gpu_gemm_silu.py tests/test_gpu_decode_e2e.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -2
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_soname.json 2>/dev/null; python -c "import json;d=json.load(open('gpurun_out/bench_soname.json'));print(d['ms_per_step'], d['e2e']['ms_per_step'], d['roofline']['frac'], d['tok_s_by_batch'], d['prefill_2k'])"
