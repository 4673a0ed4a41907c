#!/bin/bash
# the ablate variant INSTALLED as the in-tree library (no second copy of the kernels in the process)
cp build/variants/ablate/libb200ops.so nm_vllm_b200/libb200ops.so
for e in 0 1 2 3 5 6 8 9; do B200_GEMM_EXP=$e python tools/variant_check.py; done
mkdir -p gpurun_out
{
echo "=== warm-up (discard)"; timeout 200 python tools/gemm_perf.py 64 bf16 o 2>&1 | grep "M="
for e in 0 1 2 3 4 5 6 8 9; do
  echo "=== ablate build installed in-tree, B200_GEMM_EXP=$e"
  B200_GEMM_EXP=$e timeout 200 python tools/gemm_perf.py 16,64 bf16 2>&1 | grep "M="
done
} > gpurun_out/gemm_ablate5.txt 2>&1
cat gpurun_out/gemm_ablate5.txt
