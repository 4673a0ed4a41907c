#!/bin/bash
for w in 4 3 2; do echo "== B200_ATTN_WARPS=$w"; B200_ATTN_WARPS=$w timeout 300 python tools/attn_perf.py 64x256x32x8,32x256x32x8,64x512x32x8,64x1024x32x8,64x2048x32x8 2>&1 | tail -5; done
echo "== auto bf16"; timeout 300 python tools/attn_perf.py 2>&1 | tail -6
echo "== auto fp8"; timeout 300 python tools/attn_perf.py 64x256x32x8,64x2048x32x8,8x8192x32x8,64x8192x8x1,1x32768x32x8,64x32768x8x1 fp8 2>&1 | tail -6
