#!/bin/bash
mkdir -p gpurun_out
python tools/prefill_perf.py 2048 w4a16 > gpurun_out/plain_p.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:prefill_paged_attn_tiled -s 2 -c 1 -f -o gpurun_out/prefill_attn python tools/prefill_perf.py 2048 w4a16 > gpurun_out/ncu_p.log 2>&1
echo "rc=$?"; tail -2 gpurun_out/ncu_p.log; ls -la gpurun_out/prefill_attn.ncu-rep
