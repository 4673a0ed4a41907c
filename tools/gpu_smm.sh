#!/bin/bash
for mu in 1 4 8 16 32; do echo "== B200_SMM_MIN_UNITS=$mu"; B200_SMM_MIN_UNITS=$mu timeout 300 python tools/smm_perf.py 16,64 fp8 qkv_tp8,o_tp8,gate_up_tp8,down_tp8,qkv_tp1 2>&1 | tail -10; done
