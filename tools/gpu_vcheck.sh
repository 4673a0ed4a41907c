#!/bin/bash
python tools/variant_check.py
B200_LIB_OPS=build/variants/ablate/libb200ops.so B200_GEMM_EXP=0 python tools/variant_check.py
B200_LIB_OPS=build/variants/ablate/libb200ops.so B200_GEMM_EXP=8 python tools/variant_check.py
B200_LIB_OPS=build/variants/ablate/libb200ops.so B200_GEMM_EXP=6 python tools/variant_check.py
B200_LIB_OPS=build/variants/ablate/libb200ops.so B200_GEMM_EXP=2 python tools/variant_check.py
