#!/bin/bash
export DEPTH_B200_LIB=$PWD/scripts/_bin/libdepthb200_trace.so
timeout 120 python scripts/gemm_trace.py 262144 128 64 bf16 2>&1 | grep -v Loading
timeout 120 python scripts/gemm_trace.py 262144 144 24 bf16 2>&1 | grep -v Loading
timeout 120 python scripts/gemm_trace.py 4096 816 136 bf16 2>&1 | grep -v Loading
