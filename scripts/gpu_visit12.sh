#!/bin/bash
mkdir -p gpurun_out
DEPTH_B200_LIB=$PWD/scripts/_bin/libdepthb200_head.so timeout 600 python scripts/gemm_small_bench.py "HEAD library" 2>&1 | grep -v Loading | tee gpurun_out/gsb_head.md
timeout 600 python scripts/gemm_small_bench.py "working tree" 2>&1 | grep -v Loading | tee gpurun_out/gsb_new.md
DEPTH_B200_LIB=$PWD/scripts/_bin/libdepthb200_sb2.so timeout 600 python scripts/gemm_small_bench.py "working tree, 2 store buffers per group" 2>&1 | grep -v Loading | tee gpurun_out/gsb_sb2.md
