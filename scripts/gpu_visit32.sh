#!/bin/bash
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -p no:cacheprovider -k "gemm or conv3x3 or splitk" 2>&1 | tail -3
DEPTH_B200_LIB=$PWD/scripts/_bin/libdepthb200_trace.so timeout 100 python scripts/gemm_trace.py 262144 128 64 bf16 2>&1 | grep -v Loading
DEPTH_B200_LIB=$PWD/scripts/_bin/libdepthb200_trace.so timeout 100 python scripts/gemm_trace.py 18464 1024 1024 f32 acc 2>&1 | grep -v Loading
timeout 120 python scripts/proj_probe.py 2>&1 | grep -v Loading
timeout 120 python scripts/fc1_probe.py 2>&1 | grep -v Loading
