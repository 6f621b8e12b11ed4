#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -p no:cacheprovider -k "gemm or gelu" 2>&1 | tail -2
for v in prev new prev new; do
  unset DEPTH_B200_LIB
  if [ $v = prev ]; then export DEPTH_B200_LIB=$PWD/scripts/_bin/libdepthb200_prev.so; fi
  echo "== $v"; timeout 300 python scripts/fc1_probe.py 2>&1 | grep -v Loading
done
