#!/bin/bash
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_kernel --launch-skip 4 -c 1 -f -o gpurun_out/r02_gemm_shortk \
  python scripts/probe_gemm.py 262144 64 128 > gpurun_out/ncu_shortk.log 2>&1; echo "ncu exit $?"; tail -2 gpurun_out/ncu_shortk.log
