#!/bin/bash
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k 'regex:gemm_kernel<\(int\)256, \(int\)2, \(int\)1>' --launch-skip 4 -c 1 -f -o gpurun_out/r02_conv_epi1_v2 \
  python scripts/gemm_small_bench.py convonly > gpurun_out/ncu_conv.log 2>&1; echo "ncu exit $?"; tail -2 gpurun_out/ncu_conv.log
