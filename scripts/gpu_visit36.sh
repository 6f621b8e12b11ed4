#!/bin/bash
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:head_up3 --launch-skip 3 -c 1 -f -o gpurun_out/r02_head_up3 python scripts/head_bench.py > gpurun_out/ncu_head.log 2>&1; echo "ncu exit $?"
tail -3 gpurun_out/ncu_head.log
