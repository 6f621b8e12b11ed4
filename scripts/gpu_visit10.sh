#!/bin/bash
# round-2 ncu evidence: launch lists (time + DRAM bytes per launch) of the three models, full capture of the depthwise kernels
mkdir -p gpurun_out
M="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum"
B="--steps 1 --warmup 3 --no-e2e --no-cpu-baseline --no-latency"
timeout 300 python bench.py --model MiDaS_small --width 1280 --height 720 $B > gpurun_out/plain_small.log 2>&1; echo "plain small exit $?"
timeout 900 ncu --metrics $M --clock-control none -c 6000 --csv --log-file gpurun_out/r02_launches_small_720p_b32.csv \
  python bench.py --model MiDaS_small --width 1280 --height 720 $B > gpurun_out/ncu_small.log 2>&1; echo "ncu small exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:depthwise --launch-skip 72 -c 48 -f -o gpurun_out/r02_dw_full \
  python bench.py --model MiDaS_small --width 1280 --height 720 $B > gpurun_out/ncu_dw_full.log 2>&1; echo "ncu dw full exit $?"
timeout 900 ncu --metrics $M --clock-control none -c 6000 --csv --log-file gpurun_out/r02_launches_dpt_large_b32.csv \
  python bench.py $B > gpurun_out/ncu_large.log 2>&1; echo "ncu large exit $?"
timeout 900 ncu --metrics $M --clock-control none -c 6000 --csv --log-file gpurun_out/r02_launches_dpt_hybrid_b1.csv \
  python bench.py --model DPT_Hybrid --batch 1 $B > gpurun_out/ncu_hybrid.log 2>&1; echo "ncu hybrid exit $?"
ls -la gpurun_out | head -30
