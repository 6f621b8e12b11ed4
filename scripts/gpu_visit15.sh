#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/pytest_gpu_full.log 2>&1; echo "full gpu suite exit $?"; tail -3 gpurun_out/pytest_gpu_full.log
for lib in head new; do
  if [ $lib = head ]; then export DEPTH_B200_LIB=$PWD/scripts/_bin/libdepthb200_head.so; else unset DEPTH_B200_LIB; fi
  timeout 300 python bench.py --model MiDaS_small --width 1280 --height 720 --no-cpu-baseline --no-latency --no-e2e > gpurun_out/ab_small_$lib.json 2>/dev/null
  timeout 300 python bench.py --model MiDaS_small --width 640 --height 480 --batch 1 --no-cpu-baseline --no-e2e > gpurun_out/ab_small1_$lib.json 2>/dev/null
  timeout 300 python bench.py --no-cpu-baseline --no-e2e --no-latency > gpurun_out/ab_large_$lib.json 2>/dev/null
  timeout 300 python bench.py --width 1920 --height 1080 --no-cpu-baseline --no-e2e --no-latency > gpurun_out/ab_1080_$lib.json 2>/dev/null
  timeout 300 python bench.py --model DPT_Hybrid --no-cpu-baseline --no-e2e --no-latency > gpurun_out/ab_hybrid32_$lib.json 2>/dev/null
done
unset DEPTH_B200_LIB
timeout 600 python scripts/timeline.py MiDaS_small 32 1280 720 gpurun_out/tl3_small_b32.jsonl > gpurun_out/tl3_small_b32.txt 2>&1
timeout 600 python scripts/timeline.py DPT_Large 32 384 384 gpurun_out/tl3_large_b32.jsonl > gpurun_out/tl3_large_b32.txt 2>&1
