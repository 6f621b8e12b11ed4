#!/bin/bash
mkdir -p gpurun_out
timeout 600 python scripts/timeline.py DPT_Large 32 384 384 gpurun_out/tl_large_b32.jsonl > gpurun_out/tl_large_b32.txt 2>&1; echo "exit $?"; head -60 gpurun_out/tl_large_b32.txt
timeout 600 python scripts/timeline.py DPT_Hybrid 1 384 384 gpurun_out/tl_hybrid_b1.jsonl > gpurun_out/tl_hybrid_b1.txt 2>&1; echo "exit $?"; head -50 gpurun_out/tl_hybrid_b1.txt
timeout 600 python scripts/timeline.py MiDaS_small 32 1280 720 gpurun_out/tl_small_b32.jsonl > gpurun_out/tl_small_b32.txt 2>&1; echo "exit $?"; head -50 gpurun_out/tl_small_b32.txt
