#!/bin/bash
mkdir -p gpurun_out
timeout 200 python scripts/timeline.py MiDaS_small 32 1280 720 gpurun_out/tl_small.jsonl 2>&1 | grep -v Loading | head -40
timeout 200 python scripts/timeline.py DPT_Hybrid 1 384 384 gpurun_out/tl_hybrid1.jsonl 2>&1 | grep -v Loading | head -34
timeout 200 python scripts/timeline.py DPT_Large 32 384 384 gpurun_out/tl_large.jsonl 2>&1 | grep -v Loading | head -24
