#!/bin/bash
# visit 11: general-epilogue residual prefetch + MiDaS_small lanes: parity first, then timings
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/pytest_gpu_full.log 2>&1; echo "full gpu suite exit $?"; tail -4 gpurun_out/pytest_gpu_full.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_large.json 2> gpurun_out/bench_large.err; echo "bench exit $?"; cut -c1-400 gpurun_out/bench_large.json
for lanes in 1 2 4; do
  DEB_SMALL_LANES=$lanes timeout 300 python bench.py --model MiDaS_small --width 1280 --height 720 --no-cpu-baseline --no-latency > gpurun_out/bench_small_l$lanes.json 2>/dev/null; echo "small lanes=$lanes exit $?"
  python -c "
import json;d=json.loads(open('gpurun_out/bench_small_l$lanes.json').read().strip().splitlines()[-1]);print('lanes=$lanes', d['value'], d['ms_per_step'], d['e2e']['value'])"
done
for lanes in 1 4; do
  DEB_SMALL_LANES=$lanes timeout 300 python bench.py --workload stream720p --batch 16 --no-cpu-baseline > gpurun_out/bench_stream_l$lanes.json 2>/dev/null; echo "stream lanes=$lanes exit $?"; cut -c1-300 gpurun_out/bench_stream_l$lanes.json
done
timeout 600 python scripts/timeline.py DPT_Large 32 384 384 gpurun_out/tl2_large_b32.jsonl > gpurun_out/tl2_large_b32.txt 2>&1; head -24 gpurun_out/tl2_large_b32.txt | tail -18
timeout 600 python scripts/timeline.py MiDaS_small 32 1280 720 gpurun_out/tl2_small_b32.jsonl > gpurun_out/tl2_small_b32.txt 2>&1; head -8 gpurun_out/tl2_small_b32.txt | tail -3
