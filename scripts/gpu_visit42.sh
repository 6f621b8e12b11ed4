#!/bin/bash
# smem-tiled stem kernel: parity, A/B against the previous library
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -p no:cacheprovider -k "stem" 2>&1 | tail -2
timeout 600 python -m pytest tests/test_e2e_gpu.py tests/test_engine_gpu.py -m gpu -x -q -p no:cacheprovider -k "MiDaS_small or batch_is_independent or engine" 2>&1 | tail -2
for v in prev new prev new; do
  unset DEPTH_B200_LIB
  if [ $v = prev ]; then export DEPTH_B200_LIB=$PWD/scripts/_bin/libdepthb200_prev.so; fi
  timeout 300 python bench.py --model MiDaS_small --width 1280 --height 720 --no-cpu-baseline --no-e2e --no-latency > gpurun_out/ms_$v.json 2>/dev/null
  timeout 300 python bench.py --model MiDaS_small --width 640 --height 480 --batch 1 --no-cpu-baseline --no-e2e --no-latency > gpurun_out/ms1_$v.json 2>/dev/null
  python -c "
import json
for n in ('ms_$v','ms1_$v'):
    d=json.loads(open('gpurun_out/%s.json'%n).read().strip().splitlines()[-1]);print(n, round(d['value'],1),'f/s', round(d['ms_per_step'],4), d['clocks']['sm_mhz'])"
done
unset DEPTH_B200_LIB
timeout 200 python scripts/timeline.py MiDaS_small 32 1280 720 gpurun_out/tl_small.jsonl 2>&1 | grep -E "step span|stem|preprocess"
