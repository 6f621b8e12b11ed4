#!/bin/bash
# depthwise frame-packed tiles: parity, then A/B against the HEAD library on the same GPU
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -p no:cacheprovider -k "depthwise or dwproject" 2>&1 | tail -3
timeout 600 python -m pytest tests/test_e2e_gpu.py -m gpu -x -q -p no:cacheprovider -k "MiDaS_small or batch_is_independent or stream_loop" 2>&1 | tail -3
for v in prev new prev new; do
  unset DEPTH_B200_LIB
  if [ $v = prev ]; then export DEPTH_B200_LIB=$PWD/scripts/_bin/libdepthb200_prev.so; fi
  echo "== $v"
  timeout 300 python bench.py --model MiDaS_small --width 1280 --height 720 --no-cpu-baseline --no-e2e --no-latency > gpurun_out/ms_$v.json 2>/dev/null
  timeout 300 python bench.py --model MiDaS_small --width 640 --height 480 --batch 32 --no-cpu-baseline --no-e2e --no-latency > gpurun_out/ms480_$v.json 2>/dev/null
  python -c "
import json
for n in ('ms_$v','ms480_$v'):
    d=json.loads(open('gpurun_out/%s.json'%n).read().strip().splitlines()[-1]);print(n, round(d['value'],1),'f/s', round(d['ms_per_step'],4), d['clocks']['sm_mhz'], {k:round(v,4) for k,v in (d.get('kernel_ms_per_step') or {}).items()})"
done
