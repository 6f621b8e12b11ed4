#!/bin/bash
# narrow-N tile heuristic (N <= 32 -> 64-wide tile, TMA-store epilogue): parity + A/B against the library of the previous commit
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -p no:cacheprovider -k "gemm" 2>&1 | tail -2
timeout 600 python -m pytest tests/test_e2e_gpu.py -m gpu -x -q -p no:cacheprovider -k "MiDaS_small or batch_is_independent" 2>&1 | tail -2
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
