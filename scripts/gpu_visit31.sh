#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -p no:cacheprovider 2>&1 | tail -2
for v in prev new; do
  unset DEPTH_B200_LIB
  if [ $v = prev ]; then export DEPTH_B200_LIB=$PWD/scripts/_bin/libdepthb200_prev.so; fi
  echo "== $v"
  timeout 600 python scripts/gemm_small_bench.py x 2>&1 | grep -v Loading | grep -E "rcu|layer_rn|same|project|expand"
  timeout 300 python bench.py --no-cpu-baseline --no-e2e --no-latency > gpurun_out/m_$v.json 2>/dev/null
  timeout 300 python bench.py --model MiDaS_small --width 1280 --height 720 --no-cpu-baseline --no-e2e --no-latency > gpurun_out/ms_$v.json 2>/dev/null
  timeout 300 python bench.py --model DPT_Hybrid --batch 1 --no-cpu-baseline --no-e2e > gpurun_out/mh_$v.json 2>/dev/null
  python -c "
import json
for n in ('m_$v','ms_$v','mh_$v'):
    d=json.loads(open('gpurun_out/%s.json'%n).read().strip().splitlines()[-1]);print(n, round(d['value'],1),'f/s', round(d['ms_per_step'],4), d['clocks']['sm_mhz'], (d.get('latency') or {}).get('p50_ms'))"
done
