#!/bin/bash
# fused GroupNorm: all partials requested before the first add - parity + DPT_Hybrid batch-1 A/B
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_kernels_gpu.py tests/test_zz_no_alloc.py -m gpu -x -q -p no:cacheprovider -k "groupnorm or no_alloc or static" 2>&1 | tail -2
for v in prev new prev new; do
  unset DEPTH_B200_LIB
  if [ $v = prev ]; then export DEPTH_B200_LIB=$PWD/scripts/_bin/libdepthb200_prev.so; fi
  timeout 300 python bench.py --model DPT_Hybrid --batch 1 --no-cpu-baseline --no-e2e > gpurun_out/mh_$v.json 2>/dev/null
  python -c "
import json
for n in ('mh_$v',):
    d=json.loads(open('gpurun_out/%s.json'%n).read().strip().splitlines()[-1]);print(n, round(d['value'],1),'f/s', round(d['ms_per_step'],4), d['clocks']['sm_mhz'], (d.get('latency') or {}).get('p50_ms'))"
done
unset DEPTH_B200_LIB
timeout 600 python -m pytest tests/test_e2e_gpu.py -m gpu -x -q -p no:cacheprovider -k "DPT_Hybrid or hybrid" 2>&1 | tail -2
