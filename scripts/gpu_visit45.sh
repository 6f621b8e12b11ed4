#!/bin/bash
# programmatic-dependent-launch class mask (1 GEMM, 2 attention / head, 4 LayerNorm, 8 depthwise) on the batch-1 paths
mkdir -p gpurun_out
for v in 11 0 1 10 15; do
  export DEB_PDL=$v
  timeout 300 python bench.py --model DPT_Hybrid --batch 1 --no-cpu-baseline --no-e2e --no-latency > gpurun_out/mh_$v.json 2>/dev/null
  timeout 300 python bench.py --model MiDaS_small --width 640 --height 480 --batch 1 --no-cpu-baseline --no-e2e --no-latency > gpurun_out/ms1_$v.json 2>/dev/null
  python -c "
import json
for n in ('mh_$v','ms1_$v'):
    d=json.loads(open('gpurun_out/%s.json'%n).read().strip().splitlines()[-1]);print(n, round(d['value'],1),'f/s', round(d['ms_per_step'],4), d['clocks']['sm_mhz'])"
done
