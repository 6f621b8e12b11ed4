#!/bin/bash
mkdir -p gpurun_out
for v in 148 96 74 48 32; do
  DEB_GN_BLOCKS=$v timeout 300 python bench.py --model DPT_Hybrid --batch 1 --no-cpu-baseline --no-e2e > gpurun_out/h$v.json 2>/dev/null
  python -c "
import json
d=json.loads(open('gpurun_out/h$v.json').read().strip().splitlines()[-1]);print('gn_blocks=$v', round(d['ms_per_step'],4),'ms/step device', d['latency']['p50_ms'], d['roofline']['kernel_ms_per_step'].get('groupnorm'))"
done
