#!/bin/bash
# split-K for half-filled grids with long K (DEB_SPLITK_HALF=0 restores the old rule): DPT_Hybrid batch 1 A/B + parity
mkdir -p gpurun_out
for v in 0 1 0 1; do
  export DEB_SPLITK_HALF=$v
  timeout 300 python bench.py --model DPT_Hybrid --batch 1 --no-cpu-baseline --no-e2e > gpurun_out/mh_$v.json 2>/dev/null
  timeout 300 python bench.py --model MiDaS_small --width 1280 --height 720 --no-cpu-baseline --no-e2e --no-latency > gpurun_out/ms_$v.json 2>/dev/null
  python -c "
import json
for n in ('mh_$v','ms_$v'):
    d=json.loads(open('gpurun_out/%s.json'%n).read().strip().splitlines()[-1]);print(n, round(d['value'],1),'f/s', round(d['ms_per_step'],4), d['clocks']['sm_mhz'], (d.get('latency') or {}).get('p50_ms'))"
done
unset DEB_SPLITK_HALF
timeout 600 python -m pytest tests/test_e2e_gpu.py -m gpu -x -q -p no:cacheprovider -k "DPT_Hybrid or hybrid" 2>&1 | tail -2
timeout 200 python scripts/timeline.py DPT_Hybrid 1 384 384 gpurun_out/tl_hybrid1.jsonl 2>&1 | grep -v Loading | sed -n 3,16p
