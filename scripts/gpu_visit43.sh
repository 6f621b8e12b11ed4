#!/bin/bash
# stride-2 depthwise on 8x8 tiles, two outputs per thread (DEB_DW_S2_WIDE=1 = the 8x16 tile): parity + A/B
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -p no:cacheprovider -k "depthwise" 2>&1 | tail -2
for v in 1 0 1 0; do
  export DEB_DW_S2_WIDE=$v
  timeout 300 python bench.py --model MiDaS_small --width 1280 --height 720 --no-cpu-baseline --no-e2e --no-latency > gpurun_out/ms_$v.json 2>/dev/null
  timeout 300 python bench.py --model MiDaS_small --width 640 --height 480 --batch 1 --no-cpu-baseline --no-e2e --no-latency > gpurun_out/ms1_$v.json 2>/dev/null
  python -c "
import json
for n in ('ms_$v','ms1_$v'):
    d=json.loads(open('gpurun_out/%s.json'%n).read().strip().splitlines()[-1]);print(n, round(d['value'],1),'f/s', round(d['ms_per_step'],4), d['clocks']['sm_mhz'])"
done
unset DEB_DW_S2_WIDE
timeout 200 python scripts/timeline.py MiDaS_small 32 1280 720 gpurun_out/tl_small.jsonl 2>&1 | grep -E "step span|depthwise"
timeout 600 python -m pytest tests/test_e2e_gpu.py -m gpu -x -q -p no:cacheprovider -k "MiDaS_small" 2>&1 | tail -2
