#!/bin/bash
# depthwise kernel triggers its dependents early (griddepcontrol.launch_dependents): A/B via DEB_DW_PDL_TRIGGER
mkdir -p gpurun_out
for v in 0 1 0 1; do
  export DEB_DW_PDL_TRIGGER=$v
  timeout 300 python bench.py --model MiDaS_small --width 1280 --height 720 --no-cpu-baseline --no-e2e --no-latency > gpurun_out/ms_$v.json 2>/dev/null
  timeout 300 python bench.py --model MiDaS_small --width 640 --height 480 --batch 1 --no-cpu-baseline --no-e2e --no-latency > gpurun_out/ms1_$v.json 2>/dev/null
  python -c "
import json
for n in ('ms_$v','ms1_$v'):
    d=json.loads(open('gpurun_out/%s.json'%n).read().strip().splitlines()[-1]);print(n, round(d['value'],1),'f/s', round(d['ms_per_step'],4), d['clocks']['sm_mhz'])"
done
unset DEB_DW_PDL_TRIGGER
timeout 600 python -m pytest tests/test_e2e_gpu.py -m gpu -x -q -p no:cacheprovider -k "MiDaS_small or batch_is_independent" 2>&1 | tail -2
