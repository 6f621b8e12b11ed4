#!/bin/bash
mkdir -p gpurun_out
timeout 240 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -p no:cacheprovider -k "dwproject" > gpurun_out/pytest_dwp.log 2>&1; echo "dwproject tests exit $?"; tail -2 gpurun_out/pytest_dwp.log | cut -c1-200
for cfg in "0 148" "1 148" "1 296" "1 444"; do
  set -- $cfg
  DEB_MBCONV_FUSED=$1 DEB_MB_CTAS=$2 timeout 300 python bench.py --model MiDaS_small --width 1280 --height 720 --no-cpu-baseline --no-latency --no-e2e > gpurun_out/fused.json 2>/dev/null
  python -c "
import json
d=json.loads(open('gpurun_out/fused.json').read().strip().splitlines()[-1]);print('fused=$1 ctas=$2', round(d['value'],1),'f/s', round(d['ms_per_step'],4), d['roofline'].get('kernel_ms_per_step'))"
done
