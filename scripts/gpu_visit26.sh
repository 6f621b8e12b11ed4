#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -p no:cacheprovider > gpurun_out/pytest_k.log 2>&1; echo "kernel tests exit $?"; tail -2 gpurun_out/pytest_k.log
for v in 0 1; do
  echo "== DEB_WARP_STORE=$v"
  DEB_WARP_STORE=$v timeout 300 python scripts/proj_probe.py 2>&1 | grep -v Loading
  DEB_WARP_STORE=$v timeout 300 python bench.py --no-cpu-baseline --no-e2e --no-latency > gpurun_out/ws$v.json 2>/dev/null
  DEB_WARP_STORE=$v timeout 300 python bench.py --model MiDaS_small --width 1280 --height 720 --no-cpu-baseline --no-e2e --no-latency > gpurun_out/ws${v}s.json 2>/dev/null
  python -c "
import json
for n in ('ws$v','ws${v}s'):
    d=json.loads(open('gpurun_out/%s.json'%n).read().strip().splitlines()[-1]);print(n, round(d['value'],1),'f/s', round(d['ms_per_step'],4), d['clocks']['sm_mhz'])"
done
