#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -p no:cacheprovider > gpurun_out/pytest_k.log 2>&1; echo "kernel tests exit $?"; tail -2 gpurun_out/pytest_k.log
for v in 0 1; do
  echo "== DEB_EPI_DENSE=$v"
  DEB_EPI_DENSE=$v timeout 600 python scripts/gemm_small_bench.py convonly 2>&1 | grep -v Loading
  DEB_EPI_DENSE=$v timeout 300 python bench.py --no-cpu-baseline --no-e2e --no-latency > gpurun_out/v19_$v.json 2>/dev/null
  python -c "
import json;d=json.loads(open('gpurun_out/v19_$v.json').read().strip().splitlines()[-1]);print('dense=$v', round(d['value'],1),'f/s', round(d['ms_per_step'],3), d['clocks'])"
done
