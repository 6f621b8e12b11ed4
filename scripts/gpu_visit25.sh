#!/bin/bash
mkdir -p gpurun_out
for rep in 1 2; do
for mb in 1000 8; do
  DEB_COPY_MIN_MB=$mb timeout 300 python bench.py --no-cpu-baseline --no-latency > gpurun_out/e2e_$mb.json 2>/dev/null
  python -c "
import json
d=json.loads(open('gpurun_out/e2e_$mb.json').read().strip().splitlines()[-1]);print('copy_min_mb=$mb', 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'blocking', round(d['e2e']['blocking_predict_batch']['value'],1))"
done
done
timeout 300 python -m pytest tests/test_e2e_gpu.py -m gpu -x -q -p no:cacheprovider -k "stream" 2>&1 | tail -2
