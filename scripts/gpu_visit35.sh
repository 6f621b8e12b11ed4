#!/bin/bash
# head_up3 with producer groups + packed math: parity, microbenchmark (2 groups / 1 group / first formulation), whole model
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -p no:cacheprovider -k "head" 2>&1 | tail -5
echo "== v3 g2"; timeout 120 python scripts/head_bench.py 2>&1 | grep head_up
echo "== v3 g1"; DEPTH_B200_LIB=$PWD/scripts/_bin/libdepthb200_g1.so timeout 120 python scripts/head_bench.py 2>&1 | grep head_up
echo "== v1"; DEB_HEAD_UP_V1=1 timeout 120 python scripts/head_bench.py 2>&1 | grep head_up
for v in v1 new v1 new; do
  unset DEB_HEAD_UP_V1
  if [ $v = v1 ]; then export DEB_HEAD_UP_V1=1; fi
  timeout 300 python bench.py --no-cpu-baseline --no-e2e --no-latency > gpurun_out/m_$v.json 2>/dev/null
  python -c "
import json
for n in ('m_$v',):
    d=json.loads(open('gpurun_out/%s.json'%n).read().strip().splitlines()[-1]);print(n, round(d['value'],1),'f/s', round(d['ms_per_step'],4), d['clocks']['sm_mhz'])"
done
unset DEB_HEAD_UP_V1
timeout 600 python -m pytest tests/test_e2e_gpu.py -m gpu -x -q -p no:cacheprovider -k "DPT_Large or bench_configuration" 2>&1 | tail -3
