#!/bin/bash
# attention_ws2 in the models: whole-model parity with DEB_ATTN_V2=1, then DPT_Large / 1080p / DPT_Hybrid A/B
mkdir -p gpurun_out
DEB_ATTN_V2=1 timeout 600 python -m pytest tests/test_e2e_gpu.py tests/test_oracle_hf_crosscheck.py -m gpu -x -q -p no:cacheprovider -k "DPT or bench_configuration or bf16x3 or hf" 2>&1 | tail -3
for v in 0 1 0 1; do
  export DEB_ATTN_V2=$v
  timeout 300 python bench.py --no-cpu-baseline --no-e2e --no-latency > gpurun_out/m_$v.json 2>/dev/null
  python -c "
import json
d=json.loads(open('gpurun_out/m_$v.json').read().strip().splitlines()[-1]);print('large v$v', round(d['value'],1),'f/s', round(d['ms_per_step'],4), d['clocks']['sm_mhz'], d.get('model_frac_of_peak'))"
done
for v in 0 1; do
  export DEB_ATTN_V2=$v
  timeout 300 python bench.py --width 1920 --height 1080 --no-cpu-baseline --no-e2e --no-latency > gpurun_out/m1080_$v.json 2>/dev/null
  timeout 300 python bench.py --model DPT_Hybrid --batch 1 --no-cpu-baseline --no-e2e --no-latency > gpurun_out/mh_$v.json 2>/dev/null
  python -c "
import json
for n in ('m1080_$v','mh_$v'):
    d=json.loads(open('gpurun_out/%s.json'%n).read().strip().splitlines()[-1]);print(n, round(d['value'],1),'f/s', round(d['ms_per_step'],4), d['clocks']['sm_mhz'])"
done
