#!/bin/bash
# checkpoint of the tree with frame-packed depthwise tiles + head_up3: GPU suite, smoke, the three model benches, ncu of head_up3
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/ckpt_pytest_gpu.log 2>&1; echo "gpu suite exit $?"; tail -3 gpurun_out/ckpt_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/ckpt_bench_large.json 2>/dev/null; echo "bench exit $?"
timeout 300 python bench.py --model MiDaS_small --width 1280 --height 720 --no-cpu-baseline > gpurun_out/ckpt_bench_small720.json 2>/dev/null
timeout 300 python bench.py --model DPT_Hybrid --batch 1 --no-cpu-baseline > gpurun_out/ckpt_bench_hybrid1.json 2>/dev/null
timeout 300 python bench.py --width 1920 --height 1080 --no-cpu-baseline --no-latency > gpurun_out/ckpt_bench_1080p.json 2>/dev/null
python -c "
import json
for n in ('large','small720','hybrid1','1080p'):
    d=json.loads(open('gpurun_out/ckpt_bench_%s.json'%n).read().strip().splitlines()[-1]);print(n, round(d['value'],1),'f/s', round(d['ms_per_step'],4), 'e2e', round((d.get('e2e') or {}).get('value',0),1), d['clocks']['sm_mhz'], (d.get('latency') or {}).get('p50_ms'), d.get('roofline',{}).get('frac'), d.get('model_frac_of_peak'))"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:head_up3 --launch-skip 3 -c 1 -f -o gpurun_out/r02_head_up3_v2 python scripts/head_bench.py > gpurun_out/ncu_head.log 2>&1; echo "ncu exit $?"
