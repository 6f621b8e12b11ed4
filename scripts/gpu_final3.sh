#!/bin/bash
# last validation of the round-2 tree: GPU suite, the no-alloc file alone, smoke, default bench, DPT_Hybrid batch-1 line
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/final_pytest_gpu.log 2>&1; echo "gpu suite exit $?"; tail -3 gpurun_out/final_pytest_gpu.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider 2>&1 | tail -2
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/final_bench_n1.json 2> gpurun_out/final_bench.err; echo "bench exit $?"
timeout 300 python bench.py --model DPT_Hybrid --batch 1 --no-cpu-baseline > gpurun_out/final_bench_hybrid_b1.json 2>/dev/null
timeout 300 python bench.py --model MiDaS_small --width 1280 --height 720 --no-cpu-baseline --no-latency > gpurun_out/final_bench_small_720p_b32.json 2>/dev/null
timeout 300 python bench.py --model MiDaS_small --width 640 --height 480 --batch 1 --no-cpu-baseline --no-latency > gpurun_out/final_bench_small_480p_b1.json 2>/dev/null
python - <<PY
import json
for n in ('n1','hybrid_b1','small_720p_b32','small_480p_b1'):
    d=json.loads(open('gpurun_out/final_bench_%s.json'%n).read().strip().splitlines()[-1])
    print(n, round(d['value'],1), d['unit'], round(d['ms_per_step'],4), 'e2e', round((d.get('e2e') or {}).get('value',0),1), d['clocks']['sm_mhz'], d['clocks'].get('reasons'), (d.get('latency') or {}).get('p50_ms'), 'frac', d.get('roofline',{}).get('frac'), d.get('model_frac_of_peak'))
PY
