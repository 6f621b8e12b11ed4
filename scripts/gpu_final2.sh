#!/bin/bash
# final validation of the round-2 tree: GPU suite (the zz no-alloc file runs last, then alone), smoke, the bench lines of
# every single-GPU BASELINE configuration, the reference arm, and fresh ncu launch lists of DPT_Large / MiDaS_small
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/final_pytest_gpu.log 2>&1; echo "gpu suite exit $?"; tail -3 gpurun_out/final_pytest_gpu.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider 2>&1 | tail -2
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/final_bench_n1.json 2> gpurun_out/final_bench.err; echo "bench exit $?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/final_bench_ref_n1.json 2>/dev/null; echo "ref exit $?"
timeout 300 python bench.py --model MiDaS_small --width 1280 --height 720 --no-cpu-baseline --no-latency > gpurun_out/final_bench_small_720p_b32.json 2>/dev/null
timeout 300 python bench.py --model MiDaS_small --width 640 --height 480 --batch 1 --no-cpu-baseline --no-latency > gpurun_out/final_bench_small_480p_b1.json 2>/dev/null
timeout 300 python bench.py --workload stream720p --batch 16 --no-cpu-baseline --no-latency > gpurun_out/final_bench_stream720p_b16.json 2>/dev/null
timeout 300 python bench.py --model DPT_Hybrid --batch 1 --no-cpu-baseline > gpurun_out/final_bench_hybrid_b1.json 2>/dev/null
timeout 300 python bench.py --model DPT_Hybrid --no-cpu-baseline --no-latency > gpurun_out/final_bench_hybrid_b32.json 2>/dev/null
timeout 300 python bench.py --width 1920 --height 1080 --no-cpu-baseline --no-latency > gpurun_out/final_bench_1080p_b32.json 2>/dev/null
python - <<PY
import json
for n in ('n1','small_720p_b32','small_480p_b1','stream720p_b16','hybrid_b1','hybrid_b32','1080p_b32'):
    try:
        d=json.loads(open('gpurun_out/final_bench_%s.json'%n).read().strip().splitlines()[-1])
        print(n, round(d['value'],1), d['unit'], round(d['ms_per_step'],4), 'e2e', round((d.get('e2e') or {}).get('value',0),1), d['clocks']['sm_mhz'], d['clocks'].get('reasons'), (d.get('latency') or {}).get('p50_ms'), 'frac', d.get('roofline',{}).get('frac'), d.get('model_frac_of_peak'), 'launches', d.get('gpu_launches'))
    except Exception as e: print(n, 'ERR', e)
print(open('gpurun_out/final_bench_ref_n1.json').read()[:400])
PY
M="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum"
B="--steps 1 --warmup 3 --no-e2e --no-cpu-baseline --no-latency"
timeout 900 ncu --metrics $M --clock-control none -c 6000 --csv --log-file gpurun_out/r02f_launches_dpt_large_b32.csv python bench.py $B > gpurun_out/ncu_large.log 2>&1; echo "ncu large exit $?"
timeout 900 ncu --metrics $M --clock-control none -c 6000 --csv --log-file gpurun_out/r02f_launches_small_720p_b32.csv python bench.py --model MiDaS_small --width 1280 --height 720 $B > gpurun_out/ncu_small.log 2>&1; echo "ncu small exit $?"
