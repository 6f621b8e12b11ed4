#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -q -p no:cacheprovider -x > gpurun_out/pytest_k.log 2>&1; echo "kernel exit $?"; tail -5 gpurun_out/pytest_k.log
timeout 900 python -m pytest tests/test_engine_gpu.py tests/test_e2e_gpu.py -m gpu -q -p no:cacheprovider -s -x > gpurun_out/pytest_e2e.log 2>&1; echo "e2e+engine exit $?"
grep -E "rel=|passed|failed|Error|FAILED" gpurun_out/pytest_e2e.log | tail -16
for gn in 1 0; do
DEB_GN_FUSED=$gn BENCH_DETAIL=1 timeout 600 python bench.py --model DPT_Hybrid --batch 1 --no-cpu-baseline 2> gpurun_out/bench_hybrid_b1_gn$gn.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('hybrid b1 GN_FUSED=$gn', 'ms/step', round(d['ms_per_step'],3), 'latency', d['latency'], d['roofline']['kernel_ms_per_step'])"
done
BENCH_DETAIL=1 timeout 600 python bench.py --model DPT_Hybrid --batch 1 --no-cpu-baseline --no-latency > gpurun_out/bench_hybrid_b1.json 2> gpurun_out/bench_hybrid_b1.err; grep -v Loading gpurun_out/bench_hybrid_b1.err | head -40
timeout 600 python bench.py --batch 4 --no-cpu-baseline --no-latency 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('DPT_Large B=4', round(d['value'],1), 'ms/step', round(d['ms_per_step'],3))"
timeout 600 python bench.py --batch 1 --no-cpu-baseline --no-latency 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('DPT_Large B=1', round(d['value'],1), 'ms/step', round(d['ms_per_step'],3))"
timeout 600 python bench.py --model MiDaS_small --width 1280 --height 720 --no-cpu-baseline --no-latency --no-e2e 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('MiDaS_small 720p B=32', round(d['value'],1), 'ms/step', round(d['ms_per_step'],3), d['roofline']['kernel_ms_per_step'])"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches_hybrid_b1.csv \
  python bench.py --model DPT_Hybrid --batch 1 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --no-latency > gpurun_out/ncu_hybrid.log 2>&1; echo "ncu hybrid exit $?"
