#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_engine_gpu.py tests/test_e2e_gpu.py -m gpu -q -p no:cacheprovider -s -x > gpurun_out/pytest_e2e.log 2>&1; echo "e2e+engine exit $?"
grep -E "rel=|passed|failed|Error|FAILED" gpurun_out/pytest_e2e.log | tail -30
timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -q -p no:cacheprovider -x > gpurun_out/pytest_k.log 2>&1; echo "kernel exit $?"; tail -3 gpurun_out/pytest_k.log
BENCH_DETAIL=1 timeout 600 python bench.py --model MiDaS_small --width 1280 --height 720 --no-cpu-baseline --no-latency > gpurun_out/bench_small.json 2> gpurun_out/bench_small.err; echo "bench small exit $?"
cat gpurun_out/bench_small.json; grep -v Loading gpurun_out/bench_small.err | tail -30
timeout 600 python bench.py --workload stream720p --no-cpu-baseline --no-latency > gpurun_out/bench_stream.json 2> gpurun_out/bench_stream.err; echo "bench stream exit $?"
cat gpurun_out/bench_stream.json; grep -v Loading gpurun_out/bench_stream.err | tail -5
for b in 8 16 64; do timeout 600 python bench.py --workload stream720p --batch $b --no-cpu-baseline --no-latency 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('stream batch', d['stream']['frames_per_submission'], 'fused fps', round(d['stream']['fused_fps']), 'serial', d['stream']['serial_loop'], 'equal', d['stream']['ri_depth_abs_depth_equal_to_serial_loop'], 'device value', round(d['value']))"; done
