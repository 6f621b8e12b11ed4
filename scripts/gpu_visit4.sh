#!/bin/bash
# Visit 4: full GPU suite on the engine-only tree + bench lines.
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -p no:cacheprovider -s > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log
grep -E "rel=|passed|failed|Error|FAILED|launches" gpurun_out/pytest_gpu.log | tail -40
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -3 gpurun_out/smoke.log
BENCH_DETAIL=1 timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
cat gpurun_out/bench.json; grep -v Loading gpurun_out/bench.err | tail -25
BENCH_DETAIL=1 timeout 600 python bench.py --model MiDaS_small --width 1280 --height 720 --no-cpu-baseline --no-latency > gpurun_out/bench_small.json 2> gpurun_out/bench_small.err; echo "bench small exit $?"
cat gpurun_out/bench_small.json
