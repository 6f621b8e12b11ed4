#!/bin/bash
# Visit 1 of round 2: full GPU test suite (new parity gates print rel/pix/dyn), default bench line, MiDaS_small + hybrid lines.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider -s > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log
grep -E "rel=|passed|failed|Error|FAILED" gpurun_out/pytest_gpu.log | tail -40
BENCH_DETAIL=1 timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
cat gpurun_out/bench.json
BENCH_DETAIL=1 timeout 600 python bench.py --model MiDaS_small --width 1280 --height 720 --no-cpu-baseline > gpurun_out/bench_small.json 2> gpurun_out/bench_small.err; echo "bench small exit $?"
cat gpurun_out/bench_small.json
BENCH_DETAIL=1 timeout 600 python bench.py --model DPT_Hybrid --batch 1 --no-cpu-baseline > gpurun_out/bench_hybrid_b1.json 2> gpurun_out/bench_hybrid_b1.err; echo "bench hybrid exit $?"
cat gpurun_out/bench_hybrid_b1.json
