#!/bin/bash
# Short GPU-box visit: GPU tests, one bench line (+ optional micro-benchmarks with MICRO=1).
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q -p no:cacheprovider -s > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log
grep -E "rel=|passed|failed|Error" gpurun_out/pytest_gpu.log | tail -12
if [ -n "$MICRO" ]; then
  timeout 300 python scripts/gemm_bench.py > gpurun_out/gemm_bench.log 2>&1; head -12 gpurun_out/gemm_bench.log
  timeout 300 python scripts/attn_bench.py > gpurun_out/attn_bench.log 2>&1; cat gpurun_out/attn_bench.log
fi
BENCH_DETAIL=1 timeout 900 python bench.py $BENCH_ARGS > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
cat gpurun_out/bench.json
