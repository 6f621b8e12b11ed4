#!/bin/bash
# Visit 3: C++ engine bring-up.
mkdir -p gpurun_out
timeout 600 python scripts/engine_vs_pynets.py > gpurun_out/engine_vs_py.log 2>&1; echo "engine vs py exit $?"; tail -12 gpurun_out/engine_vs_py.log
timeout 900 python -m pytest tests/test_engine_gpu.py -m gpu -q -p no:cacheprovider -s -x > gpurun_out/pytest_engine.log 2>&1; echo "engine tests exit $?"; tail -15 gpurun_out/pytest_engine.log
timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -q -p no:cacheprovider -x -k "split3 or upsample" > gpurun_out/pytest_k.log 2>&1; echo "kernel tests exit $?"; tail -5 gpurun_out/pytest_k.log
timeout 1200 python -m pytest tests/test_e2e_gpu.py tests/test_oracle_hf_crosscheck.py -m gpu -q -p no:cacheprovider -s > gpurun_out/pytest_e2e.log 2>&1; echo "e2e exit $?"
grep -E "rel=|passed|failed|Error|FAILED" gpurun_out/pytest_e2e.log | tail -30
BENCH_DETAIL=1 timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
cat gpurun_out/bench.json; tail -3 gpurun_out/bench.err
