#!/bin/bash
# final validation of the tree as committed: GPU suite (the zz no-alloc file runs last), smoke, default bench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/final_pytest_gpu.log 2>&1; echo "gpu suite exit $?"; tail -3 gpurun_out/final_pytest_gpu.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider 2>&1 | tail -2
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/final_bench_n1.json 2> gpurun_out/final_bench.err; echo "bench exit $?"; cut -c1-330 gpurun_out/final_bench_n1.json
