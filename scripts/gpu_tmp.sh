mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_kernels_gpu.py -x -q -p no:cacheprovider -k "upsample or head" 2>&1 | tail -3
timeout 300 python scripts/head_bench.py 2>&1 | tail -4
BENCH_DETAIL=1 timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
cat gpurun_out/bench.json
