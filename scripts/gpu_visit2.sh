#!/bin/bash
# Visit 2: new depthwise kernel + split3 decoder; MiDaS_small launch list and depthwise captures.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -q -p no:cacheprovider -x -k "depthwise or split3 or upsample or groupnorm or gemm" > gpurun_out/pytest_k.log 2>&1; echo "kernel tests exit $?"; tail -5 gpurun_out/pytest_k.log
timeout 1200 python -m pytest tests/test_e2e_gpu.py -m gpu -q -p no:cacheprovider -s > gpurun_out/pytest_e2e.log 2>&1; echo "e2e exit $?"
grep -E "rel=|passed|failed|Error|FAILED" gpurun_out/pytest_e2e.log | tail -30
BENCH_DETAIL=1 timeout 600 python bench.py --model MiDaS_small --width 1280 --height 720 --no-cpu-baseline --no-latency > gpurun_out/bench_small.json 2> gpurun_out/bench_small.err; echo "bench small exit $?"
cat gpurun_out/bench_small.json
BENCH_DETAIL=1 timeout 600 python bench.py --model DPT_Hybrid --batch 1 --no-cpu-baseline > gpurun_out/bench_hybrid_b1.json 2> gpurun_out/bench_hybrid_b1.err; echo "bench hybrid b1 exit $?"
cat gpurun_out/bench_hybrid_b1.json
BENCH_DETAIL=1 timeout 600 python bench.py --model DPT_Hybrid --no-cpu-baseline --no-latency > gpurun_out/bench_hybrid_b32.json 2> gpurun_out/bench_hybrid_b32.err; echo "bench hybrid b32 exit $?"
cat gpurun_out/bench_hybrid_b32.json
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches_small.csv \
  python bench.py --model MiDaS_small --width 1280 --height 720 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --no-latency > gpurun_out/ncu_small.log 2>&1; echo "ncu small exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:depthwise --launch-skip 78 -c 26 -f -o gpurun_out/dw_full \
  python bench.py --model MiDaS_small --width 1280 --height 720 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --no-latency > gpurun_out/ncu_dw_full.log 2>&1; echo "ncu dw full exit $?"
