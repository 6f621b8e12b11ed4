#!/bin/bash
# One GPU-box visit: GPU tests, smoke, both bench arms, ncu launch list, per-launch DRAM traffic of the GEMM kernel,
# one --set full capture of the ViT-block GEMMs.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" | tee -a gpurun_out/smoke.log
BENCH_DETAIL=1 timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref exit $?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --no-latency > gpurun_out/ncu_list.log 2>&1; echo "ncu list exit $?"
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:gemm_kernel \
  -c 1200 --csv --log-file gpurun_out/gemm_traffic.csv \
  python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --no-latency > gpurun_out/ncu_traffic.log 2>&1; echo "ncu traffic exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_kernel --launch-skip 40 -c 12 -f -o gpurun_out/gemm_full \
  python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --no-latency > gpurun_out/ncu_full.log 2>&1; echo "ncu full exit $?"
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/bench.json
