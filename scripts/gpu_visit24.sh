#!/bin/bash
# round-2 ncu evidence for the dominant kernel with the final tree: per-launch time + DRAM bytes of the gemm_kernel launches of
# one DPT_Large step, and one --set full capture of the ViT shapes
mkdir -p gpurun_out
B="--steps 1 --warmup 3 --no-e2e --no-cpu-baseline --no-latency"
timeout 300 python bench.py $B > gpurun_out/plain_large.log 2>&1; echo "plain exit $?"
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:gemm_kernel -c 4000 --csv --log-file gpurun_out/r02_gemm_traffic.csv \
  python bench.py $B > gpurun_out/ncu_gt.log 2>&1; echo "ncu traffic exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_kernel --launch-skip 960 -c 12 -f -o gpurun_out/r02_gemm_full \
  python bench.py $B > gpurun_out/ncu_gf.log 2>&1; echo "ncu full exit $?"
