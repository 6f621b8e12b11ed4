#!/bin/bash
# Multi-GPU visit (N = $1): config #3 strong scaling, NCCL gather inside the step, config #5 1080p sweep points, weak-scaling line.
N=$1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8 > gpurun_out/gpus_n$N.txt
timeout 600 $TR bench.py --gpus $N --no-latency --no-cpu-baseline > gpurun_out/bench_n${N}_weak.json 2> gpurun_out/bench_n${N}_weak.err; echo "weak exit $?"; cat gpurun_out/bench_n${N}_weak.json | cut -c1-600
timeout 600 $TR bench.py --gpus $N --no-latency --no-cpu-baseline --gather > gpurun_out/bench_n${N}_gather.json 2> gpurun_out/bench_n${N}_gather.err; echo "gather exit $?"; cat gpurun_out/bench_n${N}_gather.json | cut -c1-900
timeout 600 $TR bench.py --gpus $N --no-latency --no-cpu-baseline --scaling strong --batch 32 > gpurun_out/bench_n${N}_strong.json 2> gpurun_out/bench_n${N}_strong.err; echo "strong exit $?"; cat gpurun_out/bench_n${N}_strong.json | cut -c1-600
timeout 900 $TR scripts/sweep_config5.py --batches $2 > gpurun_out/sweep5_n${N}.jsonl 2> gpurun_out/sweep5_n${N}.err; echo "sweep exit $?"; cat gpurun_out/sweep5_n${N}.jsonl
timeout 300 python -m pytest tests/test_e2e_gpu.py -m gpu -q -p no:cacheprovider -k nccl_gather 2>&1 | tail -2
