#!/bin/bash
# config #5 sweep points at N = $1 GPUs, batch list $2
N=$1
mkdir -p gpurun_out
if [ "$N" = "1" ]; then
  timeout 1500 python scripts/sweep_config5.py --batches $2 > gpurun_out/sweep5_n1.jsonl 2> gpurun_out/sweep5_n1.err; echo "sweep exit $?"
else
  timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 scripts/sweep_config5.py --batches $2 > gpurun_out/sweep5_n$N.jsonl 2> gpurun_out/sweep5_n$N.err; echo "sweep exit $?"
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --no-latency --no-cpu-baseline --scaling strong --batch 32 > gpurun_out/bench_n${N}_strong.json 2> gpurun_out/bench_n${N}_strong.err; echo "strong exit $?"; cut -c1-400 gpurun_out/bench_n${N}_strong.json
fi
cat gpurun_out/sweep5_n$N.jsonl; tail -3 gpurun_out/sweep5_n$N.err
