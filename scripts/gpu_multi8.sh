#!/bin/bash
# 8-GPU refresh of config #3 with the final tree: weak, weak + NCCL gather, strong
N=8
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 400 $TR bench.py --gpus $N --no-latency --no-cpu-baseline > gpurun_out/r02f_bench_n${N}_weak.json 2> gpurun_out/err_w.txt; echo "weak exit $?"
timeout 400 $TR bench.py --gpus $N --no-latency --no-cpu-baseline --gather > gpurun_out/r02f_bench_n${N}_gather.json 2> gpurun_out/err_g.txt; echo "gather exit $?"
timeout 400 $TR bench.py --gpus $N --no-latency --no-cpu-baseline --scaling strong --batch 32 > gpurun_out/r02f_bench_n${N}_strong.json 2> gpurun_out/err_s.txt; echo "strong exit $?"
python - <<PY
import json
for k in ("weak","gather","strong"):
    d=json.loads(open("gpurun_out/r02f_bench_n${N}_%s.json"%k).read().strip().splitlines()[-1])
    print(k, d["n_gpus"], round(d["value"],1), d["unit"], round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"],1), d["scaling"], d["config"]["parallelism"][:110], d["clocks"])
PY
