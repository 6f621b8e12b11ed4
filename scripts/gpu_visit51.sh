#!/bin/bash
# evidence for attention_ws2: ncu --set full of one launch at the bench shape, fresh launch list of the DPT_Large step
mkdir -p gpurun_out
timeout 100 python scripts/attn_bench.py > gpurun_out/attn_plain.log 2>&1; echo "plain exit $?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:attention_ws2 --launch-skip 5 -c 1 -f -o gpurun_out/r02_attention_ws2 python scripts/attn_bench.py > gpurun_out/ncu_attn.log 2>&1; echo "ncu attn exit $?"
M="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum"
B="--steps 1 --warmup 3 --no-e2e --no-cpu-baseline --no-latency"
timeout 300 python bench.py $B > gpurun_out/plain_large.log 2>&1; echo "plain large exit $?"
timeout 900 ncu --metrics $M --clock-control none -c 6000 --csv --log-file gpurun_out/r02f_launches_dpt_large_b32.csv python bench.py $B > gpurun_out/ncu_large.log 2>&1; echo "ncu large exit $?"
