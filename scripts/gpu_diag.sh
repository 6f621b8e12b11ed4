#!/bin/bash
# Diagnostics visit: HBM read/write/copy probe, DPT_Hybrid stage-by-stage drift, K1/K8 bandwidth cases with one
# ncu --set full capture each (after the plain run has exited 0).
mkdir -p gpurun_out
timeout 300 python scripts/hbm_probe.py > gpurun_out/hbm_probe.json 2> gpurun_out/hbm_probe.err; echo "hbm exit $?"; cat gpurun_out/hbm_probe.json
timeout 600 python tests/diag_hybrid.py > gpurun_out/diag_hybrid.log 2>&1; echo "diag exit $?"; tail -40 gpurun_out/diag_hybrid.log
timeout 600 python scripts/prepost_bench.py > gpurun_out/prepost.log 2>&1; echo "prepost exit $?"; cat gpurun_out/prepost.log
for c in k1a k1c k1d k8a k8d; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:'preprocess|postprocess' --launch-skip 3 -c 1 -f \
    -o gpurun_out/pp_$c python scripts/prepost_bench.py $c > gpurun_out/ncu_pp_$c.log 2>&1; echo "ncu $c exit $?"
done
