#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_engine_gpu.py -m gpu -x -q -p no:cacheprovider > gpurun_out/pytest_k.log 2>&1; echo "kernel+engine tests exit $?"; tail -3 gpurun_out/pytest_k.log
timeout 600 python scripts/gemm_small_bench.py "working tree" 2>&1 | grep -v Loading | tee gpurun_out/gsb_new.md
for lib in head new; do
  if [ $lib = head ]; then export DEPTH_B200_LIB=$PWD/scripts/_bin/libdepthb200_head.so; else unset DEPTH_B200_LIB; fi
  timeout 300 python bench.py --model MiDaS_small --width 1280 --height 720 --no-cpu-baseline --no-latency --no-e2e > gpurun_out/ab_small_$lib.json 2>/dev/null
  timeout 300 python bench.py --model DPT_Hybrid --batch 1 --no-cpu-baseline --no-e2e > gpurun_out/ab_hybrid_$lib.json 2>/dev/null
  timeout 300 python bench.py --no-cpu-baseline --no-e2e --no-latency > gpurun_out/ab_large_$lib.json 2>/dev/null
  python - <<PY
import json
for m in ("small","hybrid","large"):
    d=json.loads(open("gpurun_out/ab_%s_$lib.json"%m).read().strip().splitlines()[-1])
    print("$lib", m, round(d["value"],1), "f/s", round(d["ms_per_step"],4), "ms/step", d.get("latency",{}).get("p50_ms"), d["clocks"]["sm_mhz"])
PY
done
