#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -p no:cacheprovider > gpurun_out/pytest_k.log 2>&1; echo "kernel tests exit $?"; tail -2 gpurun_out/pytest_k.log
timeout 600 python scripts/gemm_small_bench.py "convonly" 2>&1 | grep -v Loading | tee gpurun_out/gsb_conv.md
timeout 300 python bench.py --no-cpu-baseline --no-e2e --no-latency > gpurun_out/ab_large_new.json 2>/dev/null
DEPTH_B200_LIB=$PWD/scripts/_bin/libdepthb200_head.so timeout 300 python bench.py --no-cpu-baseline --no-e2e --no-latency > gpurun_out/ab_large_head.json 2>/dev/null
timeout 300 python bench.py --no-cpu-baseline --no-e2e --no-latency > gpurun_out/ab_large_new2.json 2>/dev/null
python - <<PY
import json
for f in ("ab_large_new","ab_large_head","ab_large_new2"):
    d=json.loads(open("gpurun_out/%s.json"%f).read().strip().splitlines()[-1])
    print(f, round(d["value"],1), "f/s", round(d["ms_per_step"],4), d["clocks"])
PY
