#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_e2e_gpu.py -m gpu -x -q -p no:cacheprovider > gpurun_out/pytest_pp.log 2>&1; echo "kernel+e2e tests exit $?"; tail -2 gpurun_out/pytest_pp.log
for v in head new; do
  unset DEPTH_B200_LIB
  if [ $v = head ]; then export DEPTH_B200_LIB=$PWD/scripts/_bin/libdepthb200_head.so; fi
  echo "== $v"
  timeout 300 python scripts/prepost_bench.py 2>&1 | grep -v Loading | tee gpurun_out/prepost_$v.jsonl | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: print(l.rstrip()); continue
    print(d['kernel'][:3], d['case'][:55], round(d['us'],1),'us', round(d['GBps']), 'GB/s', round(d['frac_of_hbm_peak'],3))"
done
