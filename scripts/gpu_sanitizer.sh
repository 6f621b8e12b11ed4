#!/bin/bash
# compute-sanitizer over the kernel unit tests (SURVEY section 5 plan): memcheck on everything, racecheck + synccheck on the
# kernels with hand-rolled mbarrier / TMEM / arrival-counter protocols.  Summaries -> gpurun_out/sanitizer_*.log
mkdir -p gpurun_out
CS=/usr/local/cuda/bin/compute-sanitizer
K1="gemm or attention or head or depthwise or groupnorm or splitk"
timeout 1500 $CS --tool memcheck --error-exitcode 86 --launch-timeout 0 --print-limit 20 \
  python -m pytest tests/test_kernels_gpu.py -m gpu -q -p no:cacheprovider -x > gpurun_out/sanitizer_memcheck.log 2>&1; echo "memcheck exit $?"
grep -E "ERROR SUMMARY|passed|failed|Invalid|out of bounds|misaligned" gpurun_out/sanitizer_memcheck.log | tail -8
timeout 1200 $CS --tool racecheck --error-exitcode 86 --print-limit 20 \
  python -m pytest tests/test_kernels_gpu.py -m gpu -q -p no:cacheprovider -x -k "$K1" > gpurun_out/sanitizer_racecheck.log 2>&1; echo "racecheck exit $?"
grep -E "RACECHECK SUMMARY|passed|failed|hazard" gpurun_out/sanitizer_racecheck.log | tail -8
timeout 900 $CS --tool synccheck --error-exitcode 86 --print-limit 20 \
  python -m pytest tests/test_kernels_gpu.py -m gpu -q -p no:cacheprovider -x -k "$K1" > gpurun_out/sanitizer_synccheck.log 2>&1; echo "synccheck exit $?"
grep -E "ERROR SUMMARY|passed|failed|Barrier" gpurun_out/sanitizer_synccheck.log | tail -8
timeout 900 $CS --tool memcheck --error-exitcode 86 --print-limit 20 \
  python -m pytest tests/test_e2e_gpu.py -m gpu -q -p no:cacheprovider -x -k "predict_matches_oracle and (MiDaS_small or Hybrid-384)" > gpurun_out/sanitizer_memcheck_e2e.log 2>&1; echo "memcheck e2e exit $?"
grep -E "ERROR SUMMARY|passed|failed" gpurun_out/sanitizer_memcheck_e2e.log | tail -5
