#!/bin/bash
# Runs every GPU kernel test function in its own process (a trapped kernel poisons the CUDA context),
# collecting logs under gpurun_out/.
mkdir -p gpurun_out
LOG=gpurun_out/kernel_tests.log
: > $LOG
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv >> $LOG 2>&1
FUNCS=${FUNCS:-$(grep -oE "^def (test_[a-z0-9_]+)" tests/test_kernels_gpu.py | awk '{print $2}')}
for f in $FUNCS; do
  echo "=== $f" >> $LOG
  timeout 300 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "$f" --no-header -p no:cacheprovider 2>&1 | tail -40 >> $LOG
  echo "=== $f exit ${PIPESTATUS[0]}" >> $LOG
done
grep -E "^=== .* exit|passed|failed" $LOG
