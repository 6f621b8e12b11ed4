#!/bin/bash
# attention with two softmax threads per row (DEB_ATTN_V2=1): parity under a short timeout, then microbenchmark v1 / v2
mkdir -p gpurun_out
DEB_ATTN_V2=1 timeout 120 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -p no:cacheprovider -k "test_attention" 2>&1 | tail -4
echo "== v1"; timeout 100 python scripts/attn_bench.py 2>&1 | grep -v Loading
echo "== v2"; DEB_ATTN_V2=1 timeout 100 python scripts/attn_bench.py 2>&1 | grep -v Loading
