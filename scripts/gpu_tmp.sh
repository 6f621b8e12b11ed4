timeout 300 python -m pytest tests/test_kernels_gpu.py -x -q -p no:cacheprovider -k "attention" 2>&1 | tail -2
timeout 300 python scripts/attn_bench.py 2>&1 | grep tcgen05
