#!/bin/bash
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -p no:cacheprovider -k "conv3x3" 2>&1 | tail -12
