#!/bin/bash
timeout 900 python -m pytest tests/test_e2e_gpu.py -m gpu -x -q -p no:cacheprovider -s -k "predict_matches_oracle" 2>&1 | grep -E "rel=|passed|failed|Error" | tail -14
