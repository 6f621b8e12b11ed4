timeout 900 python -m pytest tests/test_e2e_gpu.py -x -q -p no:cacheprovider -s -k "predict_matches or independent" 2>&1 | grep -E "rel=|passed|failed|Error|assert" | tail -14
