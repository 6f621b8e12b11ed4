timeout 600 python -m pytest tests/test_kernels_gpu.py -x -q -p no:cacheprovider -k "post or vis or sample or metric or k8" 2>&1 | tail -3
timeout 300 python scripts/prepost_bench.py k8a k8b k8c k8d 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except: print(l.strip()); continue
    print(d['case'][:50], round(d['us'],1), round(d['GBps']), round(d['frac_of_hbm_peak'],3))"
