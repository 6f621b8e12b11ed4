#!/bin/bash
# visit 8: the allocation-free entry point alone in a fresh process, then the whole GPU suite (it runs last there too),
# then the default bench line and the bf16x3 step times
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > gpurun_out/pytest_zz_alone.log 2>&1; echo "zz alone exit $?"; tail -3 gpurun_out/pytest_zz_alone.log
timeout 2400 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/pytest_gpu_full.log 2>&1; echo "full gpu suite exit $?"; tail -8 gpurun_out/pytest_gpu_full.log
timeout 600 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench exit $?"; cut -c1-1500 gpurun_out/bench_default.json
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
for prec in bf16x3; do
timeout 600 python - <<PY 2>&1 | grep -v Loading
import time, numpy as np, torch
from depth_estimation_b200.depth_midas import DepthEstimator
for mt, B in (("DPT_Hybrid", 1), ("DPT_Large", 1), ("DPT_Large", 32)):
    est = DepthEstimator(mt, seed=0, precision="$prec")
    fr = torch.randint(0, 256, (B, 384, 384, 3), dtype=torch.uint8, device="cuda")
    for _ in range(5): est.predict_device(fr)
    torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 20 if B == 1 else 5
    e0.record()
    for _ in range(n): est.predict_device(fr)
    e1.record(); torch.cuda.synchronize()
    print(f"precision=$prec {mt} B={B}: {e0.elapsed_time(e1)/n:.3f} ms/step device-resident, {B*n/(e0.elapsed_time(e1)*1e-3):.0f} frames/s", flush=True)
    del est; torch.cuda.empty_cache()
PY
done
