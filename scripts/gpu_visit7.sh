#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -q -p no:cacheprovider -x -k "split3 or splitk or layernorm or attention or cast" > gpurun_out/pytest_k.log 2>&1; echo "kernel exit $?"; tail -5 gpurun_out/pytest_k.log
timeout 1500 python -m pytest tests/test_e2e_gpu.py tests/test_oracle_hf_crosscheck.py tests/test_engine_gpu.py -m gpu -q -p no:cacheprovider -s > gpurun_out/pytest_e2e.log 2>&1; echo "e2e exit $?"
grep -E "rel=|passed|failed|Error|FAILED" gpurun_out/pytest_e2e.log | tail -30
for prec in bf16 bf16x3; do
timeout 600 python - <<PY 2>&1 | grep -v Loading
import time, numpy as np, torch
from depth_estimation_b200.depth_midas import DepthEstimator
for mt, B in (("DPT_Hybrid", 1), ("DPT_Large", 1), ("DPT_Large", 32)):
    if "$prec" == "bf16" and mt == "DPT_Large": continue
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
