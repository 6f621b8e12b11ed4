#!/bin/bash
# evidence run of the current tree: full GPU suite, default bench (with cpu_baseline), the other BASELINE configs, timelines
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/pytest_gpu_full.log 2>&1; echo "full gpu suite exit $?"; tail -3 gpurun_out/pytest_gpu_full.log
timeout 900 python bench.py > gpurun_out/r02_bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench exit $?"
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r02_bench_ref_n1.json 2>/dev/null; echo "ref exit $?"
timeout 300 python bench.py --width 1920 --height 1080 --no-cpu-baseline --no-latency > gpurun_out/r02_bench_1080p_b32.json 2>/dev/null
timeout 300 python bench.py --model MiDaS_small --width 1280 --height 720 --no-cpu-baseline --no-latency > gpurun_out/r02_bench_small_720p_b32.json 2>/dev/null
timeout 300 python bench.py --model MiDaS_small --width 640 --height 480 --batch 1 --no-cpu-baseline > gpurun_out/r02_bench_small_480p_b1.json 2>/dev/null
timeout 300 python bench.py --model DPT_Hybrid --no-cpu-baseline --no-latency > gpurun_out/r02_bench_hybrid_b32.json 2>/dev/null
timeout 300 python bench.py --workload stream720p --batch 16 --no-cpu-baseline > gpurun_out/r02_bench_stream720p_b16.json 2>/dev/null
timeout 300 python bench.py --workload stream720p --batch 32 --no-cpu-baseline > gpurun_out/r02_bench_stream720p_b32.json 2>/dev/null
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/r02_bench_*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split("/")[-1], round(d["value"],1), d["unit"], round(d["ms_per_step"],3), "ms/step e2e", round((d.get("e2e") or {}).get("value",0),1), "lat", (d.get("latency") or {}).get("p50_ms"), d.get("clocks"))
    except Exception as e: print(f, "ERR", e)
PY
timeout 600 python scripts/timeline.py DPT_Large 32 384 384 gpurun_out/tl4_large_b32.jsonl > gpurun_out/tl4_large_b32.txt 2>&1
timeout 600 python scripts/timeline.py MiDaS_small 32 1280 720 gpurun_out/tl4_small_b32.jsonl > gpurun_out/tl4_small_b32.txt 2>&1
timeout 600 python scripts/timeline.py DPT_Hybrid 1 384 384 gpurun_out/tl4_hybrid_b1.jsonl > gpurun_out/tl4_hybrid_b1.txt 2>&1
grep "step span" gpurun_out/tl4_*.txt
