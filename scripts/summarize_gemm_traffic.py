"""gemm_traffic.csv (ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum -k regex:gemm_kernel
over `bench.py --steps 1 --warmup 3`) -> profiles/r01_gemm_traffic.json: DRAM bytes per launch of the last step's
GEMM launches (bench.py quotes `dram_bytes_per_launch` as roofline.traffic).
Usage: summarize_gemm_traffic.py <csv> <launches_per_step> <out.json>"""
import csv
import json
import sys
from collections import defaultdict

path, per_step, out = sys.argv[1], int(sys.argv[2]), sys.argv[3]
with open(path, newline="") as f:
    lines = [l for l in f if l.startswith('"')]
launch = defaultdict(dict)
names = {}
for r in csv.DictReader(lines):
    i = int(r["ID"])
    names[i] = r["Kernel Name"].split("(")[0].replace("deb::", "").strip()
    v = float(r["Metric Value"].replace(",", ""))
    u = r["Metric Unit"]
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1, "ms": 1e3}.get(u, 1)
    launch[i][r["Metric Name"]] = v * scale
ids = sorted(launch)[-per_step:]
rd = sum(launch[i]["dram__bytes_read.sum"] for i in ids)
wr = sum(launch[i]["dram__bytes_write.sum"] for i in ids)
per = defaultdict(lambda: [0, 0.0, 0.0, 0.0])
for i in ids:
    p = per[names[i]]
    p[0] += 1
    p[1] += launch[i]["dram__bytes_read.sum"]
    p[2] += launch[i]["dram__bytes_write.sum"]
    p[3] += launch[i]["gpu__time_duration.sum"]
doc = {"source": "ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none "
                 f"-k regex:gemm_kernel; python bench.py --steps 1 --warmup 3 (last {per_step} gemm_kernel launches = one "
                 "32-frame DPT_Large step)",
       "launches": len(ids), "dram_read_bytes_per_step": rd, "dram_write_bytes_per_step": wr,
       "dram_bytes_per_launch": (rd + wr) / len(ids),
       "per_variant": {k: {"launches": v[0], "read_bytes_per_launch": v[1] / v[0], "write_bytes_per_launch": v[2] / v[0],
                           "us_per_launch_cold": v[3] / v[0]} for k, v in per.items()}}
with open(out, "w") as f:
    json.dump(doc, f, indent=1)
print(json.dumps({k: doc[k] for k in ("launches", "dram_read_bytes_per_step", "dram_write_bytes_per_step", "dram_bytes_per_launch")}))
