"""Summarise an `ncu --set full` report (.ncu-rep) into a markdown table: one row per captured launch with the
metrics the roofline discussion in DESIGN.md uses.  Usage: summarize_ncu_full.py <rep> [title]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
title = sys.argv[2] if len(sys.argv) > 2 else rep
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
col = {n: i for i, n in enumerate(hdr)}


def get(r, name, scale=1.0, fmt="{:.1f}"):
    i = col.get(name)
    if i is None or r[i] in ("", "n/a"):
        return "-"
    v = float(r[i].replace(",", ""))
    u = units[i]
    if u == "Kbyte": v *= 1e3
    elif u == "Mbyte": v *= 1e6
    elif u == "Gbyte": v *= 1e9
    elif u == "ns" : v *= 1e-3
    elif u == "ms": v *= 1e3
    return fmt.format(v * scale)


COLS = [("us", "gpu__time_duration.sum", 1.0, "{:.1f}"),
        ("dram rd MB", "dram__bytes_read.sum", 1e-6, "{:.1f}"),
        ("dram wr MB", "dram__bytes_write.sum", 1e-6, "{:.1f}"),
        ("dram thr %", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", 1.0, "{:.1f}"),
        ("sm thr %", "sm__throughput.avg.pct_of_peak_sustained_elapsed", 1.0, "{:.1f}"),
        ("issue active %", "sm__inst_issued.avg.pct_of_peak_sustained_active", 1.0, "{:.1f}"),
        ("tensor pipe %", "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active", 1.0, "{:.1f}"),
        ("L2 hit %", "lts__t_sector_hit_rate.pct", 1.0, "{:.1f}"),
        ("occ %", "sm__warps_active.avg.pct_of_peak_sustained_active", 1.0, "{:.1f}"),
        ("regs", "launch__registers_per_thread", 1.0, "{:.0f}"),
        ("smem KB", "launch__shared_mem_per_block_dynamic", 1e-3, "{:.0f}")]
avail = [c for c in COLS if c[1] in col]
if "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active" not in col:
    for alt in ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
                "sm__inst_executed_pipe_tensor.avg.pct_of_peak_sustained_active"):
        if alt in col:
            avail.append(("tensor pipe %", alt, 1.0, "{:.1f}"))
            break
print(title + "\n")
print("| kernel | " + " | ".join(c[0] for c in avail) + " |")
print("|---|" + "---:|" * len(avail))
for r in data:
    name = r[col["Kernel Name"]].split("(")[0]
    print(f"| `{name}` | " + " | ".join(get(r, c[1], c[2], c[3]) for c in avail) + " |")
