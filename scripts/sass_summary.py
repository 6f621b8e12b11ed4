"""Per-kernel count of the SASS mnemonics that prove the Blackwell data path (tcgen05 MMA, TMEM load/store, TMA
load / store / reduce) in the built library.  Usage: python scripts/sass_summary.py [lib] > profiles/rNN_sass_summary.md
Mnemonics (B200_PROFILING.md): UTCHMMA / UTCQMMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st, UTMALDG / UTMASTG /
UTMAREDG = cp.async.bulk.tensor load / store / reduce, UTCBAR = tcgen05.commit, SYNCS = mbarrier ops."""
import collections
import os
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                         "depth_estimation_b200", "libdepthb200.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
MN = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UTCBAR", "SYNCS", "HMMA", "FFMA2", "MUFU.EX2",
      "LDG.E.256", "STG.E.256", "LDS.128"]
kern = None
counts = collections.OrderedDict()
sizes = {}
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        kern = m.group(1)
        counts[kern] = collections.Counter()
        sizes[kern] = 0
        continue
    if kern is None:
        continue
    m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
    if not m:
        continue
    op = m.group(1)
    sizes[kern] += 1
    for k in MN:
        if op == k or op.startswith(k + ".") or (k in ("LDG.E.256", "STG.E.256", "LDS.128", "MUFU.EX2") and op.startswith(k.split(".")[0]) and all(p in op for p in k.split(".")[1:])):
            counts[kern][k] += 1
            break


def demangle(n):
    r = subprocess.run(["cu++filt", n], capture_output=True, text=True)
    s = r.stdout.strip() or n
    s = s.replace("(int)", "")
    s = re.sub(r"\(.*", "", s)
    return s.replace("void ", "")


arch = re.findall(r"arch = (sm_\w+)", out)
print(f"# SASS instruction summary of `{os.path.basename(lib)}` ({', '.join(sorted(set(arch)))}; cuobjdump -sass, scripts/sass_summary.py)\n")
print("| kernel | SASS instr | " + " | ".join(MN) + " |")
print("|---|---:|" + "---:|" * len(MN))
tot = collections.Counter()
for k, c in counts.items():
    tot.update(c)
    print(f"| `{demangle(k)}` | {sizes[k]} | " + " | ".join(str(c[m]) if c[m] else "" for m in MN) + " |")
print(f"| **total** | {sum(sizes.values())} | " + " | ".join(str(tot[m]) for m in MN) + " |")
