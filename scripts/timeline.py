"""GPU timeline of one replayed step (kernel start / end times from CUPTI through torch.profiler): where the step time
that is in no kernel goes.  Prints per-kernel totals and the idle gaps attributed to the kernel
that precedes them; writes the raw events as JSON lines.

    python scripts/timeline.py DPT_Large 32 384 384 [out.jsonl] [precision]
"""
import collections
import json
import sys

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, ".")
from depth_estimation_b200.depth_midas import DepthEstimator  # noqa: E402


def short(name):
    name = name.replace("deb::", "").replace("(anonymous namespace)::", "")
    i = name.find("(")
    return name[:i] if i > 0 else name


def main():
    mt = sys.argv[1] if len(sys.argv) > 1 else "DPT_Large"
    B, W, H = (int(a) for a in (sys.argv[2:5] if len(sys.argv) > 4 else (32, 384, 384)))
    out = sys.argv[5] if len(sys.argv) > 5 else None
    prec = sys.argv[6] if len(sys.argv) > 6 and sys.argv[6] != "default" else None
    est = DepthEstimator(mt, seed=0, precision=prec)
    fr = torch.randint(0, 256, (B, H, W, 3), dtype=torch.uint8, device="cuda")
    for _ in range(5):
        est.predict_device(fr)
    torch.cuda.synchronize()
    steps = 3
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(steps):
            est.predict_device(fr)
        torch.cuda.synchronize()
    evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and "memcpy" not in e.name.lower()
           and "memset" not in e.name.lower()]
    evs.sort(key=lambda e: e.time_range.start)
    n = len(evs) // steps
    evs = evs[-n:]                                  # the last step
    t0, t1 = evs[0].time_range.start, evs[-1].time_range.end
    busy = collections.Counter()
    cnt = collections.Counter()
    gap_after = collections.Counter()
    overlap = 0.0
    rows = []
    prev_end = None
    for i, e in enumerate(evs):
        s, en = e.time_range.start, e.time_range.end
        nm = short(e.name)
        busy[nm] += en - s
        cnt[nm] += 1
        if prev_end is not None:
            g = s - prev_end
            if g >= 0:
                gap_after[short(evs[i - 1].name) + " -> " + nm] += g
            else:
                overlap += -g
        prev_end = en if prev_end is None else max(prev_end, en)
        rows.append({"i": i, "name": nm, "start_us": s - t0, "dur_us": en - s})
    span = t1 - t0
    tot_busy = sum(busy.values())
    print(f"{mt} B={B} {W}x{H} precision={prec}: step span {span / 1e3:.3f} ms, {len(evs)} kernels, summed kernel time "
          f"{tot_busy / 1e3:.3f} ms, overlap (PDL) {overlap / 1e3:.3f} ms, idle {sum(gap_after.values()) / 1e3:.3f} ms")
    print("-- kernel totals (ms, launches, us/launch)")
    for nm, t in busy.most_common(25):
        print(f"{t / 1e3:9.3f} {cnt[nm]:5d} {t / cnt[nm]:9.1f}  {nm[:110]}")
    print("-- idle gaps by (previous -> next) (ms)")
    for nm, t in gap_after.most_common(20):
        print(f"{t / 1e3:9.3f}  {nm[:150]}")
    if out:
        with open(out, "w") as f:
            for r in rows:
                f.write(json.dumps(r) + "\n")


if __name__ == "__main__":
    main()
