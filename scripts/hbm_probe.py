"""What does HBM give a kernel that only reads, only writes, or copies?  K8 and the x2 upsample write 4-7x more
than they read; the copy figure in MEASURED_PEAKS.json (read+write bytes) need not be reachable for them."""
import json
import torch

dev = torch.device("cuda:0")
n = 1 << 30                                     # 2 GiB of bf16 per buffer, far beyond the 126 MB L2
a = torch.empty(n, dtype=torch.bfloat16, device=dev).normal_()
b = torch.empty_like(a)


def t(fn, iters=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1e-3)
    return best


by = a.numel() * 2
res = {"copy_GBps(read+write)": 2 * by / t(lambda: b.copy_(a)) / 1e9,
       "fill_GBps(write only)": by / t(lambda: b.zero_()) / 1e9,
       "fill1_GBps(write only)": by / t(lambda: b.fill_(1.0)) / 1e9,
       "sum_GBps(read only)": by / t(lambda: a.view(torch.int32).sum()) / 1e9,
       "expand_1to4_GBps": 0.0}
src = a[: n // 4]
res["expand_1to4_GBps"] = (by + by // 4) / t(lambda: b.view(4, -1).copy_(src.view(1, -1).expand(4, -1))) / 1e9
print(json.dumps(res))
