"""Clock-stamp trace of one epilogue warp (CTA 0, warp 4) of the store epilogue: where a column unit's ~2 us go.
Needs the developer build with -DGEMM_TRACE (DEPTH_B200_LIB=scripts/_bin/libdepthb200_trace.so).

    python scripts/gemm_trace.py M N K [f32|bf16] [acc]
"""
import ctypes as C
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from depth_estimation_b200 import _lib, ops

M, N, K = (int(v) for v in sys.argv[1:4]) if len(sys.argv) > 3 else (18464, 1024, 1024)
f32 = len(sys.argv) > 4 and sys.argv[4] == "f32"
acc = len(sys.argv) > 5 and sys.argv[5] == "acc"
dev = torch.device("cuda:0")
a = (torch.randn((M, K), device=dev) * 0.5).bfloat16()
w = (torch.randn((N, K), device=dev) * 0.05).bfloat16()
b = torch.randn((N,), device=dev)
out = torch.zeros((M, N), dtype=torch.float32 if f32 else torch.bfloat16, device=dev)
for _ in range(3):
    ops.gemm(a, w, out, bias=b, accumulate=acc)
torch.cuda.synchronize()
lib = C.CDLL(_lib.LIB_PATH)
buf = (C.c_longlong * 16384)()
assert lib.de_gemm_trace_read(buf, 16384) == 0
t = np.array(buf, dtype=np.int64).reshape(-1, 16)
t = t[(t[:, 2] != 0)]
print(f"M={M} N={N} K={K} out={'f32' if f32 else 'bf16'} acc={acc}: {len(t)} units traced by warp 4 of CTA 0")
names = ["tile top->tfull", "", "(bias bar) ->unit", "tmem_ld", "math+release", "wait_read", "bar D", "STS", "fence", "bar F", "TMA issue"]
# per-unit deltas: 2->3 tmem, 3->4 math, 4->5 wait_read, 5->6 bar, 6->7 sts, 7->8 fence, 8->9 bar, 9->10 issue; unit period = stamp2[i+1]-stamp2[i]
d = {}
d["wait for the accumulator (tfull)"] = t[:, 1] - t[:, 0]
d["tmem_ld+wait"] = t[:, 3] - t[:, 2]
d["math (+acc release)"] = t[:, 4] - t[:, 3]
d["wait_read (prev store)"] = t[:, 5] - t[:, 4]
d["named barrier D"] = t[:, 6] - t[:, 5]
d["8x STS.128"] = t[:, 7] - t[:, 6]
d["fence.proxy.async"] = t[:, 8] - t[:, 7]
d["named barrier F"] = t[:, 9] - t[:, 8]
d["TMA store issue+commit"] = t[:, 10] - t[:, 9]
period = t[1:, 2] - t[:-1, 2]
for k, v in d.items():
    v = v[2:]
    print(f"  {k:34s} median {int(np.median(v)):6d}  p90 {int(np.percentile(v, 90)):6d} cycles")
print(f"  {'unit period (start->start)':26s} median {int(np.median(period[2:])):6d}  p90 {int(np.percentile(period[2:], 90)):6d} cycles")
print(f"  {'sum of the parts':26s} median {int(np.median(sum(v for v in d.values())[2:])):6d}")
