"""proj-GEMM probe (M=18464, N=1024, K=1024): fp32 reduce-add epilogue vs plain fp32 store vs bf16 store."""
import sys
import torch
sys.path.insert(0, ".")
from depth_estimation_b200 import ops
dev = torch.device("cuda:0")
M, N, K = 18464, 1024, 1024
a = (torch.randn((M, K), device=dev) * 0.5).bfloat16()
w = (torch.randn((N, K), device=dev) * 0.05).bfloat16()
b = torch.randn((N,), device=dev)
x = torch.zeros((M, N), dtype=torch.float32, device=dev)
o16 = torch.empty((M, N), dtype=torch.bfloat16, device=dev)


def t(fn, n=30):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n


fl = 2.0 * M * N * K
for name, fn in (("fp32 reduce-add (proj as shipped)", lambda: ops.gemm(a, w, x, bias=b, accumulate=True)),
                 ("fp32 plain store", lambda: ops.gemm(a, w, x, bias=b)),
                 ("bf16 store", lambda: ops.gemm(a, w, o16, bias=b))):
    us = t(fn)
    print(f"{name:36s} {us:7.1f} us  {fl / us / 1e6:6.0f} TFLOP/s", flush=True)
