"""Head tail (conv3x3 128->32 + ReLU + 1x1 + ReLU) at the DPT_Large bench shape: halo-tiled kernel vs the general
implicit-GEMM kernel's head-dot epilogue."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from depth_estimation_b200 import ops
from depth_estimation_b200.weights import pack_conv3x3

dev = torch.device("cuda:0")
B, H, W, Cin = 32, 384, 384, 128
x = torch.randn((B, H, W, Cin), device=dev).bfloat16()
w = pack_conv3x3((torch.randn((32, Cin, 3, 3), device=dev) * 0.05))
bias, w2 = torch.randn(32, device=dev), torch.randn(32, device=dev)
out = torch.empty((B, H, W), dtype=torch.float32, device=dev)


def timeit(fn, iters=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


fl = 2.0 * B * H * W * 32 * 9 * Cin
t = timeit(lambda: ops.head_conv(x, w, bias, w2, 0.7, out))
print(f"head_conv (halo, resident filters): {t:.3f} ms  {fl / t / 1e9:.0f} TFLOP/s  input {x.numel() * 2 / t / 1e6:.0f} GB/s")
t = timeit(lambda: ops.gemm(x, w, out, conv=(B, H, W, Cin), bias=bias, act=1, head=(w2, 0.7), N=32))
print(f"gemm head-dot epilogue            : {t:.3f} ms  {fl / t / 1e9:.0f} TFLOP/s")

lo = torch.randn((B, H // 2, W // 2, Cin), device=dev).bfloat16()
t = timeit(lambda: ops.head_up_conv(lo, w, bias, w2, 0.7, out, True))
print(f"head_up_conv (fused x2 upsample)  : {t:.3f} ms  {fl / t / 1e9:.0f} TFLOP/s")

# x2 bilinear upsample at the two big decoder shapes
for (h, w, c) in ((192, 192, 128), (96, 96, 256)):
    src = torch.randn((B, h, w, c), device=dev).bfloat16()
    dst = torch.empty((B, 2 * h, 2 * w, c), dtype=torch.bfloat16, device=dev)
    t = timeit(lambda: ops.upsample2x(src, dst, True))
    by = src.numel() * 2 + dst.numel() * 2
    print(f"upsample2x {h}x{w}x{c}: {t:.3f} ms  {by / t / 1e6:.0f} GB/s")
