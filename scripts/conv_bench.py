"""Implicit-GEMM 3x3 conv micro-benchmark on the DPT decoder shapes (B=32): cta_group / block_n variants."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from depth_estimation_b200 import ops
from depth_estimation_b200.weights import pack_conv3x3

dev = torch.device("cuda:0")
B = 32


def bench(name, H, W, Cin, Cout, cg, bn=0, dual=False, add=False):
    x = torch.randn((B, H, W, Cin), device=dev).bfloat16()
    w = pack_conv3x3(torch.randn((Cout, Cin, 3, 3), device=dev) * 0.02)
    bias = torch.randn(Cout, device=dev)
    out = torch.empty((B, H, W, Cout), dtype=torch.bfloat16, device=dev)
    kw = {}
    if dual:
        kw["out_relu"] = torch.empty_like(out)
    if add:
        kw["add1"] = torch.randn((B * H * W, Cout), device=dev).bfloat16()
    f = lambda: ops.gemm(x, w, out.view(-1, Cout), conv=(B, H, W, Cin), N=Cout, bias=bias, block_n=bn, cta_group=cg, **kw)
    for _ in range(3):
        f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 10
    e0.record()
    for _ in range(n):
        f()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    print(f"{name:34s} {H}x{W} {Cin}->{Cout} cg={cg} bn={bn}: {ms * 1e3:8.1f} us  {2.0 * B * H * W * Cout * 9 * Cin / ms / 1e9:7.1f} TFLOP/s", flush=True)


for cg in (1, 2):
    bench("head conv0 (store epilogue)", 192, 192, 256, 128, cg, 128)
bench("head conv0 as bn=64", 192, 192, 256, 128, 1, 64)
for cg in (1, 2):
    bench("rcu conv1 (store epilogue)", 96, 96, 256, 256, cg)
    bench("rcu conv2 + add (general)", 96, 96, 256, 256, cg, add=True)
    bench("layer_rn dual store (general)", 96, 96, 256, 256, cg, dual=True)
