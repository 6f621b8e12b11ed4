"""fc1 probe (M=18464, N=4096, K=1024, bias + GELU, bf16 out) and the other ViT-L shapes."""
import sys
import torch
sys.path.insert(0, ".")
from depth_estimation_b200 import ops
dev = torch.device("cuda:0")
M = 18464


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


for name, N, K, act in (("fc1 + GELU", 4096, 1024, ops.ACT_GELU), ("qkv", 3072, 1024, ops.ACT_NONE), ("fc1 without GELU", 4096, 1024, ops.ACT_NONE)):
    a = (torch.randn((M, K), device=dev) * 0.5).bfloat16()
    w = (torch.randn((N, K), device=dev) * 0.05).bfloat16()
    b = torch.randn((N,), device=dev)
    o = torch.empty((M, N), dtype=torch.bfloat16, device=dev)
    us = t(lambda: ops.gemm(a, w, o, bias=b, act=act))
    ref = torch.nn.functional.linear(a[:256].float(), w.float(), b)
    if act == ops.ACT_GELU:
        ref = torch.nn.functional.gelu(ref)
    err = float((o[:256].float() - ref).abs().max() / ref.abs().max())
    print(f"{name:20s} {us:7.1f} us  {2.0 * M * N * K / us / 1e6:6.0f} TFLOP/s  err {err:.1e}", flush=True)
