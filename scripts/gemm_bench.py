"""Micro-benchmark of the tcgen05 GEMM on the DPT_Large shapes (run on the GPU box)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from depth_estimation_b200 import ops

dev = torch.device("cuda:0")
M = int(os.environ.get("GB_M", 18464))
def bench(name, N, K, cg, bn=256, **kw):
    a = torch.randn(M, K, device=dev).bfloat16()
    w = (torch.randn(N, K, device=dev) * 0.02).bfloat16()
    bias = torch.randn(N, device=dev)
    out_dtype = kw.pop("out_dtype", torch.bfloat16)
    out = torch.empty(M, N, dtype=out_dtype, device=dev)
    add1 = out if kw.pop("inplace_add", False) else None
    if add1 is not None:
        out.zero_()
    f = lambda: ops.gemm(a, w, out, bias=bias, add1=add1, block_n=bn, cta_group=cg, **kw)
    for _ in range(3):
        f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 20
    e0.record()
    for _ in range(n):
        f()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    print(f"{name:28s} N={N:5d} K={K:5d} cg={cg} bn={bn}: {ms*1e3:8.1f} us  {2.0*M*N*K/ms/1e9:7.1f} TFLOP/s", flush=True)

for dbg in (0, 1):
    print("epilogue", dbg, "(1 = force general epilogue)")
    for cg in (1, 2):
        bench("qkv  bf16+bias", 3072, 1024, cg, epilogue=dbg)
        bench("fc1  bf16+bias+gelu", 4096, 1024, cg, act=2, epilogue=dbg)
        bench("proj f32+bias (no res)", 1024, 1024, cg, out_dtype=torch.float32, epilogue=dbg)
        bench("fc2  f32+bias (no res)", 1024, 4096, cg, out_dtype=torch.float32, epilogue=dbg)
bench("proj f32+bias+res(inplace)", 1024, 1024, 2, out_dtype=torch.float32, inplace_add=True)
bench("fc2  f32+bias+res(inplace)", 1024, 4096, 2, out_dtype=torch.float32, inplace_add=True)
