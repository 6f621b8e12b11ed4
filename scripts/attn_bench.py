"""Micro-benchmark of the attention kernels on the DPT_Large shape (run on the GPU box)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from depth_estimation_b200 import ops
dev = torch.device("cuda:0")
for (B, T, H) in ((32, 577, 16), (8, 1009, 16), (1, 577, 12)):
    qkv = torch.randn(B * T, 3 * H * 64, device=dev).bfloat16()
    out = torch.empty(B * T, H * 64, device=dev, dtype=torch.bfloat16)
    for legacy in (False,):
        f = lambda: ops.attention(qkv, out, B, T, H, 0.125)
        for _ in range(3): f()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20): f()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 20
        print(f"B={B} T={T} heads={H} {'mma.sync' if legacy else 'tcgen05 '}: {ms*1e3:8.1f} us  {4.0*B*H*T*T*64/ms/1e9:7.1f} TFLOP/s", flush=True)
