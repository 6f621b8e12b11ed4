"""Narrow-N 1x1 convolutions of MiDaS_small without residual (N = 24 / 32 / 48 / 64): the automatic tile choice
(32-wide tiles -> general epilogue with scattered 8-byte stores) against a forced 64-wide tile (TMA-store epilogue, the
columns beyond N are clipped by the tensor map).  CUDA events around 50 back-to-back launches."""
import sys

import torch

sys.path.insert(0, ".")
from depth_estimation_b200 import ops  # noqa: E402

SHAPES = [("ds pw 32->24 @64x128", 262144, 32, 24), ("mb project 144->24 @64x128", 262144, 144, 24),
          ("mb project 144->32 @32x64", 65536, 144, 32), ("mb project 192->48 @16x32", 16384, 192, 48),
          ("mb project 288->96 @8x16", 4096, 288, 96), ("dec 1x1 64->64 @32x64", 65536, 64, 64)]
dev = torch.device("cuda:0")
print("| shape | M | K | N | auto us | block_n=64 us |")
print("|---|---:|---:|---:|---:|---:|")
for name, M, K, N in SHAPES:
    g = torch.Generator().manual_seed(1)
    a = (torch.randn((M, K), generator=g) * 0.5).to(dev).bfloat16()
    w = (torch.randn((N, K), generator=g) * 0.1).to(dev).bfloat16()
    b = torch.randn((N,), generator=g).to(dev)
    ref = torch.nn.functional.linear(a.float(), w.float(), b)
    res = []
    for bn in (0, 64):
        out = torch.zeros((M, N), dtype=torch.bfloat16, device=dev)

        def run():
            ops.gemm(a, w, out, bias=b, act=ops.ACT_NONE, block_n=bn)
        for _ in range(3):
            run()
        torch.cuda.synchronize()
        err = float((out.float() - ref).abs().max() / ref.abs().max())
        assert err < 1e-2, (name, bn, err)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(50):
            run()
        e1.record()
        torch.cuda.synchronize()
        res.append(e0.elapsed_time(e1) * 1e3 / 50)
    print(f"| {name} | {M} | {K} | {N} | {res[0]:.1f} | {res[1]:.1f} |", flush=True)
