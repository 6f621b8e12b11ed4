"""Low-resolution 1x1 expand convolutions of MiDaS_small (short K, small M, wide N): automatic tile width against forced
128 / 64.  CUDA events around 50 launches inside a CUDA graph (PDL active, as in the model)."""
import sys

import torch

sys.path.insert(0, ".")
from depth_estimation_b200 import ops  # noqa: E402

SHAPES = [("expand 96->576 @8x16", 4096, 96, 576), ("expand 136->816 @8x16", 4096, 136, 816),
          ("expand 232->1392 @4x8", 1024, 232, 1392), ("expand 48->288 @16x32", 16384, 48, 288),
          ("expand 32->192 @32x64", 65536, 32, 192)]
dev = torch.device("cuda:0")
print("| shape | M | K | N | auto us | block_n=128 us | block_n=64 us |")
print("|---|---:|---:|---:|---:|---:|---:|")
for name, M, K, N in SHAPES:
    g = torch.Generator().manual_seed(1)
    a = (torch.randn((M, K), generator=g) * 0.5).to(dev).bfloat16()
    w = (torch.randn((N, K), generator=g) * 0.1).to(dev).bfloat16()
    b = torch.randn((N,), generator=g).to(dev)
    ref = torch.nn.functional.linear(a.float(), w.float(), b).clamp(0, 6)
    res = []
    for bn in (0, 128, 64):
        out = torch.zeros((M, N), dtype=torch.bfloat16, device=dev)

        def run():
            ops.gemm(a, w, out, bias=b, act=ops.ACT_RELU6, block_n=bn)
        for _ in range(3):
            run()
        torch.cuda.synchronize()
        err = float((out.float() - ref).abs().max() / ref.abs().max())
        assert err < 1e-2, (name, bn, err)
        s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=s):
                for _ in range(50):
                    run()
        torch.cuda.synchronize()
        gr.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        gr.replay()
        e1.record()
        torch.cuda.synchronize()
        res.append(e0.elapsed_time(e1) * 1e3 / 50)
    print(f"| {name} | {M} | {K} | {N} | {res[0]:.1f} | {res[1]:.1f} | {res[2]:.1f} |", flush=True)
