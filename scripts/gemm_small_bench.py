"""Per-shape timing of the small / short-K GEMMs that dominate MiDaS_small (720p, 32 frames) and DPT_Hybrid batch 1:
CUDA events around N back-to-back launches of ONE shape (L2-resident like in the model), and the same inside a CUDA
graph (PDL active).  Used to tune the tile / split-K heuristics of de_gemm.

    python scripts/gemm_small_bench.py [label]
"""
import sys

import torch

sys.path.insert(0, ".")
from depth_estimation_b200 import ops  # noqa: E402

# (name, M, K, N, residual, act)
SHAPES = [
    ("mb expand 24->144 @64x128", 262144, 24, 144, False, ops.ACT_RELU6),
    ("mb project 144->24 @64x128", 262144, 144, 24, False, ops.ACT_NONE),
    ("mb expand 32->192 @32x64", 65536, 32, 192, False, ops.ACT_RELU6),
    ("mb project 192->32 @32x64 +res", 65536, 192, 32, True, ops.ACT_NONE),
    ("mb expand 48->288 @16x32", 16384, 48, 288, False, ops.ACT_RELU6),
    ("mb project 288->48 @16x32 +res", 16384, 288, 48, True, ops.ACT_NONE),
    ("mb expand 96->576 @8x16", 4096, 96, 576, False, ops.ACT_RELU6),
    ("mb project 576->96 @8x16 +res", 4096, 576, 96, True, ops.ACT_NONE),
    ("mb expand 136->816 @8x16", 4096, 136, 816, False, ops.ACT_RELU6),
    ("mb project 816->136 @8x16 +res", 4096, 816, 136, True, ops.ACT_NONE),
    ("mb expand 232->1392 @4x8", 1024, 232, 1392, False, ops.ACT_RELU6),
    ("mb project 1392->232 @4x8 +res", 1024, 1392, 232, True, ops.ACT_NONE),
    ("mb project 1392->384 @4x8", 1024, 1392, 384, False, ops.ACT_NONE),
    ("probe K=64 N=128", 262144, 64, 128, False, ops.ACT_RELU6),
    ("probe K=24 N=128", 262144, 24, 128, False, ops.ACT_RELU6),
    ("probe K=64 N=144", 262144, 64, 144, False, ops.ACT_RELU6),
    ("probe K=64 N=256", 262144, 64, 256, False, ops.ACT_RELU6),
    ("probe K=64 N=64", 262144, 64, 64, False, ops.ACT_RELU6),
    ("probe K=256 N=128", 262144, 256, 128, False, ops.ACT_RELU6),
    ("probe K=1024 N=128", 262144, 1024, 128, False, ops.ACT_RELU6),
    ("vit-b qkv B=1", 577, 768, 2304, False, ops.ACT_NONE),
    ("vit-b fc2 B=1", 577, 3072, 768, False, ops.ACT_NONE),
]


# 3x3 convolutions of the DPT decoder at 32 frames: (name, B, H, W, C, residuals, relu copy)
CONVS = [
    ("rcu conv1 96x96 (store epilogue)", 32, 96, 96, 256, 0, False),
    ("same, forced through the general epilogue", 32, 96, 96, 256, -1, False),
    ("layer_rn style: raw + relu copy, no residual", 32, 96, 96, 256, 0, True),
    ("rcu conv2 96x96 +x +path, relu copy", 32, 96, 96, 256, 2, True),
    ("rcu conv2 96x96 +x", 32, 96, 96, 256, 1, False),
    ("rcu conv2 48x48 +x +path, relu copy", 32, 48, 48, 256, 2, True),
    ("rcu conv2 24x24 +x +path, relu copy", 32, 24, 24, 256, 2, True),
]


def time_it(run, n=20):
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        run()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n


def convs(dev):
    print("| conv | us/launch | TFLOP/s |")
    print("|---|---:|---:|")
    for name, B, H, W, C, nres, relu in CONVS:
        g = torch.Generator().manual_seed(2)
        x = (torch.randn((B, H, W, C), generator=g) * 0.5).to(dev).bfloat16()
        w = (torch.randn((C, 9 * C), generator=g) * 0.02).to(dev).bfloat16()
        b = torch.randn((C,), generator=g).to(dev)
        r1 = torch.randn((B, H, W, C), generator=g).to(dev).bfloat16() if nres >= 1 else None
        r2 = torch.randn((B, H, W, C), generator=g).to(dev).bfloat16() if nres >= 2 else None
        out = torch.empty((B, H, W, C), dtype=torch.bfloat16, device=dev)
        orl = torch.empty_like(out) if relu else None

        def run():
            ops.gemm(x, w, out, N=C, conv=(B, H, W, C), bias=b, act=ops.ACT_NONE if nres > 0 else ops.ACT_RELU,
                     add1=r1, ld_add1=C, add2=r2, out_relu=orl, ldo=C, epilogue=1 if nres < 0 else 0)
        t = time_it(run)
        print(f"| {name} | {t:.1f} | {2.0 * B * H * W * C * 9 * C / t / 1e6:.0f} |", flush=True)


def main():
    label = sys.argv[1] if len(sys.argv) > 1 else ""
    dev = torch.device("cuda:0")
    ws = torch.empty(64 << 20, dtype=torch.uint8, device=dev)
    print(f"# {label}")
    convs(dev)
    if label == "convonly":
        return
    print(f"# {label}")
    print("| shape | M | K | N | us/launch (stream) | us/launch (graph, PDL) |")
    print("|---|---:|---:|---:|---:|---:|")
    for name, M, K, N, res, act in SHAPES:
        g = torch.Generator().manual_seed(1)
        a = (torch.randn((M, K), generator=g) * 0.5).to(dev).bfloat16()
        w = (torch.randn((N, K), generator=g) * 0.1).to(dev).bfloat16()
        b = torch.randn((N,), generator=g).to(dev)
        r = (torch.randn((M, N), generator=g)).to(dev).bfloat16() if res else None
        out = torch.empty((M, N), dtype=torch.bfloat16, device=dev)

        def run():
            ops.gemm(a, w, out, bias=b, act=act, add1=r, splitk_ws=ws)
        for _ in range(3):
            run()
        torch.cuda.synchronize()
        ref = torch.nn.functional.linear(a.float(), w.float(), b)
        if act == ops.ACT_RELU6:
            ref = ref.clamp(0, 6)
        if res:
            ref = ref + r.float()
        err = float((out.float() - ref).abs().max() / ref.abs().max())
        assert err < 1e-2, (name, err)
        n = 50
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            run()
        e1.record()
        torch.cuda.synchronize()
        t_stream = e0.elapsed_time(e1) * 1e3 / n
        st = torch.cuda.Stream()
        with torch.cuda.stream(st):
            run()
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=st):
                for _ in range(n):
                    run()
            gr.replay()
            torch.cuda.synchronize()
            e0.record(st)
            gr.replay()
            e1.record(st)
        torch.cuda.synchronize()
        t_graph = e0.elapsed_time(e1) * 1e3 / n
        print(f"| {name} | {M} | {K} | {N} | {t_stream:.1f} | {t_graph:.1f} |", flush=True)


if __name__ == "__main__":
    main()
