"""Bandwidth benchmark of the fused pre-/post-processing kernels (K1 / K8) on working sets far larger than the
126 MB L2 (algorithmic bytes K1 = 3*H*W + 2*3*h*w, K8 = in + 4*H*W [+ 2*H*W]).  Prints one JSON
line per case: achieved GB/s against MEASURED_PEAKS.json's HBM copy bandwidth."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from depth_estimation_b200 import ops

dev = torch.device("cuda:0")
peak = 6453.0
try:
    with open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")) as f:
        peak = float(json.load(f)["hbm_gbs"])
except Exception:
    pass


def timeit(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e-3


def k1(B, H, W, mt, layout, name):
    nw, nh = ops.get_size(W, H, mt)
    frames = torch.randint(0, 256, (B, H, W, 3), dtype=torch.uint8, device=dev)
    mean, std = ((0.5,) * 3, (0.5,) * 3) if mt != "MiDaS_small" else ((0.485, 0.456, 0.406), (0.229, 0.224, 0.225))
    out = ops.preprocess(frames, nh, nw, mean, std, layout)
    t = timeit(lambda: ops.preprocess(frames, nh, nw, mean, std, layout, out=out))
    by = B * (3 * H * W + 2 * 3 * nh * nw)
    print(json.dumps({"kernel": "K1 preprocess", "case": name, "B": B, "src": [H, W], "net": [nh, nw], "layout": layout,
                      "us": t * 1e6, "algorithmic_MB": by / 1e6, "GBps": by / t / 1e9, "frac_of_hbm_peak": by / t / 1e9 / peak}),
          flush=True)


def k8(B, h, w, H, W, in_dtype, metric, name):
    pred = (torch.rand((B, h, w), device=dev) * 10).to(in_dtype)
    out_map = torch.empty((B, H, W), dtype=torch.float32, device=dev)
    out_metric = torch.empty((B, H, W), dtype=torch.bfloat16, device=dev) if metric else None
    t = timeit(lambda: ops.postprocess(pred, H, W, want_map=True, want_metric=metric, out_map=out_map, out_metric=out_metric))
    by = B * (pred.element_size() * h * w + 4 * H * W + (2 * H * W if metric else 0))
    print(json.dumps({"kernel": "K8 postprocess", "case": name, "B": B, "net": [h, w], "dst": [H, W],
                      "in_dtype": str(in_dtype), "metric_map": metric, "us": t * 1e6, "algorithmic_MB": by / 1e6,
                      "GBps": by / t / 1e9, "frac_of_hbm_peak": by / t / 1e9 / peak}), flush=True)


CASES = {
    "k1a": lambda: k1(64, 1080, 1920, "DPT_Large", ops.LAYOUT_PATCH16_BF16, "1080p -> 384x672 patch matrix (DPT, config #5)"),
    "k1b": lambda: k1(64, 1080, 1920, "DPT_Large", ops.LAYOUT_NCHW_BF16, "1080p -> 384x672 NCHW (DPT_Hybrid)"),
    "k1c": lambda: k1(128, 720, 1280, "MiDaS_small", ops.LAYOUT_NCHW_BF16, "720p -> 128x256 NCHW (MiDaS_small, config #4)"),
    "k1d": lambda: k1(512, 384, 384, "DPT_Large", ops.LAYOUT_PATCH16_BF16, "384^2 -> 384^2 patch matrix (configs #2/#3)"),
    "k8a": lambda: k8(64, 384, 672, 1080, 1920, torch.float32, False, "384x672 -> 1080p fp32 map (config #5)"),
    "k8b": lambda: k8(64, 384, 672, 1080, 1920, torch.float32, True, "384x672 -> 1080p fp32 map + bf16 metric map"),
    "k8c": lambda: k8(128, 128, 256, 720, 1280, torch.float32, False, "128x256 -> 720p fp32 map (config #4)"),
    "k8d": lambda: k8(512, 384, 384, 384, 384, torch.float32, False, "384^2 -> 384^2 (configs #2/#3)"),
}

if __name__ == "__main__":
    only = sys.argv[1:] or list(CASES)
    for name in only:
        CASES[name]()
