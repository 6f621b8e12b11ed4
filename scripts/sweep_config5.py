"""BASELINE config #5: DPT_Large, 1920x1080 frames -> net 384x672 -> full-resolution bicubic output, batch sweep.
One process per GPU (python / torchrun); every rank runs the same per-GPU batch (weak scaling), times are the max over
ranks, frames/s is the whole-job aggregate.  Prints one JSON line per batch size on rank 0.

    python scripts/sweep_config5.py --batches 1,2,4,8,16,32,64,128,256
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \
        scripts/sweep_config5.py --batches 1,32,256
"""
import argparse
import contextlib
import io
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

ap = argparse.ArgumentParser()
ap.add_argument("--batches", default="1,2,4,8,16,32,64,128,256")
ap.add_argument("--model", default="DPT_Large")
ap.add_argument("--width", type=int, default=1920)
ap.add_argument("--height", type=int, default=1080)
ap.add_argument("--max-gb", type=float, default=130.0, help="skip batch sizes whose planned buffers would exceed this")
args = ap.parse_args()
rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
from depth_estimation_b200.depth_midas import DepthEstimator

with contextlib.redirect_stdout(io.StringIO()):
    est = DepthEstimator(args.model, seed=0)
H, W = args.height, args.width
rng = np.random.default_rng(20245 + rank)


def sync():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def maxr(x):
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


per_frame_gb = None
for B in [int(b) for b in args.batches.split(",")]:
    free, total = torch.cuda.mem_get_info()
    if per_frame_gb is not None and per_frame_gb * B > min(args.max_gb, free / 2**30 * 0.85):
        if rank == 0:
            print(json.dumps({"batch": B, "skipped": f"would need ~{per_frame_gb * B:.0f} GB of work buffers"}), flush=True)
        continue
    frames = rng.integers(0, 256, (B, H, W, 3), dtype=np.uint8)
    d_frames = torch.from_numpy(frames).to(dev)
    f0 = torch.cuda.mem_get_info()[0]
    steps = max(3, min(20, 256 // B))
    for _ in range(3):
        est.predict_device(d_frames)
    sync()
    used = (f0 - torch.cuda.mem_get_info()[0]) / 2**30
    per_frame_gb = max(per_frame_gb or 0.0, used / B)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        est.predict_device(d_frames)
    e1.record()
    sync()
    ms = maxr(e0.elapsed_time(e1)) / steps
    # end to end: host frames in, host maps out, pipelined
    list(est.predict_stream([frames] * 2))
    sync()
    t0 = time.perf_counter()
    n = 0
    for out in est.predict_stream(frames for _ in range(steps)):
        n += 1
    torch.cuda.synchronize()
    dt = maxr(time.perf_counter() - t0) / steps
    # blocking single call (the reference's synchronous shape); p50 over the steps
    ts = []
    for _ in range(max(steps, 5)):
        t1 = time.perf_counter()
        est.predict_batch(frames)
        ts.append(time.perf_counter() - t1)
    p50 = maxr(float(np.median(ts)))
    if rank == 0:
        print(json.dumps({"model": args.model, "frame": f"{W}x{H}", "n_gpus": world, "batch_per_gpu": B,
                          "device_fps": world * B / (ms * 1e-3), "device_ms_per_step": ms,
                          "e2e_fps": world * B / dt, "e2e_ms_per_step": dt * 1e3,
                          "blocking_predict_batch_p50_ms": p50 * 1e3, "work_gb": round(used, 2)}), flush=True)
    del d_frames, frames
    est._plans.clear()
    est._pipes.clear()
    torch.cuda.empty_cache()
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
