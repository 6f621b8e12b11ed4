"""Drop-in for the reference's ``depth_midas.py`` / ``midas.py`` (byte-identical files): the same
``DepthEstimator(model_type)`` / ``.predict(img) -> np.float32[H,W]`` surface
(/root/reference/depth_midas.py:6-35), backed by the hand-written sm_100a kernels of libdepthb200.so.

What changes underneath (and nothing else): cvtColor + the MiDaS transform chain (:22-23) is one fused
kernel (K1), the hub model forward (:25-26) is the tcgen05 GEMM / implicit-conv / attention launch
sequence of the C++ engine (csrc/engine_nets.cu, reached through de_predict), and the bicubic resize back to the
frame size (:28-33) is K8.  There
is no CPU fallback: without a CUDA device or the built library the constructor raises.
"""
from __future__ import annotations

import os
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import torch

from . import ops
from .engine import Engine
from .init import random_state_dict

MODEL_TYPES = ("DPT_Large", "DPT_Hybrid", "MiDaS_small")
_DPT_NORM = ((0.5, 0.5, 0.5), (0.5, 0.5, 0.5))
_IMAGENET_NORM = ((0.485, 0.456, 0.406), (0.229, 0.224, 0.225))


# [UPSTREAM] hubconf.transforms(): name -> (target, Resize method for de_get_size, normalisation)
_TRANSFORM_SPECS = {"default_transform": (384, 0, _IMAGENET_NORM), "small_transform": (256, 0, _IMAGENET_NORM),
                    "dpt_transform": (384, 1, _DPT_NORM)}


class _Transforms:
    """Stand-in for the hub ``transforms`` namespace (depth_midas.py:12-16).  Each callable takes the RGB frame the
    reference passes (:22-23), runs K1 and returns what the hub transform returns - a float32 [1,3,h,w] tensor - except
    that it already lives on the estimator's device (the reference's `.to(self.device)` at :23 is then a no-op)."""

    def __init__(self, owner):
        self._o = owner
        self.dpt_transform = lambda rgb: owner._transform_rgb(rgb, "dpt_transform")
        self.small_transform = lambda rgb: owner._transform_rgb(rgb, "small_transform")
        self.default_transform = lambda rgb: owner._transform_rgb(rgb, "default_transform")


class DepthEstimator(object):
    def __init__(self, model_type, *, weights=None, seed=0, device=None, use_cuda_graph=True, precision=None):
        """model_type: "DPT_Large" | "DPT_Hybrid" | "MiDaS_small" (anything else gets the small transform,
        as at depth_midas.py:13-16, and is rejected by the model factory).
        weights: upstream-format state_dict, a path to one (torch.load), or None for seeded random init.
        precision: None = the model's default arithmetic (bf16 operands / fp32 accumulate; DPT_Hybrid: ResNetV2 stages
        and decoder as bf16x3); "bf16" = plain bf16 operands wherever the model tolerates them (DPT_Hybrid: decoder in
        bf16 - faster, 1.3e-3 from the fp32 reference instead of < 1e-3); "bf16x3" (DPT_Large / DPT_Hybrid) = every
        Linear / convolution on hi+lo bf16 operand pairs (~16 mantissa bits, 3x the tensor work) for maps whose dynamic
        range is wide against their maximum, where bf16 operands alone cost more than north_star's 1e-3."""
        if precision not in (None, "bf16", "bf16x3"):
            raise ValueError(f"precision {precision!r}: expected None, 'bf16' or 'bf16x3'")
        self.precision = precision
        self.model_type = model_type
        print(f'Loading model: {self.model_type}')
        if not torch.cuda.is_available():
            raise RuntimeError("depth_estimation_b200 needs a CUDA (sm_100a) device; there is no CPU path")
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        ops._lib.load()
        if weights is None:
            weights = random_state_dict(model_type, seed)
        else:
            from .weights import load_checkpoint
            weights = load_checkpoint(weights, model_type)      # path or mapping in upstream naming, validated
        if model_type not in MODEL_TYPES:
            raise ValueError(f"unknown model_type {model_type!r}; expected one of {MODEL_TYPES}")
        with torch.cuda.device(self.device):
            # the C++ engine behind the handle-level C ABI (de_create .. de_predict): packed weights, work buffers,
            # launch sequences of the hub model - what `torch.hub.load` returns at depth_midas.py:10
            self.estimator = Engine(model_type, weights, self.device, precision)
        print('Loading transforms...')
        self.midas_transforms = _Transforms(self)
        if self.model_type == "DPT_Large" or self.model_type == "DPT_Hybrid":
            self.transform = self.midas_transforms.dpt_transform
            self._norm = _DPT_NORM
        else:
            self.transform = self.midas_transforms.small_transform
            self._norm = _IMAGENET_NORM
        self.use_cuda_graph = use_cuda_graph
        self._plans = {}
        self._pipes = {}
        self._replayed_launches = 0
        self.metric_affine = (1.0 / 2.54, 0.0)                  # K8's metric map = map*scale + shift (face.py:137)

    # ------------------------------------------------------------------ reference surface
    def predict(self, img):
        """img: np.uint8 [H,W,3] BGR (cv2.VideoCapture frame).  Returns np.float32 [H,W] relative inverse
        depth at source resolution, host memory, synchronous - as depth_midas.py:21-35."""
        if not isinstance(img, np.ndarray) or img.ndim != 3 or img.shape[2] != 3 or img.dtype != np.uint8:
            raise TypeError("predict expects a uint8 HxWx3 BGR numpy frame")
        out = self.predict_batch(img[None])
        return out[0]

    # ------------------------------------------------------------------ supersets
    def predict_batch(self, frames, samples_rc=None, want_metric=False, want_vis=False):
        """frames: np.uint8 [B,H,W,3] BGR (host).  Returns np.float32 [B,H,W]; with samples_rc
        (int [B,S,2] (row,col)) returns (maps, samples[B,S]); with want_metric adds the bf16 metric map
        (face.py:137, D/2.54) as a float32 array; with want_vis adds the uint8 grey level of the reference's
        to_video_frame (midas_mediapipe.py:141-150: uint8(map)*5+10, wrapping) - all from the same K8 launch."""
        frames = np.ascontiguousarray(frames)
        B, H, W, _ = frames.shape
        plan = self._plan(B, H, W, 0 if samples_rc is None else int(np.asarray(samples_rc).shape[1]), want_metric,
                          want_vis)
        plan.h_in.numpy()[...] = frames
        if samples_rc is not None:
            plan.h_rc.numpy()[...] = np.asarray(samples_rc, dtype=np.int32)
        with torch.cuda.device(self.device):
            plan.run()
            torch.cuda.current_stream().synchronize()
        maps = plan.h_map.numpy().copy()
        res = [maps]
        if samples_rc is not None:
            res.append(plan.h_samples.numpy().copy())
        if want_metric:
            res.append(plan.h_metric.float().numpy().copy())
        if want_vis:
            res.append(plan.h_vis.numpy().copy())
        return res[0] if len(res) == 1 else tuple(res)

    # ------------------------------------------------------------------ SURVEY 8f N4
    def set_metric_affine(self, scale: float = 1.0 / 2.54, shift: float = 0.0):
        """Metric map K8 emits with want_metric: map*scale + shift.  (1/2.54, 0) is `FaceDet.rel2abs` (face.py:137)."""
        if (float(scale), float(shift)) != self.metric_affine:
            self.metric_affine = (float(scale), float(shift))
            self._plans.clear()                                 # the scalars are baked into the captured graphs
            self._pipes.clear()

    def calibrate_metric(self, pred_depths, gt_depths_cm):
        """Least-squares rel -> abs calibration of `FaceDet.rel2abs_2` (face.py:118-130): fits pred = m/gt + b, then has
        K8 emit cm_to_ft(map*m + b) = map*(m/30.48) + b/30.48 as its metric map.  Returns (m, b)."""
        gt = [1 / x for x in gt_depths_cm]
        A = np.vstack([gt, np.ones(len(gt))]).T
        m, b = np.linalg.lstsq(A, np.asarray(pred_depths, dtype=np.float64), rcond=None)[0]
        self.set_metric_affine(m / 30.48, b / 30.48)
        return float(m), float(b)

    def predict_nv12(self, nv12):
        """Frames as a hardware decoder delivers them: uint8 [H*3/2, W] (or [B, H*3/2, W]) NV12 - Y plane followed by
        the interleaved half-resolution UV plane, the array cv2.cvtColor(.., COLOR_YUV2BGR_NV12) takes.  The colour
        conversion is fused into K1; the result equals predict(cv2.cvtColor(nv12, COLOR_YUV2BGR_NV12)) bit for bit."""
        a = np.ascontiguousarray(nv12)
        single = a.ndim == 2
        if single:
            a = a[None]
        if a.dtype != np.uint8 or a.ndim != 3 or a.shape[1] % 3 != 0 or a.shape[2] % 2 != 0:
            raise TypeError("predict_nv12 expects uint8 [H*3/2, W] NV12 frames with even H and W")
        B, H, W = a.shape[0], a.shape[1] * 2 // 3, a.shape[2]
        with torch.cuda.device(self.device):
            t = torch.from_numpy(a).to(self.device)
            y = t[:, :H]
            uv = t[:, H:].reshape(B, H // 2, W // 2, 2)
            out = torch.empty((B, H, W), dtype=torch.float32, device=self.device)
            self.estimator.predict(B, H, W, y, uv=uv, out_map=out)
            res = out.cpu().numpy()
        return res[0] if single else res

    def predict_device(self, frames_u8: torch.Tensor, eager: bool = False) -> torch.Tensor:
        """Device-resident path: uint8 [B,H,W,3] BGR CUDA tensor -> fp32 [B,H,W] CUDA tensor (a buffer
        owned by the estimator, overwritten by the next call).  No host copies, no sync.  eager=True launches
        the kernels one by one instead of replaying the captured graph (profiling / per-kernel timing)."""
        B, H, W, _ = frames_u8.shape
        plan = self._plan(B, H, W, 0, False)
        with torch.cuda.device(self.device):
            return plan.run_device(frames_u8, eager)

    def kernel_launches(self) -> int:
        """Kernels of libdepthb200.so this process has put on the GPU so far: direct launches (counted by the
        library) plus, for every CUDA-graph replay, the launches that graph captured."""
        return int(ops._lib.launch_count()) + self._replayed_launches

    def predict_stream(self, batches, depth=2):
        """Pipelined form of predict_batch for a stream of equally-shaped uint8 [B,H,W,3] BGR host batches
        (the headless `VidStream.stream` loop of midas_mediapipe.py:32-129 feeds this): a generator that yields
        one np.float32 [B,H,W] result per input batch, in order, keeping up to `depth` batches in flight.

        Per slot: pinned staging buffers + a captured graph of K1 -> net -> K8.  H2D copies run on their own
        stream, compute on a second, D2H on a third, chained by events, so the host-side memcpys and the PCIe
        transfers of batch i+1 / i-1 overlap the kernels of batch i.  Every frame still crosses the host
        boundary in both directions; only the serialisation point of depth_midas.py:23,34 is gone."""
        for res in self.predict_stream_fused(((b, None) for b in batches), depth=depth, n_samples=0, want_vis=False,
                                             want_map=True):
            yield res.maps

    def predict_stream_fused(self, items, depth=2, n_samples=2, want_vis=True, want_map=False):
        """The pipelined loop with K8's fused outputs instead of the full-resolution fp32 map: `items` yields
        (frames, rc) - frames uint8 [B,H,W,3] (or a sequence of B [H,W,3] frames), rc int [B,n_samples,2] (row, col)
        points into the output map (`FaceDet.sample_indices`, face.py:94-107) - and a StreamResult comes back per
        item, in order: the sampled map values (what `FaceDet.get_depth` reads), the uint8 grey level of
        `to_video_frame` (midas_mediapipe.py:141-150) and, only with want_map, the map itself.  Without it a 720p
        frame costs 0.9 MB of device->host traffic instead of 3.7 MB and no host pass over the map at all."""
        depth = max(1, int(depth))
        pending = []                                  # (slot, future -> the slot's D2H-done event)
        pipe = None
        idx = 0
        last = None
        it = iter(items)
        with torch.cuda.device(self.device):
            while True:
                if last is not None:
                    last.result()                     # previous batch staged and enqueued: its source arrays are free again
                try:
                    frames, rc = next(it)
                except StopIteration:
                    break
                if isinstance(frames, np.ndarray):
                    frames = np.ascontiguousarray(frames)
                    shape = frames.shape
                    ok = frames.ndim == 4 and frames.shape[3] == 3 and frames.dtype == np.uint8
                else:
                    frames = list(frames)
                    ok = len(frames) > 0 and all(isinstance(f, np.ndarray) and f.dtype == np.uint8 and f.ndim == 3
                                                 and f.shape == frames[0].shape and f.shape[2] == 3 for f in frames)
                    shape = (len(frames),) + tuple(frames[0].shape) if ok else None
                if not ok:
                    raise TypeError("predict_stream expects uint8 [B,H,W,3] BGR batches")
                if n_samples:
                    rc = np.asarray(rc, dtype=np.int32)
                    if rc.shape != (shape[0], n_samples, 2):
                        raise ValueError(f"sample points must be [B,{n_samples},2] (row, col), got {rc.shape}")
                key = (tuple(shape), depth, n_samples, want_vis, want_map)
                if pipe is None or pipe.key != key:
                    while pending:                    # geometry change: drain, then rebuild the slots
                        yield pipe.collect(*pending.pop(0))
                    pipe = self._pipes.get(key)
                    if pipe is None:
                        pipe = _Pipeline(self, shape, depth, n_samples, want_vis, want_map)
                        pipe.key = key
                        self._pipes[key] = pipe
                    idx = 0
                # depth + 1 slots: the batch collected below (idx - depth) and the one being staged (idx) never share
                # buffers, so the host copy into pinned memory (worker thread) overlaps the copy out of it (this thread)
                slot = idx % pipe.nslots
                last = pipe.submit_async(slot, frames, rc if n_samples else None)
                pending.append((slot, last))
                idx += 1
                if len(pending) > depth:
                    yield pipe.collect(*pending.pop(0))
            while pending:
                yield pipe.collect(*pending.pop(0))

    # ------------------------------------------------------------------ internals
    def _transform_rgb(self, rgb, name):
        """Reference-style transform(img): takes the RGB frame the reference passes (:22-23)."""
        bgr = np.ascontiguousarray(rgb[:, :, ::-1])
        H, W = bgr.shape[:2]
        target, method, (mean, std) = _TRANSFORM_SPECS[name]
        nw, nh = ops.get_size_for(W, H, target, method)
        with torch.cuda.device(self.device):
            d = torch.from_numpy(bgr[None]).to(self.device)
            return ops.preprocess(d, nh, nw, mean, std, ops.LAYOUT_NCHW_F32)

    def _plan(self, B, H, W, n_samples, want_metric, want_vis=False):
        key = (B, H, W, n_samples, want_metric, want_vis)
        p = self._plans.get(key)
        if p is None:
            with torch.cuda.device(self.device):
                p = _Plan(self, B, H, W, n_samples, want_metric, want_vis)
            self._plans[key] = p
        return p


class _Plan:
    """Static buffers + (optionally) a captured CUDA graph for one (B,H,W) geometry:
    pinned host frame -> H2D -> K1 -> net -> K8 -> D2H pinned."""

    def __init__(self, est: DepthEstimator, B, H, W, n_samples, want_metric, want_vis=False, host_map=True):
        self.est = est
        dev = est.device
        self.B, self.H, self.W = B, H, W
        self.net_w, self.net_h = ops.get_size(W, H, est.model_type)
        self.h_in = torch.empty((B, H, W, 3), dtype=torch.uint8).pin_memory()
        self.d_in = torch.empty((B, H, W, 3), dtype=torch.uint8, device=dev)
        self.d_map = torch.empty((B, H, W), dtype=torch.float32, device=dev)
        self.h_map = torch.empty((B, H, W), dtype=torch.float32).pin_memory() if host_map else None
        self.want_metric = want_metric
        self.d_metric = torch.empty((B, H, W), dtype=torch.bfloat16, device=dev) if want_metric else None
        self.h_metric = torch.empty((B, H, W), dtype=torch.bfloat16).pin_memory() if want_metric else None
        self.want_vis = want_vis
        self.d_vis = torch.empty((B, H, W), dtype=torch.uint8, device=dev) if want_vis else None
        self.h_vis = torch.empty((B, H, W), dtype=torch.uint8).pin_memory() if want_vis else None
        self.n_samples = n_samples
        if n_samples:
            self.h_rc = torch.zeros((B, n_samples, 2), dtype=torch.int32).pin_memory()
            self.d_rc = torch.zeros((B, n_samples, 2), dtype=torch.int32, device=dev)
            self.h_samples = torch.empty((B, n_samples), dtype=torch.float32).pin_memory()
        self.graph = None
        self._dev_graph = None
        self._warm = False
        self._d_samples = torch.empty((B, n_samples), dtype=torch.float32, device=dev) if n_samples else None
        est.estimator.plan(B, H, W)                             # work buffers + warm-up now, not inside a capture

    def _compute(self, src):
        """K1 -> network -> K8 on the current stream: one call into the engine (de_predict)."""
        est = self.est
        samples = self._d_samples
        est.estimator.predict(self.B, self.H, self.W, src, out_map=self.d_map, out_metric=self.d_metric,
                              metric_scale=est.metric_affine[0], metric_shift=est.metric_affine[1], out_vis=self.d_vis,
                              sample_rc=self.d_rc if self.n_samples else None, out_samples=samples)
        return samples

    def _body(self):
        self.d_in.copy_(self.h_in, non_blocking=True)
        if self.n_samples:
            self.d_rc.copy_(self.h_rc, non_blocking=True)
        samples = self._compute(self.d_in)
        if self.h_map is not None:
            self.h_map.copy_(self.d_map, non_blocking=True)
        if self.want_metric:
            self.h_metric.copy_(self.d_metric, non_blocking=True)
        if self.want_vis:
            self.h_vis.copy_(self.d_vis, non_blocking=True)
        if self.n_samples:
            self.h_samples.copy_(samples, non_blocking=True)

    def run(self):
        if not self.est.use_cuda_graph:
            self._body()
            return
        if self.graph is None:
            # warm-up on a side stream (allocates work buffers / tap tables), then capture
            s = torch.cuda.Stream()
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                self._body()
            torch.cuda.current_stream().wait_stream(s)
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            n0 = ops._lib.launch_count()
            with torch.cuda.graph(g):
                self._body()
            self._graph_launches = int(ops._lib.launch_count() - n0)
            self.graph = g
        self.graph.replay()
        self.est._replayed_launches += self._graph_launches

    def run_device(self, frames_u8, eager=False):
        """Device-resident step.  With CUDA graphs on, the frames are copied (device to device, ~2 us per MB) into
        the plan's static input buffer and the captured K1 -> net -> K8 graph is replayed: ~230 launches become
        one, which removes the inter-kernel launch gaps (measured 0.4 ms of a 21 ms DPT_Large batch-32 step)."""
        if eager or not self.est.use_cuda_graph:
            self._compute(frames_u8)
            return self.d_map
        if self._dev_graph is None:
            s = torch.cuda.Stream()
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                self._compute(self.d_in)                        # allocates work buffers / tap tables
            torch.cuda.current_stream().wait_stream(s)
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            n0 = ops._lib.launch_count()
            with torch.cuda.graph(g):
                self._compute(self.d_in)
            self._dev_graph_launches = int(ops._lib.launch_count() - n0)
            self._dev_graph = g
        if frames_u8.data_ptr() != self.d_in.data_ptr():
            self.d_in.copy_(frames_u8, non_blocking=True)
        self._dev_graph.replay()
        self.est._replayed_launches += self._dev_graph_launches
        return self.d_map


class StreamResult:
    """One batch out of the pipelined loop.  `maps` (float32 [B,H,W], a copy) is present unless the stream was opened
    with want_map=False; `vis` is the uint8 grey level of the reference's to_video_frame (a view into pinned memory,
    valid until the next result is taken); `samples` float32 [B,S] are the map values at the requested (row, col)
    points.  map_pixel reads single elements of the device-resident map (a device sync; the rare body-pose branch of
    midas_mediapipe.py:95-117 needs two pixels whose position is not known when the frame is submitted)."""

    def __init__(self, plan, maps, vis, samples):
        self._plan, self.maps, self.vis, self.samples = plan, maps, vis, samples

    def map_pixel(self, b, r, c):
        return float(self._plan.d_map[b, r, c].item())


class _Pipeline:
    """`depth` in-flight slots for predict_stream: per slot a pinned input / output pair, device input / output
    buffers and a CUDA graph of the compute; three streams (H2D, compute, D2H) chained by events.  The network's
    work buffers are shared between slots - safe because all compute is ordered on the one compute stream."""

    def __init__(self, est: DepthEstimator, shape, depth, n_samples=0, want_vis=False, want_map=True):
        self.est, self.shape, self.depth = est, tuple(shape), depth
        self.n_samples, self.want_vis, self.want_map = n_samples, want_vis, want_map
        B, H, W, _ = shape
        dev = est.device
        self.nslots = n = depth + 1
        self.plans = [_Plan(est, B, H, W, n_samples, False, want_vis, host_map=want_map) for _ in range(n)]
        self.s_in, self.s_compute, self.s_out = (torch.cuda.Stream(dev) for _ in range(3))
        self.ev_in = [torch.cuda.Event() for _ in range(n)]
        self.ev_done = [torch.cuda.Event() for _ in range(n)]
        self.ev_out = [torch.cuda.Event() for _ in range(n)]
        self.graphs = [None] * n
        # one worker thread stages frames into pinned memory and enqueues the slot's work (numpy's memcpy releases the
        # GIL).  Worth it when the copies are long: MiDaS_small at 720p (88 MB in, 118 MB out per 32 frames) goes from
        # 990 to 1235 frames/s end to end; at 14 MB per batch (DPT_Large 384^2) the hand-off costs more than it hides,
        # so small batches are staged inline.  DEB_STAGE_THREAD=0 / 1 forces either.  Large batches are additionally
        # split over a few copy threads (a single memcpy stream moves ~10 GB/s; 720p frames at 5000 frames/s need 14).
        mode = os.environ.get("DEB_STAGE_THREAD")
        threaded = mode == "1" or (mode != "0" and B * H * W * 3 >= (32 << 20))
        self._pool = ThreadPoolExecutor(1, "deb-stage") if threaded else None
        ncopy = int(os.environ.get("DEB_COPY_THREADS", "0")) or min(4, max(1, (os.cpu_count() or 2) // 2))
        # the copy threads also serve inline staging: a 14-19 MB memcpy split four ways shortens the pipeline's fill and
        # drain (DPT_Large 384^2, 32 frames: first staging 1.4 ms, last copy-out 1.9 ms on one thread)
        self._copy_min = int(float(os.environ.get("DEB_COPY_MIN_MB", "8")) * (1 << 20))
        big = threaded or B * H * W * 3 >= self._copy_min
        self._copiers = ThreadPoolExecutor(ncopy, "deb-copy") if (big and ncopy > 1 and B > 1) else None
        self._ncopy = ncopy
        cur = torch.cuda.current_stream(dev)
        self.s_compute.wait_stream(cur)
        with torch.cuda.stream(self.s_compute):
            self.plans[0]._compute(self.plans[0].d_in)          # allocate work buffers / tap tables
        self.s_compute.synchronize()
        if est.use_cuda_graph:
            for i, p in enumerate(self.plans):
                g = torch.cuda.CUDAGraph()
                n0 = ops._lib.launch_count()
                with torch.cuda.graph(g, stream=self.s_compute):
                    p._compute(p.d_in)
                self.graph_launches = int(ops._lib.launch_count() - n0)
                self.graphs[i] = g
        for e in self.ev_done + self.ev_out:                    # so the first waits are no-ops
            e.record(self.s_compute)

    def submit_async(self, slot, frames, rc=None):
        """Future of submit(slot, frames) (already completed when the worker thread is disabled)."""
        if self._pool is not None:
            return self._pool.submit(self._submit_on_device, slot, frames, rc)
        return _Done(self.submit(slot, frames, rc))

    def _submit_on_device(self, slot, frames, rc):
        with torch.cuda.device(self.est.device):
            return self.submit(slot, frames, rc)

    def _stage(self, dst, frames):
        """frames: one [B,H,W,3] array or a sequence of B [H,W,3] frames -> the slot's pinned buffer."""
        B = dst.shape[0]
        if self._copiers is None:
            if isinstance(frames, np.ndarray):
                dst[...] = frames
            else:
                for i, f in enumerate(frames):
                    dst[i] = f
            return
        step = -(-B // self._ncopy)

        def part(lo):
            for i in range(lo, min(B, lo + step)):
                np.copyto(dst[i], frames[i])
        for f in [self._copiers.submit(part, lo) for lo in range(0, B, step)]:
            f.result()

    def submit(self, slot, frames, rc=None):
        p = self.plans[slot]
        self._stage(p.h_in.numpy(), frames)                     # slot is free: its previous result was collected
        if self.n_samples:
            p.h_rc.numpy()[...] = rc
        self.s_in.wait_event(self.ev_done[slot])                # previous compute on this slot has consumed d_in
        with torch.cuda.stream(self.s_in):
            p.d_in.copy_(p.h_in, non_blocking=True)
            if self.n_samples:
                p.d_rc.copy_(p.h_rc, non_blocking=True)
            self.ev_in[slot].record()
        self.s_compute.wait_event(self.ev_in[slot])
        self.s_compute.wait_event(self.ev_out[slot])            # previous D2H of this slot has drained its outputs
        with torch.cuda.stream(self.s_compute):
            if self.graphs[slot] is not None:
                self.graphs[slot].replay()
                self.est._replayed_launches += self.graph_launches
            else:
                p._compute(p.d_in)
            self.ev_done[slot].record()
        self.s_out.wait_event(self.ev_done[slot])
        with torch.cuda.stream(self.s_out):
            if self.want_map:
                p.h_map.copy_(p.d_map, non_blocking=True)
            if self.want_vis:
                p.h_vis.copy_(p.d_vis, non_blocking=True)
            if self.n_samples:
                p.h_samples.copy_(p._d_samples, non_blocking=True)
            self.ev_out[slot].record()
        return self.ev_out[slot]

    def _copy_out(self, src):
        """Pinned result -> a fresh array the caller owns; large maps are split over the copy threads."""
        if self._copiers is None or src.nbytes < self._copy_min:
            return src.copy()
        out = np.empty_like(src)
        B = src.shape[0]
        step = -(-B // self._ncopy)

        def part(lo):
            np.copyto(out[lo:lo + step], src[lo:lo + step])
        for f in [self._copiers.submit(part, lo) for lo in range(0, B, step)]:
            f.result()
        return out

    def collect(self, slot, fut):
        fut.result().synchronize()                              # enqueued (worker exceptions surface here), then D2H done
        p = self.plans[slot]
        return StreamResult(p, self._copy_out(p.h_map.numpy()) if self.want_map else None,
                            p.h_vis.numpy() if self.want_vis else None,
                            p.h_samples.numpy().copy() if self.n_samples else None)


class _Done:
    """A completed future (inline staging)."""

    def __init__(self, value):
        self._value = value

    def result(self):
        return self._value


class VidStream(object):
    """Headless counterpart of the reference wrapper (`depth_midas.py:38-85`): same constructor and method
    names, no HighGUI (`waitKey`/`destroyAllWindows` have no meaning on a compute box).  `src` is a video path
    (opened with cv2.VideoCapture when OpenCV is present) or any iterable of uint8 BGR frames; `output` is a
    path for an MJPG file (needs OpenCV) or None, in which case the 3-channel uint8 frames are kept in
    `self.outputs`."""

    def __init__(self, estimator, src=None, output=None):
        self.estimator = estimator
        self.FPS = 1 / 30
        self.FPS_MS = int(self.FPS * 1000)
        self.video = None
        self._frames = None
        self.w = self.h = 0
        if isinstance(src, str):
            import cv2
            self.video = cv2.VideoCapture(src)
            self.video.set(cv2.CAP_PROP_BUFFERSIZE, 2)
            self.w, self.h = int(self.video.get(3)), int(self.video.get(4))
        elif src is not None:
            self._frames = iter(src)
        self.output = output
        self.writer = None
        self.outputs = []
        self.status = None
        self.frame = None
        self.cnt = 0

    def _read(self):
        if self.video is not None:
            return self.video.read()
        try:
            return True, next(self._frames)
        except (StopIteration, TypeError):
            return False, None

    def update(self):
        while True:
            self.status, self.frame = self._read()
            if not self.status:
                break
            self.cnt += 1
            self.write_output()
        if self.video is not None:
            self.video.release()
        if self.writer is not None:
            self.writer.release()

    def write_output(self):
        prediction = self.estimator.predict(self.frame)
        with np.errstate(invalid="ignore"):
            output = prediction.astype(np.uint8)                # wraps, exactly like depth_midas.py:83
        three_c = np.stack([output, output, output], axis=-1)   # cv2.merge([o, o, o])
        if self.output is not None:
            import cv2
            if self.writer is None:
                h, w = output.shape
                self.writer = cv2.VideoWriter(str(self.output), cv2.VideoWriter_fourcc(*'MJPG'), 20, (w, h))
            self.writer.write(three_c)
        else:
            self.outputs.append(three_c)
