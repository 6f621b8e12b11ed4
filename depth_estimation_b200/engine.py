"""ctypes shim over the handle-level C ABI (include/depth_b200.h: de_create ... de_destroy) - the C++ engine that owns
the packed weights, the per-geometry work buffers and the launch sequences of the three hub models.  This object is
what `DepthEstimator.estimator` holds in place of the reference's hub `nn.Module` (/root/reference/depth_midas.py:10)."""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from ._lib import DE_DTYPE_BF16, DE_DTYPE_F32, DE_PRECISION_BF16, DE_PRECISION_BF16X3, DE_PRECISION_DEFAULT, DePredictIO


class Engine:
    def __init__(self, model_type: str, state_dict: dict, device: torch.device, precision=None):
        """state_dict: upstream-format tensors (any float dtype / device); copied, folded and repacked by the library."""
        lib = _lib.load()
        self._lib = lib
        self.model_type = model_type
        self.device = device
        self._h = C.c_void_p()
        prec = {None: DE_PRECISION_DEFAULT, "bf16": DE_PRECISION_BF16, "bf16x3": DE_PRECISION_BF16X3}[precision]
        idx = device.index if device.index is not None else torch.cuda.current_device()
        _lib.check(lib.de_create(model_type.encode(), int(idx), prec, C.byref(self._h)), "de_create")
        try:
            for k, v in state_dict.items():
                t = v.detach()
                if t.dtype == torch.bfloat16:
                    t, dt = t.to("cpu").contiguous(), DE_DTYPE_BF16
                    ptr = t.view(torch.int16).data_ptr()
                else:
                    t, dt = t.to("cpu", torch.float32).contiguous(), DE_DTYPE_F32
                    ptr = t.data_ptr()
                shape = (C.c_int64 * max(1, t.dim()))(*t.shape)
                _lib.check(lib.de_set_weight(self._h, k.encode(), ptr, shape, t.dim(), dt), f"de_set_weight({k})")
            _lib.check(lib.de_finalize_weights(self._h), "de_finalize_weights")
        except Exception:
            lib.de_destroy(self._h)
            self._h = None
            raise
        lay, wb, kb = C.c_int(), C.c_int64(), C.c_int64()
        _lib.check(lib.de_engine_info(self._h, C.byref(lay), C.byref(wb), C.byref(kb)), "de_engine_info")
        self.preprocess_layout = lay.value
        self.weight_bytes = wb.value

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                self._lib.de_destroy(h)
            except Exception:
                pass

    def plan(self, B: int, H: int, W: int):
        """Work buffers + eager warm-up for B frames of H x W (synchronous).  Returns (net_h, net_w)."""
        nh, nw = C.c_int(), C.c_int()
        _lib.check(self._lib.de_plan(self._h, B, H, W, C.byref(nh), C.byref(nw)), "de_plan")
        return nh.value, nw.value

    def predict(self, B: int, H: int, W: int, frames: torch.Tensor, *, uv: torch.Tensor | None = None,
                out_map: torch.Tensor | None = None, out_metric: torch.Tensor | None = None,
                metric_scale: float = 1.0 / 2.54, metric_shift: float = 0.0, out_vis: torch.Tensor | None = None,
                sample_rc: torch.Tensor | None = None, out_samples: torch.Tensor | None = None, use_graph: bool = False):
        """K1 -> network -> K8 on the current torch stream.  All tensors live on the engine's device."""
        for t in (frames, uv, out_map, out_metric, out_vis, sample_rc, out_samples):
            if t is not None and not t.is_cuda:
                raise RuntimeError("Engine.predict takes CUDA tensors only (no CPU fallback)")
        io = DePredictIO()
        io.frames = frames.data_ptr()
        if uv is None:
            assert frames.dtype == torch.uint8 and frames.dim() == 4 and frames.stride(3) == 1 and frames.stride(2) == 3
            io.row_stride = frames.stride(1)
            io.frame_stride = frames.stride(0) if B > 1 else frames.stride(1) * H
        else:
            assert frames.dtype == torch.uint8 and frames.dim() == 3 and frames.stride(2) == 1
            assert uv.dtype == torch.uint8 and tuple(uv.shape) == (B, H // 2, W // 2, 2) and uv.stride(3) == 1 and uv.stride(2) == 2
            io.row_stride = frames.stride(1)
            io.frame_stride = frames.stride(0) if B > 1 else frames.stride(1) * H
            io.uv = uv.data_ptr()
            io.uv_row_stride = uv.stride(1)
            io.uv_frame_stride = uv.stride(0) if B > 1 else uv.stride(1) * (H // 2)
        if out_map is not None:
            assert out_map.dtype == torch.float32 and out_map.is_contiguous() and tuple(out_map.shape) == (B, H, W)
            io.out_map = out_map.data_ptr()
        if out_metric is not None:
            assert out_metric.dtype == torch.bfloat16 and out_metric.is_contiguous() and tuple(out_metric.shape) == (B, H, W)
            io.out_metric = out_metric.data_ptr()
        io.metric_scale, io.metric_shift = float(metric_scale), float(metric_shift)
        if out_vis is not None:
            assert out_vis.dtype == torch.uint8 and out_vis.is_contiguous() and tuple(out_vis.shape) == (B, H, W)
            io.out_vis = out_vis.data_ptr()
        if sample_rc is not None:
            assert sample_rc.dtype == torch.int32 and sample_rc.is_contiguous() and sample_rc.shape[0] == B
            assert out_samples is not None and out_samples.dtype == torch.float32 and out_samples.is_contiguous()
            io.sample_rc = sample_rc.data_ptr()
            io.n_samples = int(sample_rc.shape[1])
            io.out_samples = out_samples.data_ptr()
        _lib.check(self._lib.de_predict(self._h, B, H, W, C.byref(io), 1 if use_graph else 0,
                                        torch.cuda.current_stream().cuda_stream), "de_predict")
