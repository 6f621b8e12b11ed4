"""Thin torch-tensor front-ends over the C ABI (include/depth_b200.h).  torch supplies device memory
and the current stream only; every op below is one call into libdepthb200.so and raises on failure."""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import ACT_GELU, ACT_NONE, ACT_RELU, ACT_RELU6, DE_DTYPE_BF16, DE_DTYPE_F32, DeGemm  # noqa: F401


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t):
    return None if t is None else t.data_ptr()


def _dt(t) -> int:
    if t.dtype == torch.bfloat16:
        return DE_DTYPE_BF16
    if t.dtype == torch.float32:
        return DE_DTYPE_F32
    raise TypeError(f"unsupported dtype {t.dtype}")


def _need_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError("depth_estimation_b200 ops run on CUDA tensors only (no CPU fallback)")


# ------------------------------------------------------------------------------- host helpers
def get_size_for(width: int, height: int, target: int, method: int):
    """[UPSTREAM] Resize.get_size: method 0 = "upper_bound", 1 = "minimal".  Returns (net_w, net_h)."""
    w, h = C.c_int(), C.c_int()
    _lib.check(_lib.load().de_get_size(width, height, target, method, C.byref(w), C.byref(h)), "de_get_size")
    return w.value, h.value


def get_size(width: int, height: int, model_type: str):
    """(net_w, net_h) the reference transform would resize to (depth_midas.py:13-16,23)."""
    dpt = model_type in ("DPT_Large", "DPT_Hybrid")
    return get_size_for(width, height, 384 if dpt else 256, 1 if dpt else 0)


def cubic_taps(src: int, dst: int, mode: int):
    idx = np.empty((dst, 4), dtype=np.int32)
    w = np.empty((dst, 4), dtype=np.float32)
    _lib.check(_lib.load().de_cubic_taps(src, dst, mode, idx.ctypes.data, w.ctypes.data), "de_cubic_taps")
    return idx, w


# ------------------------------------------------------------------------------- K1 / K8
LAYOUT_NCHW_BF16, LAYOUT_PATCH16_BF16, LAYOUT_NCHW_F32 = 0, 1, 2


def preprocess(frames_u8: torch.Tensor, net_h: int, net_w: int, mean, std, layout: int = LAYOUT_NCHW_BF16,
               out: torch.Tensor | None = None) -> torch.Tensor:
    """frames_u8: uint8 [B,H,W,3] BGR on the GPU (contiguous rows)."""
    _need_cuda(frames_u8)
    assert frames_u8.dtype == torch.uint8 and frames_u8.dim() == 4 and frames_u8.shape[3] == 3
    assert frames_u8.stride(3) == 1 and frames_u8.stride(2) == 3
    B, H, W, _ = frames_u8.shape
    if out is None:
        if layout == LAYOUT_NCHW_BF16:
            out = torch.empty((B, 3, net_h, net_w), dtype=torch.bfloat16, device=frames_u8.device)
        elif layout == LAYOUT_PATCH16_BF16:
            out = torch.empty((B * (net_h // 16) * (net_w // 16), 768), dtype=torch.bfloat16, device=frames_u8.device)
        else:
            out = torch.empty((B, 3, net_h, net_w), dtype=torch.float32, device=frames_u8.device)
    m = (C.c_float * 3)(*[float(v) for v in mean])
    s = (C.c_float * 3)(*[float(v) for v in std])
    fstride = frames_u8.stride(0) if B > 1 else frames_u8.stride(1) * H
    _lib.check(_lib.load().de_preprocess(frames_u8.data_ptr(), B, H, W, frames_u8.stride(1), fstride,
                                         net_h, net_w, m, s, layout, out.data_ptr(), _stream()), "de_preprocess")
    return out


def _alloc_net_input(B, net_h, net_w, layout, device):
    if layout == LAYOUT_NCHW_BF16:
        return torch.empty((B, 3, net_h, net_w), dtype=torch.bfloat16, device=device)
    if layout == LAYOUT_PATCH16_BF16:
        return torch.empty((B * (net_h // 16) * (net_w // 16), 768), dtype=torch.bfloat16, device=device)
    return torch.empty((B, 3, net_h, net_w), dtype=torch.float32, device=device)


def preprocess_nv12(y: torch.Tensor, uv: torch.Tensor, net_h: int, net_w: int, mean, std,
                    layout: int = LAYOUT_NCHW_BF16, out: torch.Tensor | None = None) -> torch.Tensor:
    """K1 on NV12 frames: y uint8 [B,H,W], uv uint8 [B,H/2,W/2,2] (interleaved U,V), both on the GPU.  Bit-exact with
    cv2.cvtColor(COLOR_YUV2BGR_NV12) followed by `preprocess`."""
    _need_cuda(y, uv)
    assert y.dtype == torch.uint8 and uv.dtype == torch.uint8 and y.dim() == 3 and uv.dim() == 4 and uv.shape[3] == 2
    B, H, W = y.shape
    assert tuple(uv.shape) == (B, H // 2, W // 2, 2) and y.stride(2) == 1 and uv.stride(3) == 1 and uv.stride(2) == 2
    if out is None:
        out = _alloc_net_input(B, net_h, net_w, layout, y.device)
    m = (C.c_float * 3)(*[float(v) for v in mean])
    s = (C.c_float * 3)(*[float(v) for v in std])
    _lib.check(_lib.load().de_preprocess_nv12(y.data_ptr(), uv.data_ptr(), B, H, W, y.stride(1),
                                              y.stride(0) if B > 1 else y.stride(1) * H, uv.stride(1),
                                              uv.stride(0) if B > 1 else uv.stride(1) * (H // 2), net_h, net_w, m, s,
                                              layout, out.data_ptr(), _stream()), "de_preprocess_nv12")
    return out


def postprocess(pred: torch.Tensor, H: int, W: int, want_map: bool = True, want_metric: bool = False,
                metric_scale: float = 1.0 / 2.54, metric_shift: float = 0.0, sample_rc: torch.Tensor | None = None,
                out_map: torch.Tensor | None = None, out_metric: torch.Tensor | None = None,
                out_vis: torch.Tensor | None = None):
    """pred: [B,h,w] f32/bf16.  Returns (map f32 [B,H,W] | None, metric bf16 | None, samples f32 [B,S] | None).
    out_vis: optional uint8 [B,H,W] receiving the reference's to_video_frame grey level (uint8(map)*5+10, wrapping)."""
    _need_cuda(pred, sample_rc)
    assert pred.dim() == 3 and pred.is_contiguous()
    B, h, w = pred.shape
    if want_map and out_map is None:
        out_map = torch.empty((B, H, W), dtype=torch.float32, device=pred.device)
    if want_metric and out_metric is None:
        out_metric = torch.empty((B, H, W), dtype=torch.bfloat16, device=pred.device)
    samples, ns = None, 0
    if sample_rc is not None:
        assert sample_rc.dtype == torch.int32 and sample_rc.is_contiguous() and sample_rc.shape[0] == B
        ns = sample_rc.shape[1]
        samples = torch.empty((B, ns), dtype=torch.float32, device=pred.device)
    _lib.check(_lib.load().de_postprocess(pred.data_ptr(), _dt(pred), B, h, w, H, W, _ptr(out_map), _ptr(out_metric),
                                          float(metric_scale), float(metric_shift), _ptr(out_vis), _ptr(sample_rc), ns, _ptr(samples), _stream()),
               "de_postprocess")
    return out_map, out_metric, samples


# ------------------------------------------------------------------------------- GEMM / conv
def gemm(a: torch.Tensor, w: torch.Tensor, out: torch.Tensor, *, M=None, N=None, K=None, bias=None, act=ACT_NONE,
         add1=None, add1_rowmode=0, add1_pre=False, ld_add1=None, add2=None, out_relu=None, out_rowmode=0, grp=0,
         pad=0, off=0, ldo=None, conv=None, pixel_shuffle=None, bias_mod=0, head=None, block_n=0, cta_group=0, epilogue=0, accumulate=False,
         out_split=False, relu_split=False, ld_relu=None, k_splits=1, splitk_ws=None):
    """out = epilogue(a @ w.T) through de_gemm (accumulate=True: out += ..., fp32 dense output only).  a: bf16 [M,K] (or NHWC [B,H,W,Cin] with conv=(B,H,W,Cin));
    w: bf16 [N,K].  See include/depth_b200.h for the epilogue contract."""
    _need_cuda(a, w, out, bias, add1, add2, out_relu)
    assert a.dtype == torch.bfloat16 and w.dtype == torch.bfloat16 and w.dim() == 2 and w.stride(1) == 1
    g = DeGemm()
    g.a = a.data_ptr()
    g.w = w.data_ptr()
    g.ldw = w.stride(0)
    g.N = int(N if N is not None else w.shape[0])
    if conv is not None:
        g.a_mode = 1
        g.B, g.H, g.W, g.Cin = [int(v) for v in conv]
        assert a.is_contiguous()
    else:
        g.a_mode = 0
        assert a.dim() == 2 and a.stride(1) == 1
        g.M = int(M if M is not None else a.shape[0])
        g.K = int(K if K is not None else a.shape[1])
        g.lda = a.stride(0)
    if bias is not None:
        assert bias.dtype == torch.float32
        g.bias = bias.data_ptr()
    g.bias_mod = int(bias_mod)
    g.act = int(act)
    if add1 is not None:
        g.add1 = add1.data_ptr()
        g.add1_dtype = _dt(add1)
        g.ld_add1 = int(ld_add1 if ld_add1 is not None else add1.stride(-2))
        g.add1_rowmode = int(add1_rowmode)
        g.add1_pre = 1 if add1_pre else 0
    if add2 is not None:
        g.add2 = add2.data_ptr()
        g.add2_dtype = _dt(add2)
        g.ld_add2 = int(add2.stride(-2)) if pixel_shuffle is None else 0
    g.out = out.data_ptr()
    g.out_dtype = _dt(out)
    g.ldo = int(ldo if ldo is not None else (out.stride(-2) if out.dim() >= 2 else 1))
    if out_relu is not None:
        assert out_relu.dtype == torch.bfloat16
        g.out_relu = out_relu.data_ptr()
        g.ld_relu = int(ld_relu if ld_relu is not None else (out_relu.stride(-2) if pixel_shuffle is None else 0))
    g.out_split = 1 if out_split else 0
    g.k_splits = int(k_splits)
    if splitk_ws is not None:
        assert splitk_ws.is_cuda and splitk_ws.is_contiguous()
        g.splitk_ws = splitk_ws.data_ptr()
        g.splitk_ws_bytes = splitk_ws.numel() * splitk_ws.element_size()
    g.relu_split = 1 if relu_split else 0
    g.out_rowmode = int(out_rowmode)
    g.grp, g.pad, g.off = int(grp), int(pad), int(off)
    if pixel_shuffle is not None:
        g.out_rowmode = 2
        g.ps_gh, g.ps_gw, g.ps_s, g.ps_cout = [int(v) for v in pixel_shuffle]
    if head is not None:
        hw, hb = head
        assert hw.dtype == torch.float32 and out.dtype == torch.float32
        g.head_w = hw.data_ptr()
        g.head_b = float(hb)
    g.block_n = int(block_n)
    g.cta_group = int(cta_group)
    g.epilogue = int(epilogue)
    g.accumulate = 1 if accumulate else 0
    _lib.check(_lib.load().de_gemm(C.byref(g), _stream()), "de_gemm")
    return out


def layernorm(x: torch.Tensor, gamma, beta, eps: float, out, y: torch.Tensor | None = None):
    """out = LN(x [+ y]) in bf16; with y the fp32 stream x is updated in place (x += y).  out may be None."""
    _need_cuda(x, gamma, beta, out, y)
    assert x.dtype == torch.float32 and x.dim() == 2 and (out is None or out.dtype == torch.bfloat16)
    assert y is None or (y.dtype in (torch.float32, torch.bfloat16) and y.shape == x.shape)
    _lib.check(_lib.load().de_layernorm(x.data_ptr(), x.stride(0), _ptr(y), _dt(y) if y is not None else 0,
                                        y.stride(0) if y is not None else 0,
                                        _ptr(gamma), _ptr(beta), x.shape[0], x.shape[1], float(eps), _ptr(out),
                                        out.stride(0) if out is not None else 0, _stream()), "de_layernorm")
    return out


def layernorm_split3(x: torch.Tensor, gamma, beta, eps: float, out3: torch.Tensor):
    """out3 bf16 [rows, 3*dim] = [hi | hi | lo] of LN(x)."""
    _need_cuda(x, gamma, beta, out3)
    assert x.dtype == torch.float32 and x.dim() == 2 and out3.dtype == torch.bfloat16 and out3.shape[1] == 3 * x.shape[1]
    _lib.check(_lib.load().de_layernorm_split3(x.data_ptr(), x.stride(0), gamma.data_ptr(), beta.data_ptr(), x.shape[0],
                                               x.shape[1], float(eps), out3.data_ptr(), out3.stride(0), _stream()),
               "de_layernorm_split3")
    return out3


def cast_tokens_split3(x: torch.Tensor, out3: torch.Tensor, grp: int = 0, skip: int = 0):
    _need_cuda(x, out3)
    assert x.dtype == torch.float32 and out3.dtype == torch.bfloat16 and x.dim() == 2 and out3.is_contiguous()
    _lib.check(_lib.load().de_cast_tokens_split3(x.data_ptr(), x.stride(0), x.shape[0], x.shape[1], grp, skip,
                                                 out3.data_ptr(), _stream()), "de_cast_tokens_split3")
    return out3


def patchify_split3(x_nchw: torch.Tensor, out3: torch.Tensor):
    _need_cuda(x_nchw, out3)
    B, _, h, w = x_nchw.shape
    assert x_nchw.dtype == torch.float32 and x_nchw.is_contiguous() and out3.dtype == torch.bfloat16 and out3.is_contiguous()
    assert tuple(out3.shape) == (B * (h // 16) * (w // 16), 3 * 768)
    _lib.check(_lib.load().de_patchify_split3(x_nchw.data_ptr(), out3.data_ptr(), B, h, w, _stream()), "de_patchify_split3")
    return out3


def attention(qkv: torch.Tensor, out: torch.Tensor, B: int, T: int, heads: int, scale: float, split: bool = False):
    _need_cuda(qkv, out)
    assert qkv.dtype == torch.bfloat16 and out.dtype == torch.bfloat16 and qkv.is_contiguous() and out.is_contiguous()
    if split:
        assert out.shape[-1] == 3 * heads * 64
        _lib.check(_lib.load().de_attention_split3(qkv.data_ptr(), out.data_ptr(), B, T, heads, float(scale), _stream()),
                   "de_attention_split3")
        return out
    _lib.check(_lib.load().de_attention(qkv.data_ptr(), out.data_ptr(), B, T, heads, float(scale), _stream()), "de_attention")
    return out


def cast_tokens(x: torch.Tensor, out: torch.Tensor, grp: int = 0, skip: int = 0):
    _need_cuda(x, out)
    assert x.dtype == torch.float32 and out.dtype == torch.bfloat16 and x.dim() == 2 and out.is_contiguous()
    _lib.check(_lib.load().de_cast_tokens(x.data_ptr(), x.stride(0), x.shape[0], x.shape[1], grp, skip,
                                          out.data_ptr(), _stream()), "de_cast_tokens")
    return out


def fill_cls(x: torch.Tensor, B: int, T: int, cls: torch.Tensor, pos: torch.Tensor):
    _need_cuda(x, cls, pos)
    _lib.check(_lib.load().de_fill_cls(x.data_ptr(), x.stride(0), B, T, x.shape[1], cls.data_ptr(), pos.data_ptr(),
                                       _stream()), "de_fill_cls")
    return x


def upsample2x(x: torch.Tensor, out: torch.Tensor, align_corners: bool):
    """x: bf16 NHWC [B,H,W,C] -> out [B,2H,2W,C]."""
    _need_cuda(x, out)
    B, H, W, Cc = x.shape
    assert x.dtype == torch.bfloat16 and x.is_contiguous() and out.is_contiguous()
    _lib.check(_lib.load().de_upsample2x(x.data_ptr(), out.data_ptr(), B, H, W, Cc, 1 if align_corners else 0,
                                         _stream()), "de_upsample2x")
    return out


def upsample2x_ex(x: torch.Tensor, out: torch.Tensor, align_corners: bool, split: bool = False):
    """x: bf16 / f32 NHWC [B,H,W,C] -> out: f32 [B,2H,2W,C], bf16 [B,2H,2W,C], or (split) the split3 operand [B,2H,2W,3C]."""
    _need_cuda(x, out)
    B, H, W, Cc = x.shape
    assert x.is_contiguous() and out.is_contiguous()
    mode = 2 if out.dtype == torch.float32 else (1 if split else 0)
    assert out.shape[-1] == (3 * Cc if mode == 1 else Cc) and (mode == 2 or out.dtype == torch.bfloat16)
    _lib.check(_lib.load().de_upsample2x_ex(x.data_ptr(), _dt(x), out.data_ptr(), mode, B, H, W, Cc,
                                            1 if align_corners else 0, _stream()), "de_upsample2x_ex")
    return out


def im2col3x3(x: torch.Tensor, out: torch.Tensor, stride: int, pad=(1, 1), out_hw=None):
    """NHWC bf16 -> [B*Ho*Wo, 9*C].  Default = torch Conv2d(k3, stride, padding=1) geometry."""
    _need_cuda(x, out)
    B, H, W, Cc = x.shape
    if out_hw is None:
        out_hw = ((H + 2 - 3) // stride + 1, (W + 2 - 3) // stride + 1)
    assert x.dtype == torch.bfloat16 and x.is_contiguous() and out.is_contiguous()
    _lib.check(_lib.load().de_im2col3x3(x.data_ptr(), out.data_ptr(), B, H, W, Cc, stride, pad[0], pad[1], out_hw[0],
                                        out_hw[1], _stream()), "de_im2col3x3")
    return out


def dwproject(x: torch.Tensor, dw_w: torch.Tensor, dw_b: torch.Tensor, pw_w: torch.Tensor, pw_b: torch.Tensor,
              out: torch.Tensor, k: int, ws: torch.Tensor, residual: torch.Tensor | None = None):
    """Fused depthwise k x k (stride 1, + bias, ReLU6) -> 1x1 projection (+ bias [+ residual]).  x NHWC bf16 [B,H,W,C];
    dw_w bf16 [k*k, C]; pw_w bf16 [cout, C]; out NHWC bf16 [B,H,W,cout]; ws: zeroed torch.uint8 scratch."""
    _need_cuda(x, dw_w, dw_b, pw_w, pw_b, out, ws, residual)
    B, H, W, Cc = x.shape
    cout = pw_w.shape[0]
    assert x.dtype == torch.bfloat16 and x.is_contiguous() and out.is_contiguous() and pw_w.is_contiguous() and dw_w.is_contiguous()
    assert tuple(out.shape) == (B, H, W, cout) and ws.dtype == torch.uint8
    _lib.check(_lib.load().de_dwproject(x.data_ptr(), dw_w.data_ptr(), dw_b.data_ptr(), pw_w.data_ptr(), pw_b.data_ptr(),
                                        _ptr(residual), out.data_ptr(), B, H, W, Cc, cout, k, ws.data_ptr(), ws.numel(),
                                        _stream()), "de_dwproject")
    return out


def dwproject_ws_bytes(B: int, H: int, W: int, cout: int) -> int:
    return int(_lib.load().de_dwproject_ws_bytes(B, H, W, cout))


def subsample(x: torch.Tensor, out: torch.Tensor, stride: int):
    _need_cuda(x, out)
    B, H, W, Cc = x.shape
    _lib.check(_lib.load().de_subsample(x.data_ptr(), out.data_ptr(), B, H, W, Cc, stride, _stream()), "de_subsample")
    return out


def groupnorm_ws_bytes(B: int, HW: int, C_: int, groups: int = 32) -> int:
    n = int(_lib.load().de_groupnorm_ws_bytes(B, HW, C_, groups))
    if n <= 0:
        raise RuntimeError(f"de_groupnorm_ws_bytes: unsupported shape B={B} HW={HW} C={C_} groups={groups}")
    return n


def groupnorm(x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, out: torch.Tensor | None, ws: torch.Tensor,
              act: int = ACT_RELU, add: torch.Tensor | None = None, out_f32: torch.Tensor | None = None,
              groups: int = 32, eps: float = 1e-5, split: bool = False, out_plain: torch.Tensor | None = None):
    """x: bf16 or f32 NHWC [B,H,W,C]; out: bf16 [B,H,W,C], or the split3 operand [B,H,W,3C] when `split`;
    ws: byte scratch (torch.uint8) - see include/depth_b200.h for the size."""
    _need_cuda(x, gamma, beta, out, ws, add, out_f32, out_plain)
    B, H, W, Cc = x.shape
    assert x.dtype in (torch.bfloat16, torch.float32) and x.is_contiguous() and ws.dtype == torch.uint8
    assert out is None or (out.dtype == torch.bfloat16 and out.shape[-1] == (3 * Cc if split else Cc))
    _lib.check(_lib.load().de_groupnorm(x.data_ptr(), _dt(x), gamma.data_ptr(), beta.data_ptr(), _ptr(add),
                                        _dt(add) if add is not None else 0, _ptr(out), 1 if split else 0, _ptr(out_f32),
                                        _ptr(out_plain), B, H * W, Cc, groups, float(eps), act, ws.data_ptr(),
                                        ws.numel(), _stream()),
               "de_groupnorm")
    return out


def stem_conv(x_nchw: torch.Tensor, w: torch.Tensor, bias, out: torch.Tensor, k: int, stride: int, act: int):
    """x: bf16 or f32 NCHW [B,3,H,W]; w: f32 [Cout, k*k*3]; out: bf16 or f32 NHWC."""
    _need_cuda(x_nchw, w, bias, out)
    B, _, H, W = x_nchw.shape
    assert x_nchw.dtype in (torch.bfloat16, torch.float32) and x_nchw.is_contiguous() and w.dtype == torch.float32
    _lib.check(_lib.load().de_stem_conv(x_nchw.data_ptr(), _dt(x_nchw), w.data_ptr(), _ptr(bias), out.data_ptr(), _dt(out),
                                        B, H, W, w.shape[0], k, stride, act, _stream()), "de_stem_conv")
    return out


def stem_im2col_split3(x_nchw: torch.Tensor, out: torch.Tensor, k: int, stride: int, kp: int):
    """x: f32 NCHW [B,3,H,W] -> out bf16 [B*Ho*Wo, 3*kp] = [hi | hi | lo] im2col rows (TF-same padding), see depth_b200.h."""
    _need_cuda(x_nchw, out)
    B, _, H, W = x_nchw.shape
    Ho, Wo = (H + stride - 1) // stride, (W + stride - 1) // stride
    assert x_nchw.dtype == torch.float32 and x_nchw.is_contiguous() and out.dtype == torch.bfloat16 and out.is_contiguous()
    assert out.numel() == B * Ho * Wo * 3 * kp
    _lib.check(_lib.load().de_stem_im2col_split3(x_nchw.data_ptr(), out.data_ptr(), B, H, W, k, stride, kp, _stream()),
               "de_stem_im2col_split3")
    return out


def head_conv(x: torch.Tensor, w: torch.Tensor, bias, head_w: torch.Tensor, head_b: float, out: torch.Tensor):
    """relu(head_b + head_w . relu(conv3x3(x, w) + bias)): x bf16 NHWC [B,H,W,Cin] (Cin % 64 == 0), w bf16 [32, 9*Cin]
    tap-major, out fp32 [B,H,W]."""
    _need_cuda(x, w, bias, head_w, out)
    B, H, W, Cin = x.shape
    assert x.dtype == torch.bfloat16 and x.is_contiguous() and w.dtype == torch.bfloat16 and w.shape[0] == 32
    assert out.dtype == torch.float32 and out.is_contiguous() and tuple(out.shape) == (B, H, W)
    assert head_w.dtype == torch.float32 and head_w.numel() == 32 and (bias is None or bias.dtype == torch.float32)
    _lib.check(_lib.load().de_head_conv(x.data_ptr(), w.data_ptr(), w.stride(0), _ptr(bias), head_w.data_ptr(),
                                        float(head_b), out.data_ptr(), B, H, W, Cin, _stream()), "de_head_conv")
    return out


def head_up_conv(x: torch.Tensor, w: torch.Tensor, bias, head_w: torch.Tensor, head_b: float, out: torch.Tensor,
                 align_corners: bool):
    """head_conv(upsample2x(x)) without materialising the upsampled tensor: x bf16 NHWC [B,Hs,Ws,Cin], out fp32
    [B,2Hs,2Ws]."""
    _need_cuda(x, w, bias, head_w, out)
    B, Hs, Ws, Cin = x.shape
    assert x.dtype == torch.bfloat16 and x.is_contiguous() and w.dtype == torch.bfloat16 and w.shape[0] == 32
    assert out.dtype == torch.float32 and out.is_contiguous() and tuple(out.shape) == (B, 2 * Hs, 2 * Ws)
    assert head_w.dtype == torch.float32 and head_w.numel() == 32 and (bias is None or bias.dtype == torch.float32)
    _lib.check(_lib.load().de_head_up_conv(x.data_ptr(), w.data_ptr(), w.stride(0), _ptr(bias), head_w.data_ptr(),
                                           float(head_b), out.data_ptr(), B, Hs, Ws, Cin, 1 if align_corners else 0,
                                           _stream()), "de_head_up_conv")
    return out


def depthwise(x: torch.Tensor, w: torch.Tensor, bias: torch.Tensor, out: torch.Tensor, k: int, stride: int, act: int):
    """x: bf16 NHWC; w: bf16 [k*k, C]; bias f32 [C]."""
    _need_cuda(x, w, bias, out)
    B, H, W, Cc = x.shape
    assert x.dtype == torch.bfloat16 and x.is_contiguous() and w.dtype == torch.bfloat16 and w.is_contiguous()
    _lib.check(_lib.load().de_depthwise(x.data_ptr(), w.data_ptr(), bias.data_ptr(), out.data_ptr(), B, H, W, Cc, k,
                                        stride, act, _stream()), "de_depthwise")
    return out


def maxpool3x3s2(x: torch.Tensor, out: torch.Tensor, split: bool = False):
    """x: bf16 or f32 NHWC; out: bf16 [B,Ho,Wo,C], or the split3 operand [B,Ho,Wo,3C] when `split`."""
    _need_cuda(x, out)
    B, H, W, Cc = x.shape
    assert out.dtype == torch.bfloat16 and out.shape[-1] == (3 * Cc if split else Cc)
    _lib.check(_lib.load().de_maxpool3x3s2(x.data_ptr(), _dt(x), out.data_ptr(), 1 if split else 0, B, H, W, Cc,
                                           _stream()), "de_maxpool3x3s2")
    return out
