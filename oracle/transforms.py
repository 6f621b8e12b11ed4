"""ORACLE (test infrastructure only - never imported by the product path).

CPU restatement of the MiDaS transform chain that the reference selects at
``/root/reference/depth_midas.py:12-16`` and applies at ``:23``.  The chain itself lives in the
un-vendored, un-pinned ``intel-isl/MiDaS`` hub repo (``hubconf.transforms()`` +
``midas/transforms.py``: ``Resize``, ``NormalizeImage``, ``PrepareForNet``), era v3.0 - see SURVEY.md
section 0.3 / 8a row A2.  PARITY UNPINNED by the reference itself (it has no tests); the
pins we add are (a) ``cv2.resize(INTER_CUBIC)`` of the locally installed OpenCV, against which
``cubic_resize_f64`` below is checked to 1e-12, and (b) the golden vectors under ``tests/golden``.

Everything here is plain numpy; python floats are used in ``get_size`` so numpy-2 scalar promotion
cannot change the half-to-even rounding.
"""
from __future__ import annotations

import numpy as np

# ---------------------------------------------------------------------------------------------
# hubconf.transforms(): the three Compose chains.  (depth_midas.py:13-16 picks dpt_ or small_.)
# ---------------------------------------------------------------------------------------------
TRANSFORM_SPECS = {
    # name: (target, resize_method, mean(RGB), std(RGB))
    "dpt_transform": (384, "minimal", (0.5, 0.5, 0.5), (0.5, 0.5, 0.5)),
    "small_transform": (256, "upper_bound", (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)),
    "default_transform": (384, "upper_bound", (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)),
}


def transform_name_for(model_type: str) -> str:
    """depth_midas.py:13-16 - anything that is not one of the two DPT names gets small_transform."""
    if model_type == "DPT_Large" or model_type == "DPT_Hybrid":
        return "dpt_transform"
    return "small_transform"


def _constrain_to_multiple_of(x: float, multiple_of: int = 32, min_val: int = 0, max_val=None) -> int:
    """[UPSTREAM] Resize.constrain_to_multiple_of: np.round is half-to-even."""
    y = int(np.round(x / multiple_of) * multiple_of)
    if max_val is not None and y > max_val:
        y = int(np.floor(x / multiple_of) * multiple_of)
    if y < min_val:
        y = int(np.ceil(x / multiple_of) * multiple_of)
    return y


def get_size(width: int, height: int, target: int, method: str, multiple_of: int = 32):
    """[UPSTREAM] Resize.get_size with keep_aspect_ratio=True. Returns (new_width, new_height)."""
    scale_height = target / height
    scale_width = target / width
    if method == "lower_bound":
        if scale_width > scale_height:
            scale_height = scale_width
        else:
            scale_width = scale_height
    elif method == "upper_bound":
        if scale_width < scale_height:
            scale_height = scale_width
        else:
            scale_width = scale_height
    elif method == "minimal":
        if abs(1 - scale_width) < abs(1 - scale_height):
            scale_height = scale_width
        else:
            scale_width = scale_height
    else:
        raise ValueError(f"resize_method {method} not implemented")
    if method == "lower_bound":
        new_height = _constrain_to_multiple_of(scale_height * height, multiple_of, min_val=target)
        new_width = _constrain_to_multiple_of(scale_width * width, multiple_of, min_val=target)
    elif method == "upper_bound":
        new_height = _constrain_to_multiple_of(scale_height * height, multiple_of, max_val=target)
        new_width = _constrain_to_multiple_of(scale_width * width, multiple_of, max_val=target)
    else:
        new_height = _constrain_to_multiple_of(scale_height * height, multiple_of)
        new_width = _constrain_to_multiple_of(scale_width * width, multiple_of)
    return new_width, new_height


# ---------------------------------------------------------------------------------------------
# cv2.resize(INTER_CUBIC) restated (OpenCV imgproc/resize.cpp: resizeGeneric_ + HResizeCubic /
# VResizeCubic, interpolateCubic with A = -0.75).  Tap tables are what the CUDA kernel consumes.
# ---------------------------------------------------------------------------------------------
def cubic_taps(src: int, dst: int):
    """Per-destination-index tap tables for one axis.

    Returns (idx int32 [dst,4] - already border-clamped source indices, w float32 [dst,4]).
    Follows OpenCV: scale = 1/(dst/src) in double; f = (float)((d+0.5)*scale-0.5); s=floor(f);
    f -= s (float); weights evaluated in float with A=-0.75; the 4th is 1-w0-w1-w2.
    """
    inv_scale = float(dst) / float(src)
    scale = 1.0 / inv_scale
    d = np.arange(dst, dtype=np.float64)
    f = ((d + 0.5) * scale - 0.5).astype(np.float32)
    s = np.floor(f).astype(np.int32)
    x = (f - s.astype(np.float32)).astype(np.float32)
    A = np.float32(-0.75)
    one = np.float32(1.0)
    w = np.empty((dst, 4), dtype=np.float32)
    w[:, 0] = ((A * (x + one) - np.float32(5) * A) * (x + one) + np.float32(8) * A) * (x + one) - np.float32(4) * A
    w[:, 1] = ((A + np.float32(2)) * x - (A + np.float32(3))) * x * x + one
    w[:, 2] = ((A + np.float32(2)) * (one - x) - (A + np.float32(3))) * (one - x) * (one - x) + one
    w[:, 3] = one - w[:, 0] - w[:, 1] - w[:, 2]
    idx = s[:, None] + np.arange(-1, 3, dtype=np.int32)[None, :]
    idx = np.clip(idx, 0, src - 1).astype(np.int32)
    return idx, w


def cubic_resize_f64(img: np.ndarray, new_w: int, new_h: int) -> np.ndarray:
    """cv2.resize(img_f64, (new_w,new_h), interpolation=cv2.INTER_CUBIC) restated for HxWxC float64.

    Same-size requests are a plain copy (OpenCV short-circuits dsize == ssize).  Horizontal pass
    first, then vertical, double accumulation with float weights - OpenCV's order.
    """
    assert img.dtype == np.float64 and img.ndim == 3
    H, W, _ = img.shape
    if (new_w, new_h) == (W, H):
        return img.copy()
    xi, xw = cubic_taps(W, new_w)
    yi, yw = cubic_taps(H, new_h)
    xw = xw.astype(np.float64)
    yw = yw.astype(np.float64)
    hor = (img[:, xi[:, 0]] * xw[None, :, 0, None] + img[:, xi[:, 1]] * xw[None, :, 1, None]
           + img[:, xi[:, 2]] * xw[None, :, 2, None] + img[:, xi[:, 3]] * xw[None, :, 3, None])
    out = (hor[yi[:, 0]] * yw[:, 0, None, None] + hor[yi[:, 1]] * yw[:, 1, None, None]
           + hor[yi[:, 2]] * yw[:, 2, None, None] + hor[yi[:, 3]] * yw[:, 3, None, None])
    return out


def midas_transform(img_rgb_u8: np.ndarray, name: str) -> np.ndarray:
    """The Compose chain: /255 -> Resize -> NormalizeImage -> PrepareForNet -> unsqueeze(0).

    img_rgb_u8: uint8 [H,W,3] RGB (depth_midas.py:22 already swapped BGR->RGB).
    Returns float32 [1,3,h,w] (a numpy array; the reference wraps it with torch.from_numpy).
    """
    target, method, mean, std = TRANSFORM_SPECS[name]
    img = img_rgb_u8 / 255.0                                   # float64 HWC
    H, W = img.shape[:2]
    new_w, new_h = get_size(W, H, target, method)
    img = cubic_resize_f64(img, new_w, new_h)
    img = (img - np.asarray(mean, dtype=np.float64)) / np.asarray(std, dtype=np.float64)
    img = np.ascontiguousarray(np.transpose(img, (2, 0, 1))).astype(np.float32)
    return img[None]


def midas_transform_bgr(img_bgr_u8: np.ndarray, model_type: str) -> np.ndarray:
    """depth_midas.py:22-23: cvtColor(BGR2RGB) then self.transform(img)."""
    rgb = np.ascontiguousarray(img_bgr_u8[:, :, ::-1])
    return midas_transform(rgb, transform_name_for(model_type))


def nv12_to_bgr(nv12: np.ndarray) -> np.ndarray:
    """NV12 [H*3/2, W] uint8 -> BGR [H, W, 3] uint8, restating OpenCV's cvtColor(COLOR_YUV2BGR_NV12) (ITU-R BT.601
    limited range, 20-bit fixed point, saturating) - the host conversion K1's NV12 path fuses.
    Pinned bit-exact against the installed cv2 in tests/test_oracle_cpu.py."""
    H = nv12.shape[0] * 2 // 3
    W = nv12.shape[1]
    Y = nv12[:H].astype(np.int64)
    UV = nv12[H:].reshape(H // 2, W // 2, 2).astype(np.int64)
    U = np.repeat(np.repeat(UV[..., 0], 2, 0), 2, 1) - 128
    V = np.repeat(np.repeat(UV[..., 1], 2, 0), 2, 1) - 128
    y = np.maximum(0, Y - 16) * 1220542 + (1 << 19)
    r = (y + 1673527 * V) >> 20
    g = (y - 852492 * V - 409993 * U) >> 20
    b = (y + 2116026 * U) >> 20
    return np.stack([b, g, r], -1).clip(0, 255).astype(np.uint8)
