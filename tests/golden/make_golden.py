"""Generates the golden fixtures under tests/golden/ (run in the BUILD container, where /root/reference
and cv2 exist; the GPU box only reads the committed .npz files).

  python tests/golden/make_golden.py

* face_tail.npz   - produced by importing the REFERENCE's own face.py (/root/reference/face.py:91-139):
                    FaceDet.get_depth + rel2abs on seeded depth maps and synthetic iris centres.
* transform.npz   - cv2.resize(INTER_CUBIC)-based upstream transform chain (the real third-party library the
                    reference calls through the hub transform) on seeded frames: sizes, tap probes, pixels.
* bicubic.npz     - torch F.interpolate(bicubic, align_corners=False) (depth_midas.py:28-33) probes.
* models.npz      - oracle model outputs (random-init seed 0) at 32 probe pixels per model/config.
"""
import importlib.util
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import transforms as T  # noqa: E402
from oracle.predict import OracleDepthEstimator, synthetic_frame  # noqa: E402

SIZES = [(640, 480), (1280, 720), (1920, 1080), (384, 384), (333, 517)]


def probes(shape, n=32, seed=7):
    rng = np.random.default_rng(seed)
    return np.stack([rng.integers(0, s, n) for s in shape], axis=1)


def face_tail():
    spec = importlib.util.spec_from_file_location("ref_face", "/root/reference/face.py")
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    rng = np.random.default_rng(11)
    out = {}
    cases = []
    for i, (H, W) in enumerate([(480, 640), (720, 1280), (1080, 1920), (300, 300)]):
        depth = rng.random((H, W)).astype(np.float32) * 50
        for j in range(6):
            # centres are (x, y) floats as cv2.minEnclosingCircle returns them; some beyond the frame
            l = (float(rng.uniform(0, W * 1.1)), float(rng.uniform(0, H * 1.1)))
            r = (float(rng.uniform(0, W * 1.1)), float(rng.uniform(0, H * 1.1)))
            f = ref.FaceDet(20, [505, 504, 675, 501])
            f.mesh = object()
            f.l_iris = {"center": l, "radius": 9.0}
            f.r_iris = {"center": r, "radius": 9.0}
            try:
                f.get_depth(depth)
                f.rel2abs()
                res = (float(f.ri_depth), float(f.abs_depth), 0)
            except IndexError:
                res = (0.0, 0.0, 1)             # the reference raises when its unclamped index is out of range
            cases.append((i, l[0], l[1], r[0], r[1]) + res)
        out[f"depth{i}"] = depth[::16, ::16].copy()          # checksum grid only; maps are regenerated from the seed
    out["cases"] = np.asarray(cases, dtype=np.float64)
    f = ref.FaceDet(20, [505, 504, 675, 501])
    out["focal"] = np.asarray([f.f, f.w_pix, f.d_2_obj])
    np.savez_compressed(os.path.join(HERE, "face_tail.npz"), **out)


def transform():
    import cv2
    out = {}
    for (W, H) in SIZES:
        for name in ("dpt_transform", "small_transform"):
            target, method, mean, std = T.TRANSFORM_SPECS[name]
            nw, nh = T.get_size(W, H, target, method)
            frame = synthetic_frame(W, H, 5, "U")
            rgb = np.ascontiguousarray(frame[:, :, ::-1])
            img = cv2.resize(rgb / 255.0, (nw, nh), interpolation=cv2.INTER_CUBIC)
            img = (img - np.asarray(mean)) / np.asarray(std)
            ref = np.ascontiguousarray(np.transpose(img, (2, 0, 1))).astype(np.float32)
            key = f"{name}_{W}x{H}"
            out[key + "_size"] = np.asarray([nw, nh])
            p = probes(ref.shape)
            out[key + "_probe_idx"] = p
            out[key + "_probe_val"] = ref[p[:, 0], p[:, 1], p[:, 2]]
            xi, xw = T.cubic_taps(W, nw)
            yi, yw = T.cubic_taps(H, nh)
            out[key + "_xidx8"], out[key + "_xw8"] = xi[:8], xw[:8]
            out[key + "_yidx8"], out[key + "_yw8"] = yi[-8:], yw[-8:]
    np.savez_compressed(os.path.join(HERE, "transform.npz"), **out)


def bicubic():
    out = {}
    for (h, w, H, W) in [(192, 256, 480, 640), (128, 256, 720, 1280), (384, 672, 1080, 1920), (384, 384, 384, 384)]:
        g = torch.Generator().manual_seed(h * 7 + W)
        pred = torch.rand((1, 1, h, w), generator=g) * 30
        ref = torch.nn.functional.interpolate(pred, size=(H, W), mode="bicubic", align_corners=False)[0, 0].numpy()
        key = f"{h}x{w}_{H}x{W}"
        p = probes(ref.shape, 64)
        out[key + "_probe_idx"] = p
        out[key + "_probe_val"] = ref[p[:, 0], p[:, 1]]
    np.savez_compressed(os.path.join(HERE, "bicubic.npz"), **out)


def models():
    out = {}
    for mt, W, H in [("MiDaS_small", 640, 480), ("MiDaS_small", 1280, 720), ("DPT_Hybrid", 384, 384),
                     ("DPT_Large", 384, 384)]:
        est = OracleDepthEstimator(mt, seed=0)
        frame = synthetic_frame(W, H, 3, "S")
        ref = est.predict(frame)
        p = probes(ref.shape, 32)
        key = f"{mt}_{W}x{H}"
        out[key + "_probe_idx"] = p
        out[key + "_probe_val"] = ref[p[:, 0], p[:, 1]]
        out[key + "_minmax"] = np.asarray([ref.min(), ref.max(), ref.mean()])
        print(key, ref.min(), ref.max())
    np.savez_compressed(os.path.join(HERE, "models.npz"), **out)


if __name__ == "__main__":
    face_tail()
    transform()
    bicubic()
    models()
    print("golden fixtures written to", HERE)
