"""CPU suite (-m "not gpu"): the oracle against its pins (cv2, torch, the reference's own face.py via
golden vectors), the host logic of the C ABI (get_size, tap tables), and the exported symbol table."""
import os
import re

import numpy as np
import pytest
import torch

from oracle import transforms as T
from oracle.predict import MODEL_TYPES, OracleDepthEstimator, build_model, get_depth_iris, rel2abs, synthetic_frame

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
GOLD = os.path.join(HERE, "golden")


# ----------------------------------------------------------------------------- get_size
SIZE_TABLE = [
    # W, H, transform, (net_w, net_h)   (SURVEY.md section 8 size table)
    (640, 480, "small_transform", (256, 192)),
    (1280, 720, "small_transform", (256, 128)),       # 144/32 = 4.5 -> half-to-even -> 4
    (1920, 1080, "small_transform", (256, 128)),
    (384, 384, "small_transform", (256, 256)),
    (640, 480, "dpt_transform", (512, 384)),
    (1280, 720, "dpt_transform", (672, 384)),
    (1920, 1080, "dpt_transform", (672, 384)),
    (384, 384, "dpt_transform", (384, 384)),
]


@pytest.mark.parametrize("W,H,name,expect", SIZE_TABLE)
def test_get_size_table(W, H, name, expect, lib):
    target, method, _, _ = T.TRANSFORM_SPECS[name]
    assert T.get_size(W, H, target, method) == expect
    from depth_estimation_b200 import ops
    assert ops.get_size(W, H, "DPT_Large" if name == "dpt_transform" else "MiDaS_small") == expect


def test_get_size_c_abi_matches_oracle_sweep(lib):
    from depth_estimation_b200 import ops
    rng = np.random.default_rng(0)
    for _ in range(400):
        W, H = int(rng.integers(16, 4000)), int(rng.integers(16, 3000))
        for mt, name in (("DPT_Hybrid", "dpt_transform"), ("MiDaS_small", "small_transform")):
            target, method, _, _ = T.TRANSFORM_SPECS[name]
            assert ops.get_size(W, H, mt) == T.get_size(W, H, target, method), (W, H, mt)


# ----------------------------------------------------------------------------- cubic resize pins
@pytest.mark.parametrize("W,H,nw,nh", [(640, 480, 256, 192), (1280, 720, 256, 128), (1920, 1080, 672, 384),
                                       (100, 37, 64, 96), (37, 100, 320, 32), (384, 384, 384, 384)])
def test_cubic_resize_restatement_equals_cv2(W, H, nw, nh):
    cv2 = pytest.importorskip("cv2")
    img = np.random.default_rng(1).integers(0, 256, (H, W, 3), dtype=np.uint8) / 255.0
    a = T.cubic_resize_f64(img, nw, nh)
    b = cv2.resize(img, (nw, nh), interpolation=cv2.INTER_CUBIC)
    assert np.abs(a - b).max() <= 1e-12


@pytest.mark.parametrize("src,dst", [(640, 256), (480, 192), (1920, 672), (1080, 384), (384, 384), (37, 96), (720, 128)])
def test_tap_tables_c_abi_bit_exact_vs_oracle(src, dst, lib):
    from depth_estimation_b200 import ops
    idx, w = ops.cubic_taps(src, dst, 0)
    oi, ow = T.cubic_taps(src, dst)
    assert np.array_equal(idx, oi)
    assert np.array_equal(w.view(np.uint32), ow.view(np.uint32))       # bit-exact fp32 weights


@pytest.mark.parametrize("h,w,H,W", [(192, 256, 480, 640), (128, 256, 720, 1280), (384, 672, 1080, 1920), (96, 96, 50, 70)])
def test_bicubic_taps_reproduce_torch(h, w, H, W, lib):
    """mode-1 tables (ATen coordinate rule) + separable dot products == F.interpolate(bicubic)."""
    from depth_estimation_b200 import ops
    pred = torch.rand((h, w), generator=torch.Generator().manual_seed(3)) * 30
    ref = torch.nn.functional.interpolate(pred[None, None], size=(H, W), mode="bicubic", align_corners=False)[0, 0].numpy()
    xi, xw = ops.cubic_taps(w, W, 1)
    yi, yw = ops.cubic_taps(h, H, 1)
    p = pred.numpy().astype(np.float64)
    hor = sum(p[:, xi[:, k]] * xw[None, :, k] for k in range(4))
    out = sum(hor[yi[:, k]] * yw[:, k, None] for k in range(4))
    assert np.abs(out - ref).max() / np.abs(ref).max() < 2e-6


def test_transform_golden():
    g = np.load(os.path.join(GOLD, "transform.npz"))
    for key in [k[:-5] for k in g.files if k.endswith("_size")]:
        name, wh = key.rsplit("_", 1)
        W, H = [int(v) for v in wh.split("x")]
        target, method, _, _ = T.TRANSFORM_SPECS[name]
        nw, nh = T.get_size(W, H, target, method)
        assert [nw, nh] == g[key + "_size"].tolist()
        frame = synthetic_frame(W, H, 5, "U")
        out = T.midas_transform(np.ascontiguousarray(frame[:, :, ::-1]), name)[0]
        p = g[key + "_probe_idx"]
        assert np.array_equal(out[p[:, 0], p[:, 1], p[:, 2]], g[key + "_probe_val"])
        xi, xw = T.cubic_taps(W, nw)
        assert np.array_equal(xi[:8], g[key + "_xidx8"]) and np.array_equal(xw[:8], g[key + "_xw8"])


def test_bicubic_golden():
    g = np.load(os.path.join(GOLD, "bicubic.npz"))
    for key in [k[:-10] for k in g.files if k.endswith("_probe_idx")]:
        a, b = key.split("_")
        h, w = [int(v) for v in a.split("x")]
        H, W = [int(v) for v in b.split("x")]
        pred = torch.rand((1, 1, h, w), generator=torch.Generator().manual_seed(h * 7 + W)) * 30
        ref = torch.nn.functional.interpolate(pred, size=(H, W), mode="bicubic", align_corners=False)[0, 0].numpy()
        p = g[key + "_probe_idx"]
        assert np.allclose(ref[p[:, 0], p[:, 1]], g[key + "_probe_val"], rtol=1e-6, atol=1e-6)


# ----------------------------------------------------------------------------- face tail (reference-generated)
def test_face_tail_matches_reference_golden():
    g = np.load(os.path.join(GOLD, "face_tail.npz"))
    rng = np.random.default_rng(11)
    shapes = [(480, 640), (720, 1280), (1080, 1920), (300, 300)]
    cases = g["cases"]
    ci = 0
    for i, (H, W) in enumerate(shapes):
        depth = rng.random((H, W)).astype(np.float32) * 50
        assert np.array_equal(depth[::16, ::16], g[f"depth{i}"])
        for j in range(6):
            l = (float(rng.uniform(0, W * 1.1)), float(rng.uniform(0, H * 1.1)))
            r = (float(rng.uniform(0, W * 1.1)), float(rng.uniform(0, H * 1.1)))
            row = cases[ci]
            ci += 1
            assert row[0] == i and np.allclose(row[1:5], [l[0], l[1], r[0], r[1]])
            try:
                ri, _ = get_depth_iris(depth, l, r)
                got = (float(ri), float(rel2abs(ri)), 0)
            except IndexError:
                got = (0.0, 0.0, 1)
            assert got[2] == int(row[7])
            assert got[0] == row[5] and got[1] == row[6]


@pytest.mark.skipif(not os.path.exists("/root/reference/face.py"), reason="reference tree not present on this box")
def test_face_tail_against_live_reference():
    import importlib.util
    spec = importlib.util.spec_from_file_location("ref_face", "/root/reference/face.py")
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    depth = np.random.default_rng(5).random((480, 640)).astype(np.float32)
    f = ref.FaceDet(20, [505, 504, 675, 501])
    f.mesh = object()
    f.l_iris = {"center": (100.7, 200.2), "radius": 5.0}
    f.r_iris = {"center": (160.1, 210.9), "radius": 5.0}
    f.get_depth(depth)
    f.rel2abs()
    ri, _ = get_depth_iris(depth, (100.7, 200.2), (160.1, 210.9))
    assert float(ri) == float(f.ri_depth) and float(rel2abs(ri)) == float(f.abs_depth)


# ----------------------------------------------------------------------------- oracle models
def test_parameter_counts_match_upstream_figures():
    counts = {mt: sum(p.numel() for p in build_model(mt).parameters()) for mt in MODEL_TYPES}
    assert abs(counts["DPT_Large"] / 1e6 - 344) < 1.0
    assert abs(counts["DPT_Hybrid"] / 1e6 - 123) < 1.0
    assert abs(counts["MiDaS_small"] / 1e6 - 21.3) < 0.5


def test_tf_same_padding_by_hand():
    from oracle.midas_small import Conv2dSame
    from oracle.vit import _pad_same
    x = torch.arange(16.0).view(1, 1, 4, 4)
    # k3 s2 on even input: total pad 1 -> (0 before, 1 after)
    p = _pad_same(x, 3, 2)
    assert p.shape[-2:] == (5, 5) and p[0, 0, 0, 0] == 0 and p[0, 0, 4, 4] == 0 and p[0, 0, 3, 3] == 15
    # k5 s2: total 3 -> (1, 2)
    p = _pad_same(x, 5, 2)
    assert p.shape[-2:] == (7, 7) and p[0, 0, 1, 1] == 0 and p[0, 0, 0, 0] == 0
    conv = Conv2dSame(1, 1, 3, 2)
    with torch.no_grad():
        conv.weight.fill_(1.0)
    y = conv(x)
    assert y.shape[-2:] == (2, 2)
    assert y[0, 0, 0, 0] == x[0, 0, 0:3, 0:3].sum() and y[0, 0, 1, 1] == x[0, 0, 2:4, 2:4].sum()


def test_param_inventory_matches_oracle_state_dict():
    from depth_estimation_b200.init import param_shapes, random_state_dict
    for mt in MODEL_TYPES:
        sd = {k: v for k, v in build_model(mt).state_dict().items() if "num_batches_tracked" not in k}
        ps = param_shapes(mt)
        assert set(sd) == set(ps)
        assert all(tuple(sd[k].shape) == tuple(ps[k]) for k in ps)
    rs = random_state_dict("MiDaS_small", 1)
    m = build_model("MiDaS_small")
    m.load_state_dict(rs, strict=False)


@pytest.mark.parametrize("key", ["MiDaS_small_640x480", "MiDaS_small_1280x720", "DPT_Hybrid_384x384"])
def test_model_golden(key):
    g = np.load(os.path.join(GOLD, "models.npz"))
    mt, wh = key.rsplit("_", 1)
    W, H = [int(v) for v in wh.split("x")]
    out = OracleDepthEstimator(mt, seed=0).predict(synthetic_frame(W, H, 3, "S"))
    p = g[key + "_probe_idx"]
    assert out.shape == (H, W) and out.dtype == np.float32
    assert np.allclose(out[p[:, 0], p[:, 1]], g[key + "_probe_val"], rtol=2e-4, atol=2e-4)
    assert (out > 0).mean() > 0.9


# ----------------------------------------------------------------------------- C ABI surface
def test_library_exports_every_declared_symbol(lib):
    from depth_estimation_b200 import _lib
    hdr = open(os.path.join(ROOT, "include", "depth_b200.h")).read()
    declared = set(re.findall(r"\b(de_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/depth_b200.h but not exported"
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    assert lib.de_version() >= 1


def test_product_fails_loudly_without_cuda():
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    from depth_estimation_b200.depth_midas import DepthEstimator
    with pytest.raises(RuntimeError):
        DepthEstimator("MiDaS_small")
    from depth_estimation_b200 import ops
    with pytest.raises(RuntimeError):
        ops.layernorm(torch.zeros(2, 128), torch.ones(128), torch.zeros(128), 1e-6, torch.zeros(2, 128, dtype=torch.bfloat16))


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "depth_estimation_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, re.M), f


def test_checkpoint_ingestion_roundtrip_and_validation(tmp_path):
    """a checkpoint written in upstream naming (here: the oracle module's own state_dict, wrapped and
    prefixed the way training scripts leave them) loads, validates and hashes identically; damaged ones fail loudly."""
    from depth_estimation_b200.weights import load_checkpoint, state_dict_checksum
    from oracle.predict import build_model
    model = build_model("MiDaS_small", seed=3)
    sd = model.state_dict()
    path = tmp_path / "midas_small.pt"
    torch.save({"state_dict": {"module." + k: v for k, v in sd.items()}}, path)
    got = load_checkpoint(str(path), "MiDaS_small")
    ref = load_checkpoint(sd, "MiDaS_small")
    assert list(got) == list(ref) and state_dict_checksum(got) == state_dict_checksum(ref)
    assert all(v.dtype == torch.float32 and v.device.type == "cpu" for v in got.values())
    assert not any(k.endswith("num_batches_tracked") for k in got)
    k0 = next(iter(ref))
    broken = dict(sd)
    del broken[k0]
    with pytest.raises(ValueError, match="missing"):
        load_checkpoint(broken, "MiDaS_small")
    with pytest.raises(ValueError, match="mis-shaped"):
        load_checkpoint({**sd, k0: sd[k0][..., :1]}, "MiDaS_small")
    with pytest.raises(ValueError, match="unexpected"):
        load_checkpoint({**sd, "pretrained.bogus": torch.zeros(1)}, "MiDaS_small")
    assert load_checkpoint({**sd, "pretrained.bogus": torch.zeros(1)}, "MiDaS_small", strict=False).keys() == ref.keys()
    with pytest.raises(ValueError):
        load_checkpoint(sd, "DPT_Hybrid")
    assert state_dict_checksum(ref) != state_dict_checksum({**ref, k0: ref[k0] + 1})


def test_nv12_restatement_matches_opencv():
    """N4: the fixed-point NV12 -> BGR rule K1 fuses is the installed OpenCV's, bit for bit (incl. saturation)."""
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(3)
    for H, W in ((16, 32), (270, 482), (2, 2)):
        nv12 = rng.integers(0, 256, (H * 3 // 2, W), dtype=np.uint8)
        assert np.array_equal(T.nv12_to_bgr(nv12), cv2.cvtColor(nv12, cv2.COLOR_YUV2BGR_NV12))
    ext = np.zeros((3, 2), dtype=np.uint8)
    for yv, u, v in ((0, 0, 0), (255, 255, 255), (255, 0, 255), (16, 255, 0), (235, 128, 128)):
        ext[:2] = yv
        ext[2] = (u, v)
        assert np.array_equal(T.nv12_to_bgr(ext), cv2.cvtColor(ext, cv2.COLOR_YUV2BGR_NV12))
