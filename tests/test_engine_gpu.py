"""The handle-level C ABI (include/depth_b200.h: de_create .. de_destroy) on the GPU box:
a host with no Python (tests/c/engine_demo.c, plain C) must produce the map `DepthEstimator.predict` returns, bit for
bit; the ABI's error behaviour (missing / mis-shaped weights, wrong call order) is checked through ctypes."""
import ctypes as C
import os
import struct
import subprocess

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEMO = os.path.join(ROOT, "depth_estimation_b200", "csrc", "build", "engine_demo")


def _need_gpu():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")


def _write_weights(path, sd):
    with open(path, "wb") as f:
        f.write(struct.pack("<i", len(sd)))
        for k, v in sd.items():
            t = v.detach().to("cpu", torch.float32).contiguous()
            kb = k.encode()
            f.write(struct.pack("<i", len(kb)))
            f.write(kb)
            f.write(struct.pack("<i", t.dim()))
            f.write(struct.pack(f"<{t.dim()}q", *t.shape))
            f.write(t.numpy().tobytes())


def _build_demo():
    if os.path.exists(DEMO):
        return
    src = os.path.join(ROOT, "tests", "c", "engine_demo.c")
    lib_dir = os.path.join(ROOT, "depth_estimation_b200")
    os.makedirs(os.path.dirname(DEMO), exist_ok=True)
    subprocess.run(["gcc", "-O2", "-std=c99", "-I" + os.path.join(ROOT, "include"), "-I/usr/local/cuda/include", "-o", DEMO, src,
                    "-L" + lib_dir, "-l:libdepthb200.so", "-L/usr/local/cuda/lib64", "-lcudart",
                    "-Wl,-rpath," + lib_dir], check=True)


@pytest.mark.parametrize("mt,B,W,H", [("MiDaS_small", 2, 640, 480), ("DPT_Hybrid", 1, 384, 384), ("DPT_Large", 2, 384, 384)])
def test_c_host_matches_depth_estimator(tmp_path, mt, B, W, H):
    _need_gpu()
    from depth_estimation_b200.depth_midas import DepthEstimator
    from depth_estimation_b200.init import random_state_dict
    _build_demo()
    sd = random_state_dict(mt, seed=4)
    rng = np.random.default_rng(8)
    frames = rng.integers(0, 256, (B, H, W, 3), dtype=np.uint8)
    wpath, fpath, opath = (str(tmp_path / n) for n in ("w.bin", "f.bin", "o.bin"))
    _write_weights(wpath, sd)
    with open(fpath, "wb") as f:
        f.write(struct.pack("<3i", B, H, W))
        f.write(frames.tobytes())
    env = dict(os.environ)
    env["LD_LIBRARY_PATH"] = os.path.join(ROOT, "depth_estimation_b200") + ":" + env.get("LD_LIBRARY_PATH", "")
    res = subprocess.run([DEMO, mt, wpath, fpath, opath, "3"], capture_output=True, text=True, env=env, timeout=600)
    assert res.returncode == 0, res.stdout + res.stderr
    print(res.stdout.strip())
    got = np.fromfile(opath, dtype=np.float32).reshape(B, H, W)
    est = DepthEstimator(mt, weights=sd)
    want = est.predict_batch(frames)
    assert np.isfinite(got).all() and got.std() > 0
    assert np.array_equal(got, want)


def test_handle_api_errors():
    _need_gpu()
    from depth_estimation_b200 import _lib
    from depth_estimation_b200.init import random_state_dict
    lib = _lib.load()
    h = C.c_void_p()
    assert lib.de_create(b"ResNeXt", 0, 0, C.byref(h)) != 0 and b"unknown model_type" in lib.de_last_error()
    assert lib.de_create(b"MiDaS_small", 99, 0, C.byref(h)) != 0
    assert lib.de_create(b"MiDaS_small", 0, 7, C.byref(h)) != 0
    assert lib.de_create(b"MiDaS_small", 0, 0, C.byref(h)) == 0
    io = _lib.DePredictIO()
    dummy = torch.zeros((1, 64, 64, 3), dtype=torch.uint8, device="cuda")
    out = torch.zeros((1, 64, 64), dtype=torch.float32, device="cuda")
    io.frames, io.out_map = dummy.data_ptr(), out.data_ptr()
    assert lib.de_predict(h, 1, 64, 64, C.byref(io), 0, None) != 0 and b"finalize" in lib.de_last_error()
    assert lib.de_finalize_weights(h) != 0 and b"never set" in lib.de_last_error()      # names the missing key
    lib.de_destroy(h)

    def feed(handle, sd):
        for k, v in sd.items():
            t = v.contiguous()
            shape = (C.c_int64 * max(1, t.dim()))(*t.shape)
            assert lib.de_set_weight(handle, k.encode(), t.data_ptr(), shape, t.dim(), _lib.DE_DTYPE_F32) == 0

    sd = random_state_dict("MiDaS_small", seed=0)
    bad = dict(sd)
    bad["scratch.layer1_rn.weight"] = torch.zeros(64, 32, 3, 1)
    assert lib.de_create(b"MiDaS_small", 0, 0, C.byref(h)) == 0
    feed(h, bad)
    assert lib.de_finalize_weights(h) != 0
    msg = lib.de_last_error()
    assert b"scratch.layer1_rn.weight" in msg and b"shape" in msg
    lib.de_destroy(h)

    assert lib.de_create(b"MiDaS_small", 0, 0, C.byref(h)) == 0
    feed(h, sd)
    assert lib.de_finalize_weights(h) == 0
    t = sd["scratch.layer1_rn.weight"]
    shape = (C.c_int64 * 4)(*t.shape)
    assert lib.de_set_weight(h, b"scratch.layer1_rn.weight", t.data_ptr(), shape, 4, _lib.DE_DTYPE_F32) != 0   # after finalize
    assert lib.de_finalize_weights(h) != 0                                                                       # twice
    io2 = _lib.DePredictIO()
    io2.frames = dummy.data_ptr()
    assert lib.de_predict(h, 1, 64, 64, C.byref(io2), 0, None) != 0 and b"no output" in lib.de_last_error()
    nh, nw = C.c_int(), C.c_int()
    assert lib.de_plan(h, 1, 480, 640, C.byref(nh), C.byref(nw)) == 0 and (nh.value, nw.value) == (192, 256)
    assert lib.de_predict(h, 1, 64, 64, C.byref(io), 1, torch.cuda.current_stream().cuda_stream) == 0
    torch.cuda.synchronize()
    assert torch.isfinite(out).all()
    lib.de_destroy(h)
    lib.de_destroy(None)                                                                                         # no-op


def test_engine_internal_graph_equals_eager():
    """de_predict(use_graph=1) (capture on first use, replay afterwards) returns what the eager launch sequence returns."""
    _need_gpu()
    from depth_estimation_b200.engine import Engine
    from depth_estimation_b200.init import random_state_dict
    dev = torch.device("cuda", 0)
    eng = Engine("MiDaS_small", random_state_dict("MiDaS_small", 1), dev)
    rng = np.random.default_rng(3)
    frames = torch.from_numpy(rng.integers(0, 256, (3, 120, 160, 3), dtype=np.uint8)).to(dev)
    eager = torch.empty((3, 120, 160), dtype=torch.float32, device=dev)
    eng.predict(3, 120, 160, frames, out_map=eager)
    g = torch.empty_like(eager)
    for _ in range(3):
        g.zero_()
        eng.predict(3, 120, 160, frames, out_map=g, use_graph=True)
        torch.cuda.synchronize()
        assert torch.equal(g, eager)
