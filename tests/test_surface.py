"""Surface conformance with the reference modules: names, signatures, attributes."""
import inspect
import os

import numpy as np
import pytest


def test_module_names_and_signatures():
    import depth_midas
    import face
    import midas
    for mod in (depth_midas, midas):
        sig = inspect.signature(mod.DepthEstimator.__init__)
        assert list(sig.parameters)[:2] == ["self", "model_type"]
        assert all(p.kind == inspect.Parameter.KEYWORD_ONLY for n, p in list(sig.parameters.items())[2:])
        assert list(inspect.signature(mod.DepthEstimator.predict).parameters) == ["self", "img"]
        assert list(inspect.signature(mod.VidStream.__init__).parameters) == ["self", "estimator", "src", "output"]
        assert hasattr(mod.VidStream, "update") and hasattr(mod.VidStream, "write_output")
    assert list(inspect.signature(face.FaceDet.__init__).parameters) == ["self", "d_2_obj", "points"]


@pytest.mark.skipif(not os.path.exists("/root/reference/face.py"), reason="reference tree not present on this box")
def test_facedet_matches_reference_attribute_for_attribute():
    import importlib.util
    from depth_estimation_b200.face import FaceDet
    spec = importlib.util.spec_from_file_location("ref_face", "/root/reference/face.py")
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    a, b = ref.FaceDet(20, [505, 504, 675, 501]), FaceDet(20, [505, 504, 675, 501])
    assert set(vars(a)) == set(vars(b))
    assert {n for n in dir(ref.FaceDet) if not n.startswith("_")} <= {n for n in dir(FaceDet) if not n.startswith("_")}
    rng = np.random.default_rng(3)
    depth = rng.random((480, 640)).astype(np.float32) * 40
    for f in (a, b):
        f.mesh = object()
        f.l_iris = {"center": (211.6, 300.9), "radius": 7.5}
        f.r_iris = {"center": (470.2, 630.0), "radius": 7.0}
        f.get_depth(depth)
        f.rel2abs()
        f.s2c_dist(f.w_iris, 2 * f.l_iris["radius"])
        f.get_headw((100, 200), (180, 210))
    for k in ("f", "w_pix", "d_2_obj", "ri_depth", "abs_depth", "s2c_d", "head_w"):
        assert getattr(a, k) == getattr(b, k), k
    assert a.s2c_ds == b.s2c_ds and a.abs_depths == b.abs_depths and a.head_measurements == b.head_measurements
    assert a.rmse() == b.rmse() and a.mae() == b.mae()
    # the fused-sample route gives the same bookkeeping
    l, r = b.sample_indices(depth.shape)
    c = FaceDet(20, [505, 504, 675, 501])
    c.set_depth_samples(depth[l[0], l[1]], depth[r[0], r[1]])
    assert c.ri_depth == a.ri_depth


def test_vidstream_headless_with_fake_estimator():
    from depth_estimation_b200.depth_midas import VidStream

    class Fake:
        def predict(self, img):
            return (img[:, :, 0].astype(np.float32) * 1.5) + 300.0     # > 255: exercises the uint8 wrap

    frames = [np.full((4, 6, 3), i, dtype=np.uint8) for i in range(3)]
    vs = VidStream(Fake(), frames, None)
    vs.update()
    assert vs.cnt == 3 and len(vs.outputs) == 3
    assert vs.outputs[1].shape == (4, 6, 3) and vs.outputs[1].dtype == np.uint8
    assert int(vs.outputs[2][0, 0, 0]) == int(np.float32(303.0).astype(np.uint8))


class _FakeDetector:
    """Stands in for detectors.PersonDetector (MediaPipe): deterministic landmarks derived from the frame content;
    every third frame has no face (body branch)."""
    HEAD = (0, 1)

    def __init__(self, face):
        self.face = face
        self.calls = []

    def findIris(self, frame):
        k = int(frame[0, 0, 0])
        self.calls.append(("iris", k))
        if k % 3 == 2:
            return
        H, W = frame.shape[:2]
        self.face.mesh = {0: (10 + k, 20), 1: (30 + k, 25)}
        self.face.l_iris = {"center": (5.0 + k, 7.5), "radius": 4.0 + 0.25 * k}
        self.face.r_iris = {"center": (H - 3.0, 9.0 + k), "radius": 4.5}

    def findBody(self, frame):
        self.calls.append(("body", int(frame[0, 0, 0])))
        return frame, ((12, 30), (40, 34))

    def visualize(self, frame):
        pass


class _FakeStreamEstimator:
    """predict + predict_stream with the semantics of DepthEstimator (in-order results, two in flight)."""

    def predict(self, img):
        return (img[:, :, 1].astype(np.float32) * 2.25) + img[:, :, 0].astype(np.float32) * 31.0 - 7.5

    def predict_stream(self, batches, depth=2):
        pend = []
        for b in batches:
            pend.append(np.stack([self.predict(f) for f in b]))
            if len(pend) == depth:
                yield pend.pop(0)
        while pend:
            yield pend.pop(0)


def _stream_frames(n, H=48, W=64):
    rng = np.random.default_rng(11)
    fr = rng.integers(0, 256, (n, H, W, 3), dtype=np.uint8)
    fr[:, 0, 0, 0] = np.arange(n)
    return list(fr)


def test_midas_mediapipe_stream_serial_equals_pipelined():
    """the pipelined headless loop hands FaceDet exactly the numbers the reference's serial loop
    would (same frame order, landmarks of the right frame), incl. the no-face / body branch."""
    import midas_mediapipe
    from depth_estimation_b200.face import FaceDet
    sig = inspect.signature(midas_mediapipe.VidStream.__init__)
    assert list(sig.parameters) == ["self", "estimator", "detector", "face", "src", "output"]
    runs = []
    for pipelined in (False, True):
        face = FaceDet(20, [505, 504, 675, 501])
        det = _FakeDetector(face)
        vs = midas_mediapipe.VidStream(_FakeStreamEstimator(), det, face, _stream_frames(7), None)
        rep = vs.stream(pipelined=pipelined)
        assert vs.cnt == 7 and len(vs.outputs) == 7
        assert set(rep) == {"iris", "body", "depth"} and len(vs.performance["depth"]) == 7
        assert len(vs.performance["body"]) == 2 and len(vs.performance["iris"]) == 7
        runs.append((face, vs))
    (fa, va), (fb, vb) = runs
    assert fa.ri_depths == fb.ri_depths and fa.abs_depths == fb.abs_depths and fa.s2c_ds == fb.s2c_ds
    assert fa.head_measurements == fb.head_measurements and fa.errors == fb.errors
    assert all(np.array_equal(x, y) for x, y in zip(va.outputs, vb.outputs))
    assert va.messages == vb.messages
    assert va.outputs[0].shape == (24, 64, 3)                     # two half-size frames side by side


def test_to_video_frame_wraps_like_numpy():
    import midas_mediapipe
    vs = midas_mediapipe.VidStream(None, None, None, [], None)
    x = np.array([[0.0, 1.9, 51.2, 255.9, 256.0, 300.7, -1.5, 1e4]], dtype=np.float32)
    out = vs.to_video_frame(x)
    with np.errstate(invalid="ignore"):
        ref = x.astype(np.uint8)
    ref = (ref.astype(np.uint32) * 5 + 10).astype(np.uint8)
    assert out.shape == (1, 8, 3) and np.array_equal(out[..., 0], ref) and np.array_equal(out[..., 2], ref)
