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
