"""End-to-end GPU parity: DepthEstimator.predict (CUDA path, through the C ABI) vs the CPU oracle with the
same random-init weights and the same synthetic frame.

Tolerances (north_star): fp32 oracle vs bf16-operand/fp32-accumulate kernels.  `rel` is
max|y - y*| / max(y*); `dyn` is the same error against the dynamic range max(y*)-min(y*), the stricter
figure because the degeneracy-guard bias (+1.0) lifts every pixel."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

CASES = [
    ("DPT_Large", 384, 384, "S"),
    ("DPT_Large", 640, 480, "U"),
    ("MiDaS_small", 640, 480, "S"),
    ("MiDaS_small", 1280, 720, "U"),
    ("DPT_Hybrid", 384, 384, "S"),
]


def _errs(y, ref):
    rel = float(np.abs(y - ref).max() / np.abs(ref).max())
    dyn = float(np.abs(y - ref).max() / (ref.max() - ref.min()))
    return rel, dyn


@pytest.mark.parametrize("mt,W,H,kind", CASES)
def test_predict_matches_oracle(mt, W, H, kind):
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from depth_estimation_b200.depth_midas import DepthEstimator
    from oracle.predict import OracleDepthEstimator, build_model, synthetic_frame
    try:
        model = build_model(mt, seed=0)
        est = DepthEstimator(mt, weights=model.state_dict())
    except ImportError as e:                       # model family not built yet
        pytest.skip(str(e))
    frame = synthetic_frame(W, H, 3, kind)
    ref = OracleDepthEstimator(mt, model=model).predict(frame)
    out = est.predict(frame)
    assert out.shape == ref.shape == (H, W) and out.dtype == np.float32
    assert (ref > 0).mean() > 0.9
    rel, dyn = _errs(out, ref)
    print(f"{mt} {W}x{H} {kind}: rel={rel:.3e} dyn={dyn:.3e}")
    assert rel < 1e-2, (rel, dyn)                  # bf16 metric-depth budget (1 %)
    # graph replay gives the same answer, and the batched / metric / sample outputs agree
    out2 = est.predict(frame)
    assert np.array_equal(out, out2)
    rc = np.array([[[H // 2, W // 3], [H - 1, W - 1]]], dtype=np.int32)
    maps, samples, metric = est.predict_batch(frame[None], samples_rc=rc, want_metric=True)
    assert np.array_equal(maps[0], out)
    assert samples[0, 0] == out[H // 2, W // 3] and samples[0, 1] == out[H - 1, W - 1]
    assert np.abs(metric[0] - ref / 2.54).max() / np.abs(ref / 2.54).max() < 1e-2
