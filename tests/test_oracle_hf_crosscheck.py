"""Parity pin for the DPT models: an INDEPENDENT implementation (HF transformers ``DPTForDepthEstimation``, whose
weights are converted from the same intel-isl/MiDaS checkpoints the reference loads at depth_midas.py:10) is given
seeded random weights; those weights are renamed to upstream MiDaS keys (tests/hf_crosscheck.py) and fed to

  * the CPU oracle (``not gpu`` tests, fp32 vs fp32: agreement to rounding), and
  * the CUDA product path through ``DepthEstimator(weights=...)`` (``gpu`` tests, north_star tolerance 1e-3),

so neither the oracle nor the kernels are only checked against a restatement written in this repo.
HF's reassemble stage assumes a square token grid (and its hybrid embedder a fixed image size), hence square
frames only; non-square geometry is covered oracle-vs-product in tests/test_e2e_gpu.py.
"""
import numpy as np
import pytest
import torch

transformers = pytest.importorskip("transformers")

from tests.hf_crosscheck import (bf16_operands, complete_for_product, hf_hybrid_config, hf_large_config, hf_to_upstream,
                                 load_into_oracle, randomize_)


def _hf(model_type, seed):
    from transformers import DPTForDepthEstimation
    hybrid = model_type == "DPT_Hybrid"
    cfg = hf_hybrid_config() if hybrid else hf_large_config()
    torch.manual_seed(seed)
    hf = randomize_(DPTForDepthEstimation(cfg).eval(), seed)
    return hf, hf_to_upstream(hf.state_dict(), hybrid=hybrid)


class _HFNet(torch.nn.Module):
    """gives the HF model the call signature of the hub model (tensor in, [B,h,w] out)."""

    def __init__(self, hf):
        super().__init__()
        self.hf = hf

    def forward(self, x):
        return self.hf(pixel_values=x).predicted_depth


@pytest.mark.parametrize("model_type,side,seed", [("DPT_Large", 384, 0), ("DPT_Large", 320, 3), ("DPT_Hybrid", 384, 1)])
def test_oracle_matches_hf(model_type, side, seed):
    from oracle.dpt import DPTDepthModel
    hf, up = _hf(model_type, seed)
    orc = DPTDepthModel("vitl16_384" if model_type == "DPT_Large" else "vitb_rn50_384").eval()
    extra = load_into_oracle(orc, up)
    # the only tensors the oracle has and HF lacks: timm's unused 1000-way classifier
    assert sorted(extra) == ["pretrained.model.head.bias", "pretrained.model.head.weight"]
    x = torch.randn(1, 3, side, side, generator=torch.Generator().manual_seed(seed + 100))
    with torch.no_grad():
        want = hf(pixel_values=x).predicted_depth
        got = orc(x)
    assert got.shape == want.shape == (1, side, side)
    assert (want > 0).float().mean() > 0.9 and want.std() > 1e-2          # not a degenerate map
    err = float((got - want).abs().max() / (want.max() - want.min()))
    print(f"{model_type} {side}: oracle vs HF, max err / dynamic range = {err:.2e}")
    assert err < 2e-5                                                      # fp32 vs fp32, different op order


def test_hf_key_inventory_is_the_upstream_one():
    """The renamed HF state_dict plus the three tensor groups HF's converter drops is exactly the key/shape inventory
    the product's checkpoint loader enforces for a real upstream checkpoint."""
    from depth_estimation_b200.init import param_shapes
    from transformers import DPTForDepthEstimation
    for mt, cfg in (("DPT_Large", hf_large_config()), ("DPT_Hybrid", hf_hybrid_config())):
        with torch.device("meta"):
            hf = DPTForDepthEstimation(cfg)
        shapes = {k: tuple(v.shape) for k, v in hf.state_dict().items()}
        fake = {k: torch.empty(s, device="meta") for k, s in shapes.items()}
        up = {k: tuple(v.shape) for k, v in hf_to_upstream(fake, hybrid=mt == "DPT_Hybrid").items()}
        want = {k: tuple(s) for k, s in param_shapes(mt).items()}
        assert all(want[k] == s for k, s in up.items())
        dropped = sorted(set(want) - set(up))
        assert all(k.startswith(("pretrained.model.head.", "scratch.refinenet4.resConfUnit1.")) for k in dropped), dropped


@pytest.mark.gpu
@pytest.mark.parametrize("model_type,side,seed", [("DPT_Large", 384, 0), ("DPT_Large", 320, 3), ("DPT_Hybrid", 384, 1)])
def test_product_matches_hf(model_type, side, seed):
    """The CUDA path against the independent implementation, whole ``predict`` (transform -> net -> bicubic).

    These maps have a wide dynamic range (max / (max - min) ~ 1.4; the maps of ``init.random_state_dict`` that the
    other e2e cases use sit at ~10, which flatters any max-normalised error).  On them the arithmetic contract
    itself - bf16 operands, fp32 accumulation - costs 6e-3 of the maximum: the HF model run with bf16-rounded
    operands differs from its own fp32 run by that much (``bf16_operands``).  north_star's 1e-3 therefore cannot be
    demanded here; the assertion is that the CUDA path is no further from the fp32 reference than 1.5x that
    precision floor (an indexing / layout / fusion bug shows up as 1e-1 .. 1e0)."""
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from depth_estimation_b200.depth_midas import DepthEstimator
    from oracle.predict import OracleDepthEstimator, synthetic_frame
    hf, up = _hf(model_type, seed)
    est = DepthEstimator(model_type, weights=complete_for_product(up, model_type))
    frame = synthetic_frame(side, side, 7, "S")
    ref = OracleDepthEstimator(model_type, model=_HFNet(hf)).predict(frame)      # reference predict body around HF
    with bf16_operands(hf):
        emu = OracleDepthEstimator(model_type, model=_HFNet(hf)).predict(frame)
    out = est.predict(frame)
    assert out.shape == ref.shape == (side, side)
    assert (ref > 0).mean() > 0.9
    scale = float(np.abs(ref).max())
    rel, floor = float(np.abs(out - ref).max() / scale), float(np.abs(emu - ref).max() / scale)
    dyn = float(np.abs(out - ref).max() / (ref.max() - ref.min()))
    print(f"{model_type} {side}: CUDA vs HF rel={rel:.3e} dyn={dyn:.3e}; bf16-operand floor of HF itself rel={floor:.3e}")
    assert rel < max(1e-3, 1.5 * floor), (rel, floor, dyn)
    assert rel < 2e-2
    # precision="bf16x3": every Linear / convolution on hi+lo bf16 operand pairs - north_star's 1e-3 on these wide-range
    # maps too (max-normalised and per pixel), which no plain-bf16-operand pipeline can reach (the floor above)
    est3 = DepthEstimator(model_type, weights=complete_for_product(up, model_type), precision="bf16x3")
    out3 = est3.predict(frame)
    d = np.abs(out3 - ref)
    rel3 = float(d.max() / scale)
    m = ref > 1e-2 * ref.max()
    pix3 = float((d[m] / ref[m]).max())
    print(f"{model_type} {side}: bf16x3 vs HF rel={rel3:.3e} pix={pix3:.3e}")
    assert rel3 < 1e-3 and rel3 < rel / 3, (rel3, rel)
