"""Parity pin for the MiDaS_small encoder: an INDEPENDENT implementation of EfficientNet (HF ``transformers``
``EfficientNetModel``, written against the TF/Keras original) is configured into tf_efficientnet_lite3 and given the
oracle's seeded weights; its four encoder taps must equal ``oracle/midas_small.py::MidasNet_small.encode`` - the hub
backbone behind /root/reference/depth_midas.py:10,26 for model_type "MiDaS_small".

HF's model is the non-Lite EfficientNet (swish, squeeze-excite, Keras `correct_pad`).  Configured into Lite3:
  * hidden_act "relu6", width/depth coefficients 1.0 with the already-scaled Lite3 channel / repeat lists
    (channel_multiplier 1.2, depth_multiplier 1.4 applied, first and last stage repeats not scaled - the Lite rule),
    batch_norm_eps 1e-3, depthwise_padding [] (every stride-2 depthwise gets Keras' asymmetric pad);
  * squeeze-excite neutralised by weights: expand conv weight 0, bias +30 => sigmoid == 1.0f exactly in fp32.
Keras `correct_pad` (k//2-1 before, k//2 after) equals TF-"same" on even extents, which every BASELINE input has at
every stride-2 level (192x256 and 128x256 are multiples of 32); the oracle's general TF-same rule on odd extents is
covered by the hand-computed padding test in tests/test_oracle_cpu.py.

The decoder of MiDaS_small is the FeatureFusionBlock / RCU code of oracle/dpt.py, which the DPT cross-check
(tests/test_oracle_hf_crosscheck.py) pins against HF's independent DPT neck.
"""
import pytest
import torch

transformers = pytest.importorskip("transformers")

STAGE_REPEATS = (1, 3, 3, 5, 5, 6, 1)
# hidden_states index (0 = stem output) at which upstream layer1..4 end: stem+stage0+stage1 | stage2 | stage3+4 | stage5+6
TAPS = (4, 7, 17, 24)


def lite3_config():
    from transformers import EfficientNetConfig
    return EfficientNetConfig(
        num_channels=3, image_size=256, width_coefficient=1.0, depth_coefficient=1.0, depth_divisor=8,
        kernel_sizes=[3, 3, 5, 3, 5, 5, 3], in_channels=[32, 24, 32, 48, 96, 136, 232],
        out_channels=[24, 32, 48, 96, 136, 232, 384], depthwise_padding=[], strides=[1, 2, 2, 2, 1, 2, 1],
        num_block_repeats=list(STAGE_REPEATS), expand_ratios=[1, 6, 6, 6, 6, 6, 6], squeeze_expansion_ratio=0.25,
        hidden_act="relu6", hidden_dim=1280, pooling_type="mean", batch_norm_eps=1e-3, drop_connect_rate=0.0)


def upstream_to_hf(sd: dict) -> dict:
    """MiDaS_small `pretrained.*` keys (blocks._make_efficientnet_backbone naming) -> HF EfficientNetModel keys."""
    where = [("layer1", 3), ("layer1", 4), ("layer2", 0), ("layer3", 0), ("layer3", 1), ("layer4", 0), ("layer4", 1)]
    out = {"embeddings.convolution.weight": sd["pretrained.layer1.0.weight"]}
    bn_keys = ("weight", "bias", "running_mean", "running_var")
    for k in bn_keys:
        out["embeddings.batchnorm." + k] = sd["pretrained.layer1.1." + k]
    n = 0
    for si, ((layer, li), reps) in enumerate(zip(where, STAGE_REPEATS)):
        for r in range(reps):
            p, h = f"pretrained.{layer}.{li}.{r}.", f"encoder.blocks.{n}."
            if si == 0:                                                 # DepthwiseSeparableConv: dw, bn1, pw, bn2
                names = {"depthwise_conv.depthwise_conv": "conv_dw", "depthwise_conv.depthwise_norm": "bn1",
                         "projection.project_conv": "conv_pw", "projection.project_bn": "bn2"}
            else:                                                       # InvertedResidual
                names = {"expansion.expand_conv": "conv_pw", "expansion.expand_bn": "bn1",
                         "depthwise_conv.depthwise_conv": "conv_dw", "depthwise_conv.depthwise_norm": "bn2",
                         "projection.project_conv": "conv_pwl", "projection.project_bn": "bn3"}
            for hn, un in names.items():
                if un.startswith("bn"):
                    for k in bn_keys:
                        out[h + hn + "." + k] = sd[p + un + "." + k]
                else:
                    out[h + hn + ".weight"] = sd[p + un + ".weight"]
            n += 1
    return out


def build_hf_lite3(oracle_model):
    from transformers import EfficientNetModel
    hf = EfficientNetModel(lite3_config()).eval()
    mapped = upstream_to_hf(oracle_model.state_dict())
    own = hf.state_dict()
    for k, v in mapped.items():
        assert k in own and own[k].shape == v.shape, (k, tuple(v.shape), tuple(own[k].shape) if k in own else None)
    missing, unexpected = hf.load_state_dict(mapped, strict=False)
    assert not unexpected
    # everything not supplied must be the SE layers, the unused top conv/bn, or BN step counters
    assert all(("squeeze_excite" in k) or k.startswith("encoder.top_") or k.endswith("num_batches_tracked")
               for k in missing), [k for k in missing if "squeeze_excite" not in k][:5]
    with torch.no_grad():
        for name, p in hf.named_parameters():
            if "squeeze_excite.expand.weight" in name:
                p.zero_()
            elif "squeeze_excite.expand.bias" in name:
                p.fill_(30.0)                                           # sigmoid(30) rounds to exactly 1.0f
    return hf


def test_every_backbone_tensor_is_consumed():
    """The key map covers the whole `pretrained.*` inventory of the oracle (nothing silently left at its init)."""
    from oracle.predict import build_model
    m = build_model("MiDaS_small", seed=3)
    have = {k for k in m.state_dict() if k.startswith("pretrained.") and not k.endswith("num_batches_tracked")}
    used = set()

    class Rec(dict):
        def __getitem__(self, k):
            used.add(k)
            return dict.__getitem__(self, k)
    upstream_to_hf(Rec(m.state_dict()))
    assert used == have, sorted(have - used)[:5]


@pytest.mark.parametrize("h,w,seed", [(192, 256, 0), (128, 256, 1), (256, 256, 2)])
def test_oracle_encoder_matches_hf_efficientnet(h, w, seed):
    from oracle.predict import build_model
    m = build_model("MiDaS_small", seed=seed)
    hf = build_hf_lite3(m)
    x = torch.randn(2, 3, h, w, generator=torch.Generator().manual_seed(10 + seed))
    with torch.no_grad():
        want = hf(pixel_values=x, output_hidden_states=True).hidden_states
        got = m.encode(x)
    assert len(want) == 1 + sum(STAGE_REPEATS)
    for lvl, (g, t) in enumerate(zip(got, TAPS)):
        ref = want[t]
        assert g.shape == ref.shape == (2, (32, 48, 136, 384)[lvl], h >> (lvl + 2), w >> (lvl + 2))
        assert float(ref.abs().max()) > 1e-2                          # not a dead feature map
        err = float((g - ref).abs().max() / ref.abs().max())
        print(f"layer{lvl + 1} {tuple(g.shape)}: oracle vs HF EfficientNet(lite3) max err / max = {err:.2e}")
        assert err < 1e-5
