"""Parameter inventories (upstream ``state_dict`` key -> shape) of the three MiDaS hub models and a
seeded random initialiser.  The reference obtains real weights from torch.hub at
/root/reference/depth_midas.py:10; there is no network here, so benchmarks use random-init weights of the
exact architecture (BASELINE.json: "random-init").  Real checkpoints load through the same key names.
"""
from __future__ import annotations

import math

import torch

LITE3_STAGES = (("ds", 24, 1, 3, 1), ("ir", 32, 3, 3, 2), ("ir", 48, 3, 5, 2), ("ir", 96, 5, 3, 2),
                ("ir", 136, 5, 5, 1), ("ir", 232, 6, 5, 2), ("ir", 384, 1, 3, 1))


def _vit_shapes(prefix, dim, depth, grid=24):
    s = {prefix + "cls_token": (1, 1, dim), prefix + "pos_embed": (1, grid * grid + 1, dim),
         prefix + "norm.weight": (dim,), prefix + "norm.bias": (dim,),
         prefix + "head.weight": (1000, dim), prefix + "head.bias": (1000,)}
    for i in range(depth):
        p = f"{prefix}blocks.{i}."
        s.update({p + "norm1.weight": (dim,), p + "norm1.bias": (dim,), p + "attn.qkv.weight": (3 * dim, dim),
                  p + "attn.qkv.bias": (3 * dim,), p + "attn.proj.weight": (dim, dim), p + "attn.proj.bias": (dim,),
                  p + "norm2.weight": (dim,), p + "norm2.bias": (dim,), p + "mlp.fc1.weight": (4 * dim, dim),
                  p + "mlp.fc1.bias": (4 * dim,), p + "mlp.fc2.weight": (dim, 4 * dim), p + "mlp.fc2.bias": (dim,)})
    return s


def _scratch_shapes(in_shape, feats, head_mid):
    s = {}
    for i in range(4):
        s[f"scratch.layer{i + 1}_rn.weight"] = (feats[i], in_shape[i], 3, 3)
        f = feats[i]
        p = f"scratch.refinenet{i + 1}."
        for u in ("resConfUnit1", "resConfUnit2"):
            for c in ("conv1", "conv2"):
                s[p + f"{u}.{c}.weight"] = (f, f, 3, 3)
                s[p + f"{u}.{c}.bias"] = (f,)
        fo = feats[i - 1] if i > 0 else feats[0]
        s[p + "out_conv.weight"] = (fo, f, 1, 1)
        s[p + "out_conv.bias"] = (fo,)
    f1 = feats[0]
    s.update({"scratch.output_conv.0.weight": (head_mid, f1, 3, 3), "scratch.output_conv.0.bias": (head_mid,),
              "scratch.output_conv.2.weight": (32, head_mid, 3, 3), "scratch.output_conv.2.bias": (32,),
              "scratch.output_conv.4.weight": (1, 32, 1, 1), "scratch.output_conv.4.bias": (1,)})
    return s


def _readout_shapes(idx, dim, cout, kind):
    p = f"pretrained.act_postprocess{idx}."
    s = {p + "0.project.0.weight": (dim, 2 * dim), p + "0.project.0.bias": (dim,),
         p + "3.weight": (cout, dim, 1, 1), p + "3.bias": (cout,)}
    if kind == "up4":
        s.update({p + "4.weight": (cout, cout, 4, 4), p + "4.bias": (cout,)})
    elif kind == "up2":
        s.update({p + "4.weight": (cout, cout, 2, 2), p + "4.bias": (cout,)})
    elif kind == "down2":
        s.update({p + "4.weight": (cout, cout, 3, 3), p + "4.bias": (cout,)})
    return s


def _resnetv2_shapes(prefix):
    s = {prefix + "stem.conv.weight": (64, 3, 7, 7), prefix + "stem.norm.weight": (64,), prefix + "stem.norm.bias": (64,)}
    cin = 64
    for si, (cout, depth) in enumerate(((256, 3), (512, 4), (1024, 9))):
        mid = cout // 4
        for b in range(depth):
            p = f"{prefix}stages.{si}.blocks.{b}."
            bi = cin if b == 0 else cout
            if b == 0:
                s.update({p + "downsample.conv.weight": (cout, bi, 1, 1), p + "downsample.norm.weight": (cout,),
                          p + "downsample.norm.bias": (cout,)})
            s.update({p + "conv1.weight": (mid, bi, 1, 1), p + "norm1.weight": (mid,), p + "norm1.bias": (mid,),
                      p + "conv2.weight": (mid, mid, 3, 3), p + "norm2.weight": (mid,), p + "norm2.bias": (mid,),
                      p + "conv3.weight": (cout, mid, 1, 1), p + "norm3.weight": (cout,), p + "norm3.bias": (cout,)})
        cin = cout
    return s


def _bn(s, p, c):
    s.update({p + "weight": (c,), p + "bias": (c,), p + "running_mean": (c,), p + "running_var": (c,)})


def _lite3_shapes():
    """[UPSTREAM] blocks._make_efficientnet_backbone key layout over tf_efficientnet_lite3."""
    s = {"pretrained.layer1.0.weight": (32, 3, 3, 3)}
    _bn(s, "pretrained.layer1.1.", 32)
    # (layer name, index inside the layer Sequential) for upstream stages 0..6
    where = [("layer1", 3), ("layer1", 4), ("layer2", 0), ("layer3", 0), ("layer3", 1), ("layer4", 0), ("layer4", 1)]
    cin = 32
    for (kind, cout, reps, k, _), (layer, li) in zip(LITE3_STAGES, where):
        for r in range(reps):
            p = f"pretrained.{layer}.{li}.{r}."
            if kind == "ds":
                s[p + "conv_dw.weight"] = (cin, 1, k, k)
                _bn(s, p + "bn1.", cin)
                s[p + "conv_pw.weight"] = (cout, cin, 1, 1)
                _bn(s, p + "bn2.", cout)
            else:
                mid = cin * 6
                s[p + "conv_pw.weight"] = (mid, cin, 1, 1)
                _bn(s, p + "bn1.", mid)
                s[p + "conv_dw.weight"] = (mid, 1, k, k)
                _bn(s, p + "bn2.", mid)
                s[p + "conv_pwl.weight"] = (cout, mid, 1, 1)
                _bn(s, p + "bn3.", cout)
            cin = cout
    return s


def param_shapes(model_type: str) -> dict:
    if model_type == "DPT_Large":
        s = _vit_shapes("pretrained.model.", 1024, 24)
        s.update({"pretrained.model.patch_embed.proj.weight": (1024, 3, 16, 16),
                  "pretrained.model.patch_embed.proj.bias": (1024,)})
        for i, (c, k) in enumerate(((256, "up4"), (512, "up2"), (1024, "same"), (1024, "down2"))):
            s.update(_readout_shapes(i + 1, 1024, c, k))
        s.update(_scratch_shapes((256, 512, 1024, 1024), (256,) * 4, 128))
        return s
    if model_type == "DPT_Hybrid":
        s = _vit_shapes("pretrained.model.", 768, 12)
        s.update(_resnetv2_shapes("pretrained.model.patch_embed.backbone."))
        s.update({"pretrained.model.patch_embed.proj.weight": (768, 1024, 1, 1),
                  "pretrained.model.patch_embed.proj.bias": (768,)})
        s.update(_readout_shapes(3, 768, 768, "same"))
        s.update(_readout_shapes(4, 768, 768, "down2"))
        s.update(_scratch_shapes((256, 512, 768, 768), (256,) * 4, 128))
        return s
    if model_type == "MiDaS_small":
        s = _lite3_shapes()
        s.update(_scratch_shapes((32, 48, 136, 384), (64, 128, 256, 512), 32))
        return s
    raise ValueError(f"unknown model_type {model_type!r}")


def random_state_dict(model_type: str, seed: int = 0) -> dict:
    """Seeded random-init weights with upstream key names (fp32, CPU)."""
    g = torch.Generator().manual_seed(seed)
    sd = {}
    for k, shape in param_shapes(model_type).items():
        if k.endswith("running_var"):
            t = torch.rand(shape, generator=g) + 0.5
        elif k.endswith("running_mean"):
            t = torch.randn(shape, generator=g) * 0.1
        elif ".norm" in k or ".bn" in k or k.startswith("pretrained.layer1.1."):
            t = (1.0 + 0.1 * torch.randn(shape, generator=g)) if k.endswith("weight") else 0.1 * torch.randn(shape, generator=g)
        elif k.endswith("cls_token"):
            t = torch.randn(shape, generator=g) * 1e-6
        elif k.endswith("pos_embed"):
            t = torch.randn(shape, generator=g) * 0.02
        elif k.endswith("bias"):
            t = torch.randn(shape, generator=g) * 0.02
        elif len(shape) == 2:
            t = torch.randn(shape, generator=g) * 0.02
        else:
            fan_in = shape[1] * shape[2] * shape[3]
            t = torch.randn(shape, generator=g) * (1.0 / math.sqrt(fan_in))
        sd[k] = t
    sd["scratch.output_conv.4.bias"] = torch.ones(1)
    sd["scratch.output_conv.2.bias"] = torch.full((32,), 0.5)
    return sd
