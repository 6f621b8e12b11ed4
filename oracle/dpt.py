"""ORACLE (test infrastructure only).  CPU fp32 restatement of intel-isl/MiDaS v3.0
``DPTDepthModel`` (midas/dpt_depth.py, midas/vit.py, midas/blocks.py) as loaded by
``/root/reference/depth_midas.py:10`` for model_type "DPT_Large" and "DPT_Hybrid".  Upstream source is
not vendored - published algorithm restated, upstream ``state_dict`` key names kept.
PARITY UNPINNED by the reference (it has no tests); the whole model (ViT-L / ResNetV2+ViT-B, reassemble, fusion,
head) is cross-checked against the independently written HF transformers ``DPTForDepthEstimation`` with
identical weights in tests/test_oracle_hf_crosscheck.py (agreement 1e-6 .. 4e-6 of the dynamic range).
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from .vit import VisionTransformer


class ResidualConvUnit(nn.Module):
    """[UPSTREAM] blocks.ResidualConvUnit_custom, bn=False, activation = nn.ReLU(False):
    out = conv2(relu(conv1(relu(x)))) + x   (skip adds the UN-activated x)."""

    def __init__(self, features):
        super().__init__()
        self.conv1 = nn.Conv2d(features, features, 3, 1, 1, bias=True)
        self.conv2 = nn.Conv2d(features, features, 3, 1, 1, bias=True)

    def forward(self, x):
        out = self.conv1(F.relu(x))
        out = self.conv2(F.relu(out))
        return out + x


class FeatureFusionBlock(nn.Module):
    """[UPSTREAM] blocks.FeatureFusionBlock_custom (deconv=False, bn=False)."""

    def __init__(self, features, expand=False, align_corners=True):
        super().__init__()
        self.align_corners = align_corners
        self.out_conv = nn.Conv2d(features, features // 2 if expand else features, 1, 1, 0, bias=True)
        self.resConfUnit1 = ResidualConvUnit(features)
        self.resConfUnit2 = ResidualConvUnit(features)

    def forward(self, *xs):
        output = xs[0]
        if len(xs) == 2:
            output = output + self.resConfUnit1(xs[1])
        output = self.resConfUnit2(output)
        output = F.interpolate(output, scale_factor=2, mode="bilinear", align_corners=self.align_corners)
        return self.out_conv(output)


class Scratch(nn.Module):
    """[UPSTREAM] blocks._make_scratch + the fusion blocks / head that DPT attaches to it."""

    def __init__(self, in_shape, features, expand=False):
        super().__init__()
        o = [features, features * 2, features * 4, features * 8] if expand else [features] * 4
        self.layer1_rn = nn.Conv2d(in_shape[0], o[0], 3, 1, 1, bias=False)
        self.layer2_rn = nn.Conv2d(in_shape[1], o[1], 3, 1, 1, bias=False)
        self.layer3_rn = nn.Conv2d(in_shape[2], o[2], 3, 1, 1, bias=False)
        self.layer4_rn = nn.Conv2d(in_shape[3], o[3], 3, 1, 1, bias=False)


class ProjectReadout(nn.Module):
    """[UPSTREAM] vit.ProjectReadout: cat(tokens, cls broadcast) -> Linear(2D,D) -> GELU."""

    def __init__(self, dim):
        super().__init__()
        self.project = nn.Sequential(nn.Linear(2 * dim, dim), nn.GELU())

    def forward(self, x):
        readout = x[:, 0].unsqueeze(1).expand_as(x[:, 1:])
        return self.project(torch.cat((x[:, 1:], readout), -1))


class _Pretrained(nn.Module):
    pass


def _act_postprocess(dim, cout, kind):
    """Sequential index layout matches upstream so keys are act_postprocessN.{0.project.0,3,4}."""
    mods = [ProjectReadout(dim), nn.Identity(), nn.Identity(), nn.Conv2d(dim, cout, 1, 1, 0)]
    if kind == "up4":
        mods.append(nn.ConvTranspose2d(cout, cout, 4, 4, 0, bias=True))
    elif kind == "up2":
        mods.append(nn.ConvTranspose2d(cout, cout, 2, 2, 0, bias=True))
    elif kind == "down2":
        mods.append(nn.Conv2d(cout, cout, 3, 2, 1))
    return nn.Sequential(*mods)


class DPTDepthModel(nn.Module):
    """backbone in {"vitl16_384" (DPT_Large), "vitb_rn50_384" (DPT_Hybrid)}, non_negative=True."""

    def __init__(self, backbone="vitl16_384", features=256):
        super().__init__()
        self.backbone = backbone
        self.pretrained = _Pretrained()
        if backbone == "vitl16_384":
            dim, self.hooks, feats = 1024, (5, 11, 17, 23), (256, 512, 1024, 1024)
            self.pretrained.model = VisionTransformer(1024, 24, 16, hybrid=False)
            self.pretrained.act_postprocess1 = _act_postprocess(dim, feats[0], "up4")
            self.pretrained.act_postprocess2 = _act_postprocess(dim, feats[1], "up2")
        elif backbone == "vitb_rn50_384":
            dim, self.hooks, feats = 768, (8, 11), (256, 512, 768, 768)
            self.pretrained.model = VisionTransformer(768, 12, 12, hybrid=True)
            self.pretrained.act_postprocess1 = nn.Sequential(nn.Identity(), nn.Identity(), nn.Identity())
            self.pretrained.act_postprocess2 = nn.Sequential(nn.Identity(), nn.Identity(), nn.Identity())
        else:
            raise ValueError(backbone)
        self.pretrained.act_postprocess3 = _act_postprocess(dim, feats[2], "same")
        self.pretrained.act_postprocess4 = _act_postprocess(dim, feats[3], "down2")
        self.scratch = Scratch(feats, features, expand=False)
        for i in (1, 2, 3, 4):
            setattr(self.scratch, f"refinenet{i}", FeatureFusionBlock(features))
        self.scratch.output_conv = nn.Sequential(
            nn.Conv2d(features, features // 2, 3, 1, 1),
            nn.Identity(),                                    # Interpolate(x2, bilinear, align_corners=True)
            nn.Conv2d(features // 2, 32, 3, 1, 1),
            nn.ReLU(),
            nn.Conv2d(32, 1, 1, 1, 0),
            nn.ReLU(),
            nn.Identity(),
        )

    def _reassemble(self, seq, tokens, gh, gw):
        x = seq[0](tokens)                                    # readout project
        x = x.transpose(1, 2).unflatten(2, (gh, gw))
        for m in list(seq)[3:]:
            x = m(x)
        return x

    def forward_vit(self, x):
        B, _, h, w = x.shape
        gh, gw = h // 16, w // 16
        acts, stage = self.pretrained.model.forward_flex(x, set(self.hooks))
        p = self.pretrained
        if self.backbone == "vitl16_384":
            l1 = self._reassemble(p.act_postprocess1, acts[self.hooks[0]], gh, gw)
            l2 = self._reassemble(p.act_postprocess2, acts[self.hooks[1]], gh, gw)
            l3 = self._reassemble(p.act_postprocess3, acts[self.hooks[2]], gh, gw)
            l4 = self._reassemble(p.act_postprocess4, acts[self.hooks[3]], gh, gw)
        else:
            l1, l2 = stage
            l3 = self._reassemble(p.act_postprocess3, acts[self.hooks[0]], gh, gw)
            l4 = self._reassemble(p.act_postprocess4, acts[self.hooks[1]], gh, gw)
        return l1, l2, l3, l4

    def decode(self, l1, l2, l3, l4):
        s = self.scratch
        l1, l2, l3, l4 = s.layer1_rn(l1), s.layer2_rn(l2), s.layer3_rn(l3), s.layer4_rn(l4)
        p4 = s.refinenet4(l4)
        p3 = s.refinenet3(p4, l3)
        p2 = s.refinenet2(p3, l2)
        p1 = s.refinenet1(p2, l1)
        oc = s.output_conv
        x = oc[0](p1)
        x = F.interpolate(x, scale_factor=2, mode="bilinear", align_corners=True)
        x = oc[5](oc[4](oc[3](oc[2](x))))
        return x.squeeze(dim=1)

    def forward(self, x):
        return self.decode(*self.forward_vit(x))


def DPT_Large():
    return DPTDepthModel("vitl16_384")


def DPT_Hybrid():
    return DPTDepthModel("vitb_rn50_384")
