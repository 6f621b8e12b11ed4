"""ORACLE (test infrastructure only).  CPU fp32 restatement of intel-isl/MiDaS ``MidasNet_small``
(midas/midas_net_custom.py, midas/blocks.py) over the ``tf_efficientnet_lite3`` backbone of
rwightman/gen-efficientnet-pytorch, as loaded by ``/root/reference/depth_midas.py:10`` for
model_type "MiDaS_small" (hub: features=64, exportable=True, non_negative=True, blocks={'expand':True}).
Neither upstream is vendored - published algorithm restated with upstream key names.
PARITY UNPINNED by the reference itself (it has no tests or fixtures).  Pins that exist: the encoder
equals HF transformers' independent EfficientNet implementation configured into Lite3 bit for bit on all four taps
(tests/test_oracle_lite3_crosscheck.py); the decoder blocks are the oracle/dpt.py code pinned against HF's DPT neck
(tests/test_oracle_hf_crosscheck.py); TF-"same" padding on odd extents is checked by hand in tests/test_oracle_cpu.py.
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from .dpt import FeatureFusionBlock, Scratch

BN_EPS_TF = 1e-3


class Conv2dSame(nn.Conv2d):
    """geffnet conv2d_layers Conv2dSame / Conv2dSameExport: TF 'SAME' padding, extra pixel bottom/right.
    For stride 1 this is the symmetric (k-1)//2 padding geffnet picks statically."""

    def __init__(self, cin, cout, k, stride=1, groups=1):
        super().__init__(cin, cout, k, stride=stride, padding=0, groups=groups, bias=False)

    def forward(self, x):
        k, s = self.kernel_size[0], self.stride[0]
        ih, iw = x.shape[-2:]
        ph = max((math.ceil(ih / s) - 1) * s + (k - 1) + 1 - ih, 0)
        pw = max((math.ceil(iw / s) - 1) * s + (k - 1) + 1 - iw, 0)
        x = F.pad(x, [pw // 2, pw - pw // 2, ph // 2, ph - ph // 2])
        return F.conv2d(x, self.weight, None, self.stride, 0, 1, self.groups)


def _bn(c):
    return nn.BatchNorm2d(c, eps=BN_EPS_TF)


class DepthwiseSeparableConv(nn.Module):
    """geffnet DepthwiseSeparableConv (no SE, pw_act=False, ReLU6)."""

    def __init__(self, cin, cout, k, stride):
        super().__init__()
        self.has_residual = stride == 1 and cin == cout
        self.conv_dw = Conv2dSame(cin, cin, k, stride, groups=cin)
        self.bn1 = _bn(cin)
        self.conv_pw = Conv2dSame(cin, cout, 1)
        self.bn2 = _bn(cout)

    def forward(self, x):
        y = F.relu6(self.bn1(self.conv_dw(x)))
        y = self.bn2(self.conv_pw(y))
        return y + x if self.has_residual else y


class InvertedResidual(nn.Module):
    """geffnet InvertedResidual (exp_ratio 6, no SE, ReLU6)."""

    def __init__(self, cin, cout, k, stride, exp=6):
        super().__init__()
        mid = cin * exp
        self.has_residual = stride == 1 and cin == cout
        self.conv_pw = Conv2dSame(cin, mid, 1)
        self.bn1 = _bn(mid)
        self.conv_dw = Conv2dSame(mid, mid, k, stride, groups=mid)
        self.bn2 = _bn(mid)
        self.conv_pwl = Conv2dSame(mid, cout, 1)
        self.bn3 = _bn(cout)

    def forward(self, x):
        y = F.relu6(self.bn1(self.conv_pw(x)))
        y = F.relu6(self.bn2(self.conv_dw(y)))
        y = self.bn3(self.conv_pwl(y))
        return y + x if self.has_residual else y


# (type, out_ch, repeats, kernel, stride) after channel_multiplier 1.2 / depth_multiplier 1.4 scaling
LITE3_STAGES = (
    ("ds", 24, 1, 3, 1),
    ("ir", 32, 3, 3, 2),
    ("ir", 48, 3, 5, 2),
    ("ir", 96, 5, 3, 2),
    ("ir", 136, 5, 5, 1),
    ("ir", 232, 6, 5, 2),
    ("ir", 384, 1, 3, 1),
)


def _make_blocks():
    blocks, cin = [], 32
    for kind, cout, reps, k, s in LITE3_STAGES:
        stage = []
        for r in range(reps):
            stride = s if r == 0 else 1
            stage.append(DepthwiseSeparableConv(cin, cout, k, stride) if kind == "ds"
                         else InvertedResidual(cin, cout, k, stride))
            cin = cout
        blocks.append(nn.Sequential(*stage))
    return blocks


class _Pretrained(nn.Module):
    def __init__(self):
        super().__init__()
        b = _make_blocks()
        # [UPSTREAM] blocks._make_efficientnet_backbone
        self.layer1 = nn.Sequential(Conv2dSame(3, 32, 3, 2), _bn(32), nn.ReLU6(), *b[0:2])
        self.layer2 = nn.Sequential(*b[2:3])
        self.layer3 = nn.Sequential(*b[3:5])
        self.layer4 = nn.Sequential(*b[5:9])


class MidasNet_small(nn.Module):
    def __init__(self, features=64):
        super().__init__()
        self.pretrained = _Pretrained()
        self.scratch = Scratch([32, 48, 136, 384], features, expand=True)
        f1, f2, f3, f4 = features, features * 2, features * 4, features * 8
        self.scratch.refinenet4 = FeatureFusionBlock(f4, expand=True)
        self.scratch.refinenet3 = FeatureFusionBlock(f3, expand=True)
        self.scratch.refinenet2 = FeatureFusionBlock(f2, expand=True)
        self.scratch.refinenet1 = FeatureFusionBlock(f1, expand=False)
        self.scratch.output_conv = nn.Sequential(
            nn.Conv2d(features, features // 2, 3, 1, 1),
            nn.Identity(),                                    # Interpolate(x2, bilinear, align_corners=False)
            nn.Conv2d(features // 2, 32, 3, 1, 1),
            nn.ReLU(),
            nn.Conv2d(32, 1, 1, 1, 0),
            nn.ReLU(),
            nn.Identity(),
        )

    def encode(self, x):
        p = self.pretrained
        l1 = p.layer1(x)
        l2 = p.layer2(l1)
        l3 = p.layer3(l2)
        l4 = p.layer4(l3)
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
        x = F.interpolate(x, scale_factor=2, mode="bilinear", align_corners=False)
        x = oc[5](oc[4](oc[3](oc[2](x))))
        return torch.squeeze(x, dim=1)

    def forward(self, x):
        return self.decode(*self.encode(x))
