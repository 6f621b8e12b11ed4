"""ORACLE (test infrastructure only).  CPU fp32 restatement of the timm 0.6.7 VisionTransformer and
ResNetV2 pieces that ``torch.hub.load("intel-isl/MiDaS", "DPT_Large"|"DPT_Hybrid")`` builds
(reference call site ``/root/reference/depth_midas.py:10``; timm pinned by ``midas.yaml:116``).
timm is not installed here and the hub repo is not vendored - this is a
restatement of the published algorithm with upstream ``state_dict`` key names.  PARITY UNPINNED by the
reference (no tests); cross-checked against HF ``transformers`` DPT in tests/test_oracle_hf_crosscheck.py.
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn
import torch.nn.functional as F


class Attention(nn.Module):
    """timm vision_transformer.Attention (qkv_bias=True)."""

    def __init__(self, dim, num_heads):
        super().__init__()
        self.num_heads = num_heads
        self.scale = (dim // num_heads) ** -0.5
        self.qkv = nn.Linear(dim, dim * 3, bias=True)
        self.proj = nn.Linear(dim, dim)

    def forward(self, x):
        B, N, C = x.shape
        qkv = self.qkv(x).reshape(B, N, 3, self.num_heads, C // self.num_heads).permute(2, 0, 3, 1, 4)
        q, k, v = qkv.unbind(0)
        attn = (q @ k.transpose(-2, -1)) * self.scale
        attn = attn.softmax(dim=-1)
        x = (attn @ v).transpose(1, 2).reshape(B, N, C)
        return self.proj(x)


class Mlp(nn.Module):
    def __init__(self, dim, hidden):
        super().__init__()
        self.fc1 = nn.Linear(dim, hidden)
        self.act = nn.GELU()              # exact erf GELU
        self.fc2 = nn.Linear(hidden, dim)

    def forward(self, x):
        return self.fc2(self.act(self.fc1(x)))


class Block(nn.Module):
    """Pre-norm block, LayerNorm eps 1e-6 (timm partial(nn.LayerNorm, eps=1e-6))."""

    def __init__(self, dim, num_heads, mlp_ratio=4.0):
        super().__init__()
        self.norm1 = nn.LayerNorm(dim, eps=1e-6)
        self.attn = Attention(dim, num_heads)
        self.norm2 = nn.LayerNorm(dim, eps=1e-6)
        self.mlp = Mlp(dim, int(dim * mlp_ratio))

    def forward(self, x):
        x = x + self.attn(self.norm1(x))
        x = x + self.mlp(self.norm2(x))
        return x


class PatchEmbed(nn.Module):
    def __init__(self, patch=16, in_chans=3, dim=768):
        super().__init__()
        self.proj = nn.Conv2d(in_chans, dim, kernel_size=patch, stride=patch)


# ------------------------------- ResNetV2 (hybrid backbone) -------------------------------------
def _pad_same(x, k, s, value=0.0):
    """timm layers/padding.py pad_same: TF 'SAME' - extra pixel goes bottom/right."""
    ih, iw = x.shape[-2:]
    ph = max((math.ceil(ih / s) - 1) * s + (k - 1) + 1 - ih, 0)
    pw = max((math.ceil(iw / s) - 1) * s + (k - 1) + 1 - iw, 0)
    if ph > 0 or pw > 0:
        x = F.pad(x, [pw // 2, pw - pw // 2, ph // 2, ph - ph // 2], value=value)
    return x


class StdConv2dSame(nn.Conv2d):
    """timm layers/std_conv.py StdConv2dSame: weight standardisation (biased var, eps) + TF-same pad."""

    def __init__(self, cin, cout, k, stride=1, eps=1e-8):
        super().__init__(cin, cout, k, stride=stride, padding=0, bias=False)
        self.eps = eps

    def std_weight(self):
        w = self.weight
        return F.batch_norm(w.reshape(1, self.out_channels, -1), None, None, training=True, momentum=0.0,
                            eps=self.eps).reshape_as(w)

    def forward(self, x):
        x = _pad_same(x, self.kernel_size[0], self.stride[0])
        return F.conv2d(x, self.std_weight(), None, self.stride, 0)


class GroupNormAct(nn.GroupNorm):
    def __init__(self, ch, apply_act=True):
        super().__init__(32, ch, eps=1e-5)
        self.apply_act = apply_act

    def forward(self, x):
        x = F.group_norm(x, self.num_groups, self.weight, self.bias, self.eps)
        return F.relu(x) if self.apply_act else x


class DownsampleConv(nn.Module):
    def __init__(self, cin, cout, stride):
        super().__init__()
        self.conv = StdConv2dSame(cin, cout, 1, stride=stride)
        self.norm = GroupNormAct(cout, apply_act=False)

    def forward(self, x):
        return self.norm(self.conv(x))


class Bottleneck(nn.Module):
    """timm resnetv2.Bottleneck (non-pre-activation variant used by vit_base_resnet50_384)."""

    def __init__(self, cin, cout, stride, has_proj):
        super().__init__()
        mid = cout // 4
        self.downsample = DownsampleConv(cin, cout, stride) if has_proj else None
        self.conv1 = StdConv2dSame(cin, mid, 1)
        self.norm1 = GroupNormAct(mid)
        self.conv2 = StdConv2dSame(mid, mid, 3, stride=stride)
        self.norm2 = GroupNormAct(mid)
        self.conv3 = StdConv2dSame(mid, cout, 1)
        self.norm3 = GroupNormAct(cout, apply_act=False)

    def forward(self, x):
        shortcut = x if self.downsample is None else self.downsample(x)
        x = self.norm1(self.conv1(x))
        x = self.norm2(self.conv2(x))
        x = self.norm3(self.conv3(x))
        return F.relu(x + shortcut)


class _Stage(nn.Module):
    def __init__(self, cin, cout, stride, depth):
        super().__init__()
        self.blocks = nn.Sequential(*[Bottleneck(cin if i == 0 else cout, cout, stride if i == 0 else 1, i == 0)
                                      for i in range(depth)])

    def forward(self, x):
        return self.blocks(x)


class _Stem(nn.Module):
    def __init__(self):
        super().__init__()
        self.conv = StdConv2dSame(3, 64, 7, stride=2)
        self.norm = GroupNormAct(64)

    def forward(self, x):
        x = self.norm(self.conv(x))
        x = _pad_same(x, 3, 2, value=float("-inf"))         # MaxPool2dSame
        return F.max_pool2d(x, 3, 2)


class ResNetV2(nn.Module):
    """layers (3,4,9), channels 256/512/1024, strides 1/2/2 -> 1024ch at /16."""

    def __init__(self):
        super().__init__()
        self.stem = _Stem()
        self.stages = nn.Sequential(_Stage(64, 256, 1, 3), _Stage(256, 512, 2, 4), _Stage(512, 1024, 2, 9))

    def forward_features(self, x):
        x = self.stem(x)
        s0 = self.stages[0](x)
        s1 = self.stages[1](s0)
        s2 = self.stages[2](s1)
        return s0, s1, s2


class HybridEmbed(nn.Module):
    def __init__(self, dim=768):
        super().__init__()
        self.backbone = ResNetV2()
        self.proj = nn.Conv2d(1024, dim, kernel_size=1, stride=1)


class VisionTransformer(nn.Module):
    """timm VisionTransformer restricted to what MiDaS forward_flex touches, incl. the unused 1000-way
    ``head`` (real checkpoints carry ``pretrained.model.head.*``)."""

    def __init__(self, dim, depth, heads, hybrid=False, grid=24):
        super().__init__()
        self.hybrid = hybrid
        self.patch_size = (1, 1) if hybrid else (16, 16)
        self.patch_embed = HybridEmbed(dim) if hybrid else PatchEmbed(16, 3, dim)
        self.cls_token = nn.Parameter(torch.zeros(1, 1, dim))
        self.pos_embed = nn.Parameter(torch.zeros(1, grid * grid + 1, dim))
        self.blocks = nn.Sequential(*[Block(dim, heads) for _ in range(depth)])
        self.norm = nn.LayerNorm(dim, eps=1e-6)
        self.head = nn.Linear(dim, 1000)
        self.start_index = 1
        nn.init.trunc_normal_(self.pos_embed, std=0.02)
        nn.init.normal_(self.cls_token, std=1e-6)
        for m in self.modules():
            if isinstance(m, nn.Linear):
                nn.init.trunc_normal_(m.weight, std=0.02)
                nn.init.zeros_(m.bias)

    def resize_pos_embed(self, gs_h, gs_w):
        """[UPSTREAM] midas/vit.py _resize_pos_embed: bilinear, align_corners=False (default)."""
        posemb = self.pos_embed
        tok, grid = posemb[:, :self.start_index], posemb[0, self.start_index:]
        gs_old = int(math.sqrt(len(grid)))
        grid = grid.reshape(1, gs_old, gs_old, -1).permute(0, 3, 1, 2)
        grid = F.interpolate(grid, size=(gs_h, gs_w), mode="bilinear")
        grid = grid.permute(0, 2, 3, 1).reshape(1, gs_h * gs_w, -1)
        return torch.cat([tok, grid], dim=1)

    def forward_flex(self, x, hooks):
        """[UPSTREAM] midas/vit.py forward_flex.  Returns the hooked activations (and, for the hybrid,
        the two ResNet stage outputs) instead of using forward hooks."""
        B, _, h, w = x.shape
        stage_feats = None
        if self.hybrid:
            pos = self.resize_pos_embed(h // 16, w // 16)
            s0, s1, s2 = self.patch_embed.backbone.forward_features(x)
            stage_feats = (s0, s1)
            x = s2
        else:
            pos = self.resize_pos_embed(h // 16, w // 16)
        x = self.patch_embed.proj(x).flatten(2).transpose(1, 2)
        x = torch.cat((self.cls_token.expand(B, -1, -1), x), dim=1)
        x = x + pos
        acts = {}
        for i, blk in enumerate(self.blocks):
            x = blk(x)
            if i in hooks:
                acts[i] = x
        return acts, stage_feats
