"""DPT_Hybrid forward pass as sm_100a kernel launches.

Restates [UPSTREAM] intel-isl/MiDaS v3.0 `DPTDepthModel(backbone="vitb_rn50_384")` (midas/dpt_depth.py,
midas/vit.py `_make_vit_b_rn50_backbone`, `forward_flex`) over timm 0.6.7 `vit_base_resnet50_384`
(ResNetV2 3/4/9 with StdConv2dSame + GroupNorm(32)+ReLU, then ViT-B/12) - the module the reference calls
at /root/reference/depth_midas.py:26 for model_type "DPT_Hybrid".

Weight standardisation (timm recomputes it every forward) is folded once at load.  1x1 convs and 3x3
stride-1 convs run on the tcgen05 GEMM / implicit-GEMM kernel, the three 3x3 stride-2 convs through an
im2col gather + GEMM, GroupNorm(+residual+ReLU) on the two-pass NHWC bandwidth kernel.
"""
from __future__ import annotations

import torch

from . import ops
from .decoder import ScratchDecoder
from .dpt import Reassemble, ViTEncoder
from .weights import pack_conv1x1, pack_conv3x3, pack_conv3x3_flat, standardize_weight


def _same_pad_before(n, k, s):
    out = (n + s - 1) // s
    return max((out - 1) * s + k - n, 0) // 2


class _Bottleneck:
    def __init__(self, sd, p, cin, cout, stride, has_proj, dev):
        std = lambda k: standardize_weight(sd[p + k].detach().to(dev, torch.float32))  # noqa: E731
        f32 = lambda k: sd[p + k].detach().to(dev, torch.float32).contiguous()  # noqa: E731
        self.cin, self.cout, self.mid, self.stride, self.has_proj = cin, cout, cout // 4, stride, has_proj
        self.w1 = pack_conv1x1(std("conv1.weight"))
        self.w2 = pack_conv3x3(std("conv2.weight")) if stride == 1 else pack_conv3x3_flat(std("conv2.weight"))
        self.w3 = pack_conv1x1(std("conv3.weight"))
        self.n = [(f32(f"norm{i}.weight"), f32(f"norm{i}.bias")) for i in (1, 2, 3)]
        if has_proj:
            self.wd = pack_conv1x1(std("downsample.conv.weight"))
            self.nd = (f32("downsample.norm.weight"), f32("downsample.norm.bias"))


class DPTHybridNet:
    hooks = (8, 11)

    def __init__(self, sd: dict, device):
        dev = device
        self.dev, self.dim = dev, 768
        self.preprocess_layout = ops.LAYOUT_NCHW_BF16
        bp = "pretrained.model.patch_embed.backbone."
        w = standardize_weight(sd[bp + "stem.conv.weight"].detach().to(dev, torch.float32))
        self.stem_w = w.permute(0, 2, 3, 1).reshape(64, 147).contiguous()
        self.stem_n = (sd[bp + "stem.norm.weight"].detach().to(dev, torch.float32).contiguous(),
                       sd[bp + "stem.norm.bias"].detach().to(dev, torch.float32).contiguous())
        self.stages = []
        cin = 64
        for si, (cout, depth) in enumerate(((256, 3), (512, 4), (1024, 9))):
            blocks = []
            for b in range(depth):
                blocks.append(_Bottleneck(sd, f"{bp}stages.{si}.blocks.{b}.", cin if b == 0 else cout, cout,
                                          (1 if si == 0 else 2) if b == 0 else 1, b == 0, dev))
            self.stages.append(blocks)
            cin = cout
        self.pe_w = pack_conv1x1(sd["pretrained.model.patch_embed.proj.weight"].detach().to(dev))
        self.pe_b = sd["pretrained.model.patch_embed.proj.bias"].detach().to(dev, torch.float32).contiguous()
        self.vit = ViTEncoder(sd, "pretrained.model.", 768, 12, 12, dev)
        self.reasm = [Reassemble(sd, 3, 768, dev), Reassemble(sd, 4, 768, dev)]
        self.dec = ScratchDecoder(sd, (256, 512, 768, 768), (256, 256, 256, 256), True, dev)
        self._bufs = {}
        self._x = {}
        self._ws = torch.zeros(4 << 20, dtype=torch.uint8, device=dev)

    def _buf(self, key, shape, dtype=torch.bfloat16):
        t = self._bufs.get(key)
        if t is None or tuple(t.shape) != tuple(shape):
            t = torch.empty(shape, dtype=dtype, device=self.dev)
            self._bufs[key] = t
        return t

    def _gn(self, x, nrm, key, act=ops.ACT_RELU, add=None, want_f32=False):
        out = self._buf(key, tuple(x.shape))
        o32 = self._buf((key, "f32"), tuple(x.shape), torch.float32) if want_f32 else None
        ops.groupnorm(x, nrm[0], nrm[1], out, self._ws, act=act, add=add, out_f32=o32)
        return out, o32

    def _bottleneck(self, tag, blk: _Bottleneck, x, x32, B, H, W):
        """x: bf16 block input (GEMM operand); x32: the same tensor in fp32 (residual stream) or None."""
        s, mid, cout, cin = blk.stride, blk.mid, blk.cout, blk.cin
        Ho, Wo = (H + s - 1) // s, (W + s - 1) // s
        if blk.has_proj:
            xs = x
            if s > 1:
                xs = self._buf((tag, "sub"), (B, Ho, Wo, cin))
                ops.subsample(x, xs, s)
            d = self._buf((tag, "dc"), (B, Ho, Wo, cout))
            ops.gemm(xs.view(-1, cin), blk.wd, d.view(-1, cout))
            _, shortcut = self._gn(d, blk.nd, (tag, "dn"), act=ops.ACT_NONE, want_f32=True)
        else:
            shortcut = x32
        t = self._buf((tag, "c1"), (B, H, W, mid))
        ops.gemm(x.view(-1, cin), blk.w1, t.view(-1, mid))
        t, _ = self._gn(t, blk.n[0], (tag, "n1"))
        u = self._buf((tag, "c2"), (B, Ho, Wo, mid))
        if s == 1:
            ops.gemm(t, blk.w2, u.view(-1, mid), conv=(B, H, W, mid), N=mid)
        else:
            col = self._buf((tag, "col"), (B * Ho * Wo, 9 * mid))
            ops.im2col3x3(t, col, s, pad=(_same_pad_before(H, 3, s), _same_pad_before(W, 3, s)), out_hw=(Ho, Wo))
            ops.gemm(col, blk.w2, u.view(-1, mid))
        u, _ = self._gn(u, blk.n[1], (tag, "n2"))
        v = self._buf((tag, "c3"), (B, Ho, Wo, cout))
        ops.gemm(u.view(-1, mid), blk.w3, v.view(-1, cout))
        y, y32 = self._gn(v, blk.n[2], (tag, "n3"), act=ops.ACT_RELU, add=shortcut, want_f32=True)
        return y, y32, Ho, Wo

    def forward(self, x_nchw: torch.Tensor, B: int, net_h: int, net_w: int) -> torch.Tensor:
        """x_nchw: bf16 [B,3,net_h,net_w] from K1.  Returns fp32 [B, net_h, net_w] (>= 0)."""
        H, W = (net_h + 1) // 2, (net_w + 1) // 2
        s0 = self._buf(("stem", B, H, W), (B, H, W, 64))
        ops.stem_conv(x_nchw, self.stem_w, None, s0, 7, 2, ops.ACT_NONE)
        s0, _ = self._gn(s0, self.stem_n, ("stemn", B, H, W))
        Hp, Wp = (H + 1) // 2, (W + 1) // 2
        x = self._buf(("pool", B, Hp, Wp), (B, Hp, Wp, 64))
        ops.maxpool3x3s2(s0, x)
        H, W = Hp, Wp
        stage_out = []
        x32 = None
        for si, blocks in enumerate(self.stages):
            for bi, blk in enumerate(blocks):
                x, x32, H, W = self._bottleneck((si, bi, B, H, W), blk, x, x32, B, H, W)
            stage_out.append(x)
        gh, gw = net_h // 16, net_w // 16
        assert (H, W) == (gh, gw)
        N, D = gh * gw, self.dim
        T = N + 1
        if (B, T) not in self._x:
            self._x[(B, T)] = torch.empty((B * T, D), dtype=torch.float32, device=self.dev)
        tok = self._x[(B, T)]
        pos = self.vit.pos_embed(gh, gw)
        ops.gemm(x.view(-1, 1024), self.pe_w, tok, bias=self.pe_b, add1=pos, add1_rowmode=2, out_rowmode=1, grp=N,
                 pad=1, off=1)
        ops.fill_cls(tok, B, T, self.vit.cls, pos)
        feats = [stage_out[0], stage_out[1], None, None]

        def on_hook(i, xs):
            j = self.hooks.index(i)
            feats[2 + j] = self.reasm[j].forward(xs, B, gh, gw)

        self.vit.run_blocks(tok, B, T, set(self.hooks), on_hook)
        return self.dec.forward(feats, B)
