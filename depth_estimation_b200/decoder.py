"""The MiDaS "scratch" decoder (layerN_rn -> refinenet4..1 -> output_conv) as a sequence of sm_100a
kernel launches.  Shared by DPT_Large / DPT_Hybrid (features 256, align_corners=True head) and
MiDaS_small (expand decoder 64/128/256/512, align_corners=False head).

Restates [UPSTREAM] midas/blocks.py `_make_scratch`, `FeatureFusionBlock_custom`,
`ResidualConvUnit_custom` and the heads of midas/dpt_depth.py / midas/midas_net_custom.py, which the
reference reaches through `self.estimator(input_batch)` (/root/reference/depth_midas.py:26).

Layout: NHWC bf16 activations, fp32 accumulation.  Fusions (all parity-checked against the oracle):
  * every 3x3 conv is one implicit-GEMM launch with bias / ReLU / skip adds in its epilogue;
  * the RCU's pre-activation ReLU is produced by the previous kernel as a second (relu'd) output;
  * `path + RCU1(l)` is folded into RCU1.conv2's epilogue (two addends);
  * the 1x1 out_conv is applied BEFORE the bilinear x2 (linear ops commute; interpolation weights sum to
    one, so the bias commutes as well) - 4x fewer FLOPs;
  * the head's upsample x2 + conv3x3(->32)+ReLU+conv1x1(32->1)+ReLU is a single launch for DPT's 128 channels
    (csrc/head_conv.cu: the upsampled activation only ever exists in shared memory); MiDaS_small's 32 channels
    use de_upsample2x + the GEMM kernel's head-dot epilogue.
"""
from __future__ import annotations

import torch

from . import ops
from .weights import pack_conv1x1, pack_conv3x3


class ScratchDecoder:
    def __init__(self, sd: dict, in_channels, features, head_align_corners: bool, device):
        """sd: upstream state_dict (fp32 tensors); keys under 'scratch.'."""
        dev = device
        f32 = lambda k: sd[k].detach().to(dev, torch.float32).contiguous()  # noqa: E731
        self.dev = dev
        self.cin = list(in_channels)
        self.feat = list(features)                     # per-level width f1..f4
        self.head_ac = head_align_corners
        self.rn_w = [pack_conv3x3(sd[f"scratch.layer{i + 1}_rn.weight"].to(dev)) for i in range(4)]
        self.ref = []
        for i in range(4):
            p = f"scratch.refinenet{i + 1}."
            blk = {}
            for u in ("resConfUnit1", "resConfUnit2"):
                for c in ("conv1", "conv2"):
                    blk[f"{u}.{c}.w"] = pack_conv3x3(sd[p + f"{u}.{c}.weight"].to(dev))
                    blk[f"{u}.{c}.b"] = f32(p + f"{u}.{c}.bias")
            blk["out.w"] = pack_conv1x1(sd[p + "out_conv.weight"].to(dev))
            blk["out.b"] = f32(p + "out_conv.bias")
            blk["cout"] = sd[p + "out_conv.weight"].shape[0]
            self.ref.append(blk)
        oc = "scratch.output_conv."
        self.h0_w = pack_conv3x3(sd[oc + "0.weight"].to(dev))
        self.h0_b = f32(oc + "0.bias")
        self.h0_c = sd[oc + "0.weight"].shape[0]
        self.h2_w = pack_conv3x3(sd[oc + "2.weight"].to(dev))
        self.h2_b = f32(oc + "2.bias")
        assert sd[oc + "2.weight"].shape[0] == 32
        self.h4_w = sd[oc + "4.weight"].detach().to(dev, torch.float32).reshape(32).contiguous()
        self.h4_b = float(sd[oc + "4.bias"].item())
        self._bufs = {}

    def _buf(self, key, shape, dtype=torch.bfloat16):
        # keyed on (name, shape, dtype): a buffer whose address is baked into a captured CUDA graph is never replaced
        key = (key, tuple(shape), dtype)
        t = self._bufs.get(key)
        if t is None:
            t = torch.empty(shape, dtype=dtype, device=self.dev)
            self._bufs[key] = t
        return t

    def _conv(self, x, w, out, B, H, W, cin, **kw):
        cout = out.shape[-1]
        return ops.gemm(x, w, out.view(-1, cout), conv=(B, H, W, cin), N=cout, **kw)

    def _rcu(self, tag, blk, unit, x_raw, x_relu, B, H, W, C, extra=None, want_relu=False):
        """conv2(relu(conv1(relu(x)))) + x [+ extra]; returns (raw, relu'd or None)."""
        t = self._buf((tag, "t", B, H, W, C), (B, H, W, C))
        self._conv(x_relu, blk[f"{unit}.conv1.w"], t, B, H, W, C, bias=blk[f"{unit}.conv1.b"], act=ops.ACT_RELU)
        out = self._buf((tag, unit, "o", B, H, W, C), (B, H, W, C))
        out_r = self._buf((tag, unit, "or", B, H, W, C), (B, H, W, C)) if want_relu else None
        self._conv(t, blk[f"{unit}.conv2.w"], out, B, H, W, C, bias=blk[f"{unit}.conv2.b"], add1=x_raw.view(-1, C),
                   add2=extra, out_relu=out_r)
        return out, out_r

    def forward(self, feats, B):
        """feats: 4 NHWC bf16 tensors [B,Hi,Wi,Ci] (levels 1..4, fine -> coarse).  Returns fp32 [B,h,w]."""
        l_raw, l_relu = [], []
        for i, x in enumerate(feats):
            _, H, W, Cc = x.shape
            f = self.feat[i]
            raw = self._buf(("rn", i, B, H, W), (B, H, W, f))
            rel = self._buf(("rnr", i, B, H, W), (B, H, W, f))
            self._conv(x, self.rn_w[i], raw, B, H, W, Cc, out_relu=rel)
            l_raw.append(raw)
            l_relu.append(rel)
        path = None
        for i in (3, 2, 1, 0):
            blk = self.ref[i]
            _, H, W, Cc = l_raw[i].shape
            if path is None:
                s_raw, s_relu = l_raw[i], l_relu[i]
            else:
                s_raw, s_relu = self._rcu(("ffb", i), blk, "resConfUnit1", l_raw[i], l_relu[i], B, H, W, Cc,
                                          extra=path, want_relu=True)
            u, _ = self._rcu(("ffb2", i), blk, "resConfUnit2", s_raw, s_relu, B, H, W, Cc)
            co = blk["cout"]
            v = self._buf(("oc", i, B, H, W), (B, H, W, co))
            ops.gemm(u.view(-1, Cc), blk["out.w"], v.view(-1, co), bias=blk["out.b"])
            path = self._buf(("up", i, B, H, W), (B, 2 * H, 2 * W, co))
            ops.upsample2x(v, path, True)
        _, H, W, Cc = path.shape
        h0 = self._buf(("h0", B, H, W), (B, H, W, self.h0_c))
        self._conv(path, self.h0_w, h0, B, H, W, Cc, bias=self.h0_b)
        pred = self._buf(("pred", B, H, W), (B, 2 * H, 2 * W), torch.float32)
        if self.h0_c % 64 == 0 and self.h0_c <= 128:        # DPT: upsample + head tail in one halo-tiled kernel
            ops.head_up_conv(h0, self.h2_w, self.h2_b, self.h4_w, self.h4_b, pred, self.head_ac)
        else:                                               # MiDaS_small (32 channels): general implicit-GEMM kernel
            h1 = self._buf(("h1", B, H, W), (B, 2 * H, 2 * W, self.h0_c))
            ops.upsample2x(h0, h1, self.head_ac)
            ops.gemm(h1, self.h2_w, pred, conv=(B, 2 * H, 2 * W, self.h0_c), bias=self.h2_b, act=ops.ACT_RELU,
                     head=(self.h4_w, self.h4_b), N=32)
        return pred
