"""MiDaS_small forward pass as sm_100a kernel launches.

Restates [UPSTREAM] intel-isl/MiDaS `MidasNet_small(features=64, backbone="efficientnet_lite3",
exportable=True, non_negative=True, blocks={'expand': True})` (midas/midas_net_custom.py, midas/blocks.py)
over rwightman/gen-efficientnet-pytorch `tf_efficientnet_lite3` - the module the reference calls at
/root/reference/depth_midas.py:26 for model_type "MiDaS_small".

BatchNorm (eps 1e-3) is folded into the preceding conv at load.  1x1 convs run on the tcgen05 GEMM with
bias / ReLU6 / residual epilogues; depthwise convs on the NHWC bandwidth kernel; the 3-channel stem on the
direct kernel; the decoder is decoder.ScratchDecoder (expand widths 64/128/256/512).
"""
from __future__ import annotations

import torch

from . import ops
from .decoder import ScratchDecoder
from .init import LITE3_STAGES
from .weights import fold_bn

BN_EPS = 1e-3
# upstream Sequential positions of efficientnet stages 0..6 (blocks._make_efficientnet_backbone)
_WHERE = [("layer1", 3), ("layer1", 4), ("layer2", 0), ("layer3", 0), ("layer3", 1), ("layer4", 0), ("layer4", 1)]
_LAYER_END = {1: 0, 2: 1, 4: 2, 6: 3}          # stage index -> decoder level it completes


class MidasSmallNet:
    def __init__(self, sd: dict, device):
        dev = device
        self.dev = dev
        self.preprocess_layout = ops.LAYOUT_NCHW_BF16

        def bn(p):
            return [sd[p + k].detach().to(dev, torch.float32) for k in ("weight", "bias", "running_mean", "running_var")]

        def folded(wkey, bnp):
            w = sd[wkey].detach().to(dev, torch.float32)
            return fold_bn(w, *bn(bnp), BN_EPS)

        w, b = folded("pretrained.layer1.0.weight", "pretrained.layer1.1.")
        self.stem_w = w.permute(0, 2, 3, 1).reshape(32, 27).contiguous()
        self.stem_b = b.contiguous()
        self.blocks = []
        cin = 32
        for si, ((kind, cout, reps, k, s), (layer, li)) in enumerate(zip(LITE3_STAGES, _WHERE)):
            for r in range(reps):
                p = f"pretrained.{layer}.{li}.{r}."
                stride = s if r == 0 else 1
                blk = {"kind": kind, "k": k, "stride": stride, "cin": cin, "cout": cout, "stage": si,
                       "last": r == reps - 1}
                if kind == "ds":
                    w, b = folded(p + "conv_dw.weight", p + "bn1.")
                    blk["dw_w"] = w.reshape(cin, k * k).t().contiguous().to(torch.bfloat16)
                    blk["dw_b"] = b.contiguous()
                    w, b = folded(p + "conv_pw.weight", p + "bn2.")
                    blk["pw_w"] = w.reshape(cout, cin).to(torch.bfloat16).contiguous()
                    blk["pw_b"] = b.contiguous()
                    blk["mid"] = cin
                else:
                    mid = cin * 6
                    w, b = folded(p + "conv_pw.weight", p + "bn1.")
                    blk["exp_w"] = w.reshape(mid, cin).to(torch.bfloat16).contiguous()
                    blk["exp_b"] = b.contiguous()
                    w, b = folded(p + "conv_dw.weight", p + "bn2.")
                    blk["dw_w"] = w.reshape(mid, k * k).t().contiguous().to(torch.bfloat16)
                    blk["dw_b"] = b.contiguous()
                    w, b = folded(p + "conv_pwl.weight", p + "bn3.")
                    blk["pw_w"] = w.reshape(cout, mid).to(torch.bfloat16).contiguous()
                    blk["pw_b"] = b.contiguous()
                    blk["mid"] = mid
                blk["residual"] = stride == 1 and cin == cout
                self.blocks.append(blk)
                cin = cout
        self.dec = ScratchDecoder(sd, (32, 48, 136, 384), (64, 128, 256, 512), False, dev)
        self._bufs = {}

    def _buf(self, key, shape):
        # keyed on (name, shape): a buffer whose address is baked into a captured CUDA graph is never replaced
        key = (key, tuple(shape))
        t = self._bufs.get(key)
        if t is None:
            t = torch.empty(shape, dtype=torch.bfloat16, device=self.dev)
            self._bufs[key] = t
        return t

    def forward(self, x_nchw: torch.Tensor, B: int, net_h: int, net_w: int) -> torch.Tensor:
        """x_nchw: bf16 [B,3,net_h,net_w] from K1 (layout 0).  Returns fp32 [B, net_h, net_w] (>= 0)."""
        H, W = (net_h + 1) // 2, (net_w + 1) // 2
        x = self._buf(("stem", B, H, W), (B, H, W, 32))
        ops.stem_conv(x_nchw, self.stem_w, self.stem_b, x, 3, 2, ops.ACT_RELU6)
        feats = [None] * 4
        for bi, blk in enumerate(self.blocks):
            s, mid, cout, cin = blk["stride"], blk["mid"], blk["cout"], blk["cin"]
            Ho, Wo = (H + s - 1) // s, (W + s - 1) // s
            t = x
            if blk["kind"] == "ir":
                t = self._buf((bi, "e", B, H, W), (B, H, W, mid))
                ops.gemm(x.view(-1, cin), blk["exp_w"], t.view(-1, mid), bias=blk["exp_b"], act=ops.ACT_RELU6)
            d = self._buf((bi, "d", B, Ho, Wo), (B, Ho, Wo, mid))
            ops.depthwise(t, blk["dw_w"], blk["dw_b"], d, blk["k"], s, ops.ACT_RELU6)
            y = self._buf((bi, "y", B, Ho, Wo), (B, Ho, Wo, cout))
            ops.gemm(d.view(-1, mid), blk["pw_w"], y.view(-1, cout), bias=blk["pw_b"],
                     add1=x.view(-1, cin) if blk["residual"] else None)
            x, H, W = y, Ho, Wo
            if blk["last"] and blk["stage"] in _LAYER_END:
                feats[_LAYER_END[blk["stage"]]] = x
        return self.dec.forward(feats, B)
