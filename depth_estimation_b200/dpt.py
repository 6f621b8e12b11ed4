"""DPT_Large / DPT_Hybrid forward pass as sm_100a kernel launches.

Restates [UPSTREAM] intel-isl/MiDaS v3.0 `DPTDepthModel.forward` (midas/dpt_depth.py) with
`forward_vit` / `forward_flex` / `_resize_pos_embed` / `ProjectReadout` (midas/vit.py) over timm 0.6.7
`VisionTransformer` - the module the reference calls at /root/reference/depth_midas.py:26.

Numerics: bf16 GEMM operands, fp32 accumulation (TMEM), fp32 residual stream, fp32 LayerNorm statistics,
fp32 softmax; decoder activations bf16 NHWC.
"""
from __future__ import annotations

import os

import torch
import torch.nn.functional as F

from . import ops
from .decoder import ScratchDecoder
from .weights import pack_conv1x1, pack_conv3x3_flat, pack_conv_transpose, pack_linear


def resize_pos_embed(pos: torch.Tensor, gh: int, gw: int) -> torch.Tensor:
    """[UPSTREAM] vit._resize_pos_embed (bilinear, align_corners=False).  One-time weight preparation at
    plan time (per input geometry), not part of the per-frame path.  pos: [1, 1+g*g, D] -> [1+gh*gw, D]."""
    tok, grid = pos[:, :1], pos[0, 1:]
    g = int(round(grid.shape[0] ** 0.5))
    grid = grid.reshape(1, g, g, -1).permute(0, 3, 1, 2)
    grid = F.interpolate(grid.float(), size=(gh, gw), mode="bilinear")
    grid = grid.permute(0, 2, 3, 1).reshape(1, gh * gw, -1)
    return torch.cat([tok.float(), grid], dim=1)[0].contiguous()


class ViTEncoder:
    """timm VisionTransformer blocks on an fp32 token stream x [B*T, D]."""

    def __init__(self, sd: dict, prefix: str, dim: int, depth: int, heads: int, device):
        dev = device
        self.dev, self.dim, self.depth, self.heads = dev, dim, depth, heads
        f32 = lambda k: sd[prefix + k].detach().to(dev, torch.float32).contiguous()  # noqa: E731
        bf = lambda k: pack_linear(sd[prefix + k].detach().to(dev))  # noqa: E731
        self.cls = f32("cls_token").reshape(dim)
        self.pos_raw = sd[prefix + "pos_embed"].detach().to(dev, torch.float32)
        self.blocks = []
        for i in range(depth):
            p = f"blocks.{i}."
            self.blocks.append(dict(
                n1w=f32(p + "norm1.weight"), n1b=f32(p + "norm1.bias"),
                qkv_w=bf(p + "attn.qkv.weight"), qkv_b=f32(p + "attn.qkv.bias"),
                proj_w=bf(p + "attn.proj.weight"), proj_b=f32(p + "attn.proj.bias"),
                n2w=f32(p + "norm2.weight"), n2b=f32(p + "norm2.bias"),
                fc1_w=bf(p + "mlp.fc1.weight"), fc1_b=f32(p + "mlp.fc1.bias"),
                fc2_w=bf(p + "mlp.fc2.weight"), fc2_b=f32(p + "mlp.fc2.bias"),
            ))
        self._pos = {}
        self._bufs = {}

    def pos_embed(self, gh, gw):
        key = (gh, gw)
        if key not in self._pos:
            self._pos[key] = resize_pos_embed(self.pos_raw, gh, gw)
        return self._pos[key]

    def _buf(self, key, shape, dtype):
        # keyed on (name, shape, dtype): a buffer whose address is baked into a captured CUDA graph is never replaced
        key = (key, tuple(shape), dtype)
        t = self._bufs.get(key)
        if t is None:
            t = torch.empty(shape, dtype=dtype, device=self.dev)
            self._bufs[key] = t
        return t

    def run_blocks(self, x: torch.Tensor, B: int, T: int, hooks, on_hook):
        """x: fp32 [B*T, D], updated in place.  on_hook(i, x) is called after each hooked block."""
        D, M = self.dim, B * T
        xn = self._buf(("xn", M), (M, D), torch.bfloat16)
        qkv = self._buf(("qkv", M), (M, 3 * D), torch.bfloat16)
        att = self._buf(("att", M), (M, D), torch.bfloat16)
        hid = self._buf(("hid", M), (M, 4 * D), torch.bfloat16)
        # branch outputs (attention proj / fc2) are ADDED into the fp32 stream by the GEMM itself (TMA reduce-add
        # epilogue, accumulate=True): no bf16 branch buffer, no rounding of the branch, and LayerNorm only reads x
        # (6 instead of 12 bytes per element).  DEB_Y_BUF=1 restores the bf16 branch buffer + add-in-LayerNorm path.
        scale = (D // self.heads) ** -0.5
        if os.environ.get("DEB_Y_BUF"):
            ybuf = self._buf(("y", M), (M, D), torch.bfloat16)
            y = None                                # pending branch output, folded into the next LayerNorm
            for i, w in enumerate(self.blocks):
                ops.layernorm(x, w["n1w"], w["n1b"], 1e-6, xn, y=y)
                if (i - 1) in hooks:                # x now holds the output of block i-1
                    on_hook(i - 1, x)
                ops.gemm(xn, w["qkv_w"], qkv, bias=w["qkv_b"])
                ops.attention(qkv, att, B, T, self.heads, scale)
                ops.gemm(att, w["proj_w"], ybuf, bias=w["proj_b"])
                ops.layernorm(x, w["n2w"], w["n2b"], 1e-6, xn, y=ybuf)
                ops.gemm(xn, w["fc1_w"], hid, bias=w["fc1_b"], act=ops.ACT_GELU)
                ops.gemm(hid, w["fc2_w"], ybuf, bias=w["fc2_b"])
                y = ybuf
            ops.layernorm(x, None, None, 1e-6, None, y=y)
            if (self.depth - 1) in hooks:
                on_hook(self.depth - 1, x)
            return
        for i, w in enumerate(self.blocks):
            ops.layernorm(x, w["n1w"], w["n1b"], 1e-6, xn)
            ops.gemm(xn, w["qkv_w"], qkv, bias=w["qkv_b"])
            ops.attention(qkv, att, B, T, self.heads, scale)
            ops.gemm(att, w["proj_w"], x, bias=w["proj_b"], accumulate=True)
            ops.layernorm(x, w["n2w"], w["n2b"], 1e-6, xn)
            ops.gemm(xn, w["fc1_w"], hid, bias=w["fc1_b"], act=ops.ACT_GELU)
            ops.gemm(hid, w["fc2_w"], x, bias=w["fc2_b"], accumulate=True)
            if i in hooks:
                on_hook(i, x)


class Reassemble:
    """[UPSTREAM] act_postprocessN: ProjectReadout -> 1x1 conv -> {ConvT k4s4, ConvT k2s2, id, conv3x3 s2}."""

    def __init__(self, sd: dict, idx: int, dim: int, device):
        dev = device
        p = f"pretrained.act_postprocess{idx}."
        self.dev, self.dim = dev, dim
        wr = sd[p + "0.project.0.weight"].detach().to(dev)                  # [D, 2D]: [tokens | cls]
        self.w_read = pack_linear(wr)                                       # views below keep ld = 2D
        self.w_tok = self.w_read[:, :dim]
        self.w_cls = self.w_read[:, dim:]
        self.b_read = sd[p + "0.project.0.bias"].detach().to(dev, torch.float32).contiguous()
        self.w1 = pack_conv1x1(sd[p + "3.weight"].detach().to(dev))
        self.b1 = sd[p + "3.bias"].detach().to(dev, torch.float32).contiguous()
        self.cout = self.w1.shape[0]
        self.kind = "same"
        k4 = p + "4.weight"
        if k4 in sd:
            w4 = sd[k4].detach().to(dev)
            self.b4 = sd[p + "4.bias"].detach().to(dev, torch.float32).contiguous()
            if w4.shape[-1] == 3:
                self.kind, self.w4 = "down2", pack_conv3x3_flat(w4)
            else:
                self.kind, self.s, self.w4 = "up", w4.shape[-1], pack_conv_transpose(w4)
        self._bufs = {}

    def _buf(self, key, shape, dtype=torch.bfloat16):
        # keyed on (name, shape, dtype): a buffer whose address is baked into a captured CUDA graph is never replaced
        key = (key, tuple(shape), dtype)
        t = self._bufs.get(key)
        if t is None:
            t = torch.empty(shape, dtype=dtype, device=self.dev)
            self._bufs[key] = t
        return t

    def forward(self, x: torch.Tensor, B: int, gh: int, gw: int) -> torch.Tensor:
        """x: fp32 token stream [B*T, D] (T = 1+gh*gw) at the hooked block.  Returns NHWC bf16."""
        D, N = self.dim, gh * gw
        T = N + 1
        tok = self._buf(("tok", B, N), (B * N, D))
        ops.cast_tokens(x, tok, grp=T, skip=1)
        cls = self._buf(("cls", B), (B, D))
        ops.cast_tokens(x.view(B, T * D)[:, :D], cls)                      # row stride T*D: the cls rows
        cproj = self._buf(("cproj", B), (B, D), torch.float32)
        ops.gemm(cls, self.w_cls, cproj, bias=self.b_read)
        r = self._buf(("r", B, N), (B * N, D))
        ops.gemm(tok, self.w_tok, r, act=ops.ACT_GELU, add1=cproj, add1_rowmode=1, add1_pre=True, grp=N)
        c = self.cout
        p = self._buf(("p", B, N), (B, gh, gw, c))
        ops.gemm(r, self.w1, p.view(-1, c), bias=self.b1)
        if self.kind == "same":
            return p
        if self.kind == "up":
            s = self.s
            out = self._buf(("o", B, N), (B, gh * s, gw * s, c))
            ops.gemm(p.view(-1, c), self.w4, out, bias=self.b4, bias_mod=c, pixel_shuffle=(gh, gw, s, c))
            return out
        ho, wo = (gh - 1) // 2 + 1, (gw - 1) // 2 + 1
        col = self._buf(("col", B, N), (B * ho * wo, 9 * c))
        ops.im2col3x3(p, col, 2)
        out = self._buf(("o", B, N), (B, ho, wo, c))
        ops.gemm(col, self.w4, out.view(-1, c), bias=self.b4)
        return out


class DPTLargeNet:
    """DPT_Large: ViT-L/16 (24 blocks, D=1024, 16 heads, hooks 5/11/17/23) + reassemble + decoder."""

    hooks = (5, 11, 17, 23)

    def __init__(self, sd: dict, device):
        self.dev = device
        self.dim = 1024
        self.vit = ViTEncoder(sd, "pretrained.model.", 1024, 24, 16, device)
        pw = sd["pretrained.model.patch_embed.proj.weight"].detach().to(device)
        self.pe_w = pack_linear(pw.reshape(pw.shape[0], -1))                # [D, 3*16*16], col = c*256+py*16+px
        self.pe_b = sd["pretrained.model.patch_embed.proj.bias"].detach().to(device, torch.float32).contiguous()
        self.reasm = [Reassemble(sd, i + 1, 1024, device) for i in range(4)]
        self.dec = ScratchDecoder(sd, (256, 512, 1024, 1024), (256, 256, 256, 256), True, device)
        self._x = {}
        self.preprocess_layout = ops.LAYOUT_PATCH16_BF16

    def forward(self, patches: torch.Tensor, B: int, net_h: int, net_w: int) -> torch.Tensor:
        """patches: bf16 [B*gh*gw, 768] from K1 (layout 1).  Returns fp32 [B, net_h, net_w] (>= 0)."""
        gh, gw = net_h // 16, net_w // 16
        N, D = gh * gw, self.dim
        T = N + 1
        key = (B, T)
        if key not in self._x:
            self._x[key] = torch.empty((B * T, D), dtype=torch.float32, device=self.dev)
        x = self._x[key]
        pos = self.vit.pos_embed(gh, gw)
        ops.gemm(patches, self.pe_w, x, bias=self.pe_b, add1=pos, add1_rowmode=2, out_rowmode=1, grp=N, pad=1, off=1)
        ops.fill_cls(x, B, T, self.vit.cls, pos)
        feats = [None] * 4

        def on_hook(i, xs):
            j = self.hooks.index(i)
            feats[j] = self.reasm[j].forward(xs, B, gh, gw)

        self.vit.run_blocks(x, B, T, set(self.hooks), on_hook)
        return self.dec.forward(feats, B)
