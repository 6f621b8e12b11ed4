"""One-time repacking of upstream-format (torch ``state_dict``) weights into the layouts the sm_100a
kernels consume: bf16, K-major [N,K] matrices, 3x3 filters as [Cout, 9*ceil64(Cin)] tap-major rows,
ConvTranspose(k==s) as [(i,j,co), ci], BatchNorm and weight-standardisation folded (SURVEY.md section 5,
"Checkpoint / resume").  Pure tensor reshuffling - runs wherever the tensors live."""
from __future__ import annotations

import torch


def _ceil64(c: int) -> int:
    return (c + 63) // 64 * 64


def pack_conv3x3(w: torch.Tensor) -> torch.Tensor:
    """[Cout,Cin,3,3] -> bf16 [Cout, 9*cin_pad], k = (kh*3+kw)*cin_pad + ci, zero padded."""
    cout, cin, kh, kw = w.shape
    assert kh == 3 and kw == 3
    cp = _ceil64(cin)
    out = torch.zeros((cout, 9, cp), dtype=torch.bfloat16, device=w.device)
    out[:, :, :cin] = w.permute(0, 2, 3, 1).reshape(cout, 9, cin).to(torch.bfloat16)
    return out.reshape(cout, 9 * cp).contiguous()


def pack_conv3x3_flat(w: torch.Tensor) -> torch.Tensor:
    """[Cout,Cin,3,3] -> bf16 [Cout, 9*Cin] with k = tap*Cin + ci (for de_im2col3x3 + plain GEMM)."""
    cout, cin = w.shape[:2]
    return w.permute(0, 2, 3, 1).reshape(cout, 9 * cin).to(torch.bfloat16).contiguous()


def pack_conv_transpose(w: torch.Tensor) -> torch.Tensor:
    """ConvTranspose2d weight [Cin,Cout,s,s] (kernel == stride) -> bf16 [(i*s+j)*Cout+co, Cin]."""
    cin, cout, s, s2 = w.shape
    assert s == s2
    return w.permute(2, 3, 1, 0).reshape(s * s * cout, cin).to(torch.bfloat16).contiguous()


def pack_linear(w: torch.Tensor) -> torch.Tensor:
    return w.to(torch.bfloat16).contiguous()


def pack_conv1x1(w: torch.Tensor) -> torch.Tensor:
    return w.reshape(w.shape[0], w.shape[1]).to(torch.bfloat16).contiguous()


def _hi_lo(w: torch.Tensor):
    hi = w.to(torch.bfloat16)
    lo = (w.float() - hi.float()).to(torch.bfloat16)
    return hi, lo


def split3_conv1x1(w: torch.Tensor) -> torch.Tensor:
    """fp32 [Cout,Cin(,1,1)] -> bf16 [Cout, 3*Cin] = [Wh | Wl | Wh] (hi = bf16(w), lo = bf16(w - hi)).  Against an
    activation row laid out [Ah | Ah | Al] (the kernels' split3 output mode) the ordinary bf16 GEMM accumulates
    Ah*Wh + Ah*Wl + Al*Wh in fp32: a bf16x3 product with ~16 mantissa bits, used where bf16's 8 bits are not enough
    (DPT_Hybrid's weight-standardised ResNetV2 - see dpt_hybrid.py)."""
    hi, lo = _hi_lo(w.reshape(w.shape[0], w.shape[1]))
    return torch.cat([hi, lo, hi], dim=1).contiguous()


def split3_conv3x3(w: torch.Tensor) -> torch.Tensor:
    """fp32 [Cout,Cin,3,3] -> bf16 [Cout, 9*3*Cin], k = tap*(3*Cin) + {Wh | Wl | Wh}: the tap-major layout of both the
    implicit-GEMM conv (Cin' = 3*Cin must be a multiple of 64) and de_im2col3x3 + GEMM, on split3 activations."""
    cout, cin = w.shape[:2]
    assert (3 * cin) % 64 == 0
    hi, lo = _hi_lo(w.permute(0, 2, 3, 1).reshape(cout, 9, cin))
    return torch.cat([hi, lo, hi], dim=2).reshape(cout, 27 * cin).contiguous()


def fold_bn(w: torch.Tensor, bn_w, bn_b, bn_mean, bn_var, eps: float):
    """conv (no bias) followed by BatchNorm(eval) -> (w', b') in fp32."""
    scale = bn_w / torch.sqrt(bn_var + eps)
    return w * scale.view(-1, *([1] * (w.dim() - 1))), bn_b - bn_mean * scale


def standardize_weight(w: torch.Tensor, eps: float = 1e-8) -> torch.Tensor:
    """timm StdConv2dSame weight standardisation (biased variance), folded once at load."""
    flat = w.reshape(w.shape[0], -1).double()
    mean = flat.mean(dim=1, keepdim=True)
    var = flat.var(dim=1, unbiased=False, keepdim=True)
    return ((flat - mean) / torch.sqrt(var + eps)).reshape_as(w).float()


# ------------------------------------------------------------------------------------------------ checkpoints
# Upstream release files the reference's torch.hub call downloads (depth_midas.py:10) - names only, for messages.
UPSTREAM_CHECKPOINTS = {"DPT_Large": "dpt_large-midas-2f21e586.pt", "DPT_Hybrid": "dpt_hybrid-midas-501f0c75.pt",
                        "MiDaS_small": "midas_v21_small-70d6b9c8.pt"}
# keys real checkpoints carry that the forward pass never reads (SURVEY.md 8a footer)
_IGNORABLE = ("num_batches_tracked",)


def state_dict_checksum(sd: dict) -> str:
    """sha256 over (key, dtype, shape, raw bytes) in sorted key order - identifies a weight set independent of the
    container format it was stored in."""
    import hashlib
    h = hashlib.sha256()
    for k in sorted(sd):
        t = sd[k].detach().cpu().contiguous()
        h.update(k.encode())
        h.update(str(t.dtype).encode())
        h.update(str(tuple(t.shape)).encode())
        h.update(t.view(torch.uint8).numpy().tobytes() if t.numel() else b"")
    return h.hexdigest()


def load_checkpoint(src, model_type: str, strict: bool = True) -> dict:
    """Checkpoint ingestion: `src` is a path (torch.load, tensors only) or an already loaded
    mapping in upstream naming - a bare state_dict, or one wrapped as {"state_dict": ...} / {"model": ...}, with or
    without a "module." prefix.  Every key of the architecture's inventory (init.param_shapes) must be present with
    the right shape; unknown keys are an error under `strict` (BatchNorm's num_batches_tracked is dropped).
    Returns fp32 CPU tensors ready for the net constructors, which fold BN / weight standardisation and repack to
    bf16 (functions above)."""
    from .init import param_shapes
    if isinstance(src, (str, bytes)) or hasattr(src, "__fspath__"):
        obj = torch.load(src, map_location="cpu", weights_only=True)
    else:
        obj = src
    for wrap in ("state_dict", "model"):
        if isinstance(obj, dict) and wrap in obj and isinstance(obj[wrap], dict):
            obj = obj[wrap]
    if not isinstance(obj, dict):
        raise TypeError(f"checkpoint for {model_type} is not a state_dict mapping")
    sd = {}
    for k, v in obj.items():
        if not torch.is_tensor(v):
            raise TypeError(f"checkpoint entry {k!r} is not a tensor")
        k = k[len("module."):] if k.startswith("module.") else k
        if k.endswith(_IGNORABLE):
            continue
        sd[k] = v.detach().to("cpu", torch.float32)
    want = param_shapes(model_type)
    missing = [k for k in want if k not in sd]
    wrong = [f"{k}: {tuple(sd[k].shape)} != {tuple(want[k])}" for k in want if k in sd and tuple(sd[k].shape) != tuple(want[k])]
    extra = [k for k in sd if k not in want]
    if missing or wrong or (strict and extra):
        hint = UPSTREAM_CHECKPOINTS.get(model_type, "?")
        raise ValueError(f"checkpoint does not match {model_type} (upstream file {hint}): "
                         f"{len(missing)} missing {missing[:3]}, {len(wrong)} mis-shaped {wrong[:3]}, "
                         f"{len(extra)} unexpected {extra[:3]}")
    return {k: sd[k] for k in want}
