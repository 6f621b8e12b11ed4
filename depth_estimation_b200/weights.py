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
