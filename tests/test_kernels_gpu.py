"""GPU parity tests for the individual sm_100a kernels, called through the C ABI (ops.py -> ctypes).
The checker is torch fp32 on the same (bf16-rounded) inputs, or the CPU oracle for K1/K8."""
import math

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch.device("cuda:0")


def _rand(shape, dev, scale=1.0, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (torch.randn(shape, generator=g) * scale).to(dev)


def _relerr(a, b):
    a, b = a.float(), b.float()
    return ((a - b).abs().max() / (b.abs().max() + 1e-12)).item()


# ------------------------------------------------------------------------------------ GEMM
GEMM_CASES = [
    # M, N, K, block_n
    (128, 256, 64, 0),
    (577, 1024, 1024, 0),
    (577, 3072, 1024, 256),
    (1000, 128, 192, 128),
    (300, 96, 144, 0),
    (4617, 64, 256, 64),
    (50, 32, 64, 32),
    (18464, 1024, 1024, 128),
    (2308, 24, 144, 0),
]


@pytest.mark.parametrize("M,N,K,bn", GEMM_CASES)
def test_gemm_plain(M, N, K, bn):
    from depth_estimation_b200 import ops
    dev = _dev()
    a = _rand((M, K), dev, seed=1).bfloat16()
    w = _rand((N, K), dev, 0.05, seed=2).bfloat16()
    out = torch.full((M, N), float("nan"), dtype=torch.float32, device=dev)
    ops.gemm(a, w, out, block_n=bn)
    torch.cuda.synchronize()
    ref = a.float() @ w.float().t()
    assert torch.isfinite(out).all()
    assert _relerr(out, ref) < 2e-5, _relerr(out, ref)


@pytest.mark.parametrize("act", [0, 1, 2, 3])
def test_gemm_epilogue_bias_act_residual(act):
    from depth_estimation_b200 import ops
    dev = _dev()
    M, N, K = 700, 512, 320
    a = _rand((M, K), dev, seed=3).bfloat16()
    w = _rand((N, K), dev, 0.08, seed=4).bfloat16()
    bias = _rand((N,), dev, seed=5)
    res = _rand((M, N), dev, seed=6)
    out = torch.empty((M, N), dtype=torch.float32, device=dev)
    ops.gemm(a, w, out, bias=bias, act=act, add1=res)
    v = a.float() @ w.float().t() + bias
    v = {0: lambda t: t, 1: F.relu, 2: F.gelu, 3: F.relu6}[act](v) + res
    torch.cuda.synchronize()
    assert _relerr(out, v) < 2e-5
    # bf16 output + pre-activation addend + relu copy
    outb = torch.empty((M, N), dtype=torch.bfloat16, device=dev)
    outr = torch.empty((M, N), dtype=torch.bfloat16, device=dev)
    resb = res.bfloat16()
    ops.gemm(a, w, outb, bias=bias, act=act, add1=resb, add1_pre=True, out_relu=outr)
    v2 = {0: lambda t: t, 1: F.relu, 2: F.gelu, 3: F.relu6}[act](a.float() @ w.float().t() + bias + resb.float())
    torch.cuda.synchronize()
    assert _relerr(outb, v2) < 6e-3
    assert torch.equal(outr, F.relu(outb.float()).bfloat16())


def test_gemm_row_maps_tokens():
    """patch-embed style: rows m=(b,p) -> b*(N+1)+1+p, addend = pos[1+p] broadcast over b."""
    from depth_estimation_b200 import ops
    dev = _dev()
    B, Np, D, K = 3, 36, 256, 768
    a = _rand((B * Np, K), dev, seed=7).bfloat16()
    w = _rand((D, K), dev, 0.05, seed=8).bfloat16()
    bias = _rand((D,), dev, seed=9)
    pos = _rand((Np + 1, D), dev, seed=10)
    x = torch.zeros((B * (Np + 1), D), dtype=torch.float32, device=dev)
    ops.gemm(a, w, x, bias=bias, add1=pos, add1_rowmode=2, out_rowmode=1, grp=Np, pad=1, off=1)
    ref = (a.float() @ w.float().t() + bias).view(B, Np, D) + pos[1:][None]
    torch.cuda.synchronize()
    xv = x.view(B, Np + 1, D)
    assert _relerr(xv[:, 1:], ref) < 2e-5
    assert (xv[:, 0] == 0).all()
    # per-group broadcast addend before the activation (readout: W_cls x_cls + b)
    cb = _rand((B, D), dev, seed=11)
    o = torch.empty((B * Np, D), dtype=torch.bfloat16, device=dev)
    ops.gemm(a, w, o, act=2, add1=cb, add1_rowmode=1, add1_pre=True, grp=Np)
    ref2 = F.gelu((a.float() @ w.float().t()).view(B, Np, D) + cb[:, None]).view(B * Np, D)
    torch.cuda.synchronize()
    assert _relerr(o, ref2) < 6e-3


@pytest.mark.parametrize("M,N,K", [(256, 256, 64), (577, 1024, 1024), (18464, 3072, 1024), (1000, 512, 320), (4617, 256, 2304)])
def test_gemm_cta_pair(M, N, K):
    """cta_group::2 path: CTA pair per 256-row tile (odd 128-row tile counts included)."""
    from depth_estimation_b200 import ops
    dev = _dev()
    a = _rand((M, K), dev, seed=70).bfloat16()
    w = _rand((N, K), dev, 0.05, seed=71).bfloat16()
    bias = _rand((N,), dev, seed=72)
    res = _rand((M, N), dev, seed=73)
    out = torch.full((M, N), float("nan"), dtype=torch.float32, device=dev)
    ops.gemm(a, w, out, bias=bias, add1=res, block_n=256, cta_group=2)
    torch.cuda.synchronize()
    ref = a.float() @ w.float().t() + bias + res
    assert torch.isfinite(out).all()
    assert _relerr(out, ref) < 2e-5, _relerr(out, ref)


@pytest.mark.parametrize("B,H,W,Cin", [(3, 24, 24, 256), (1, 96, 96, 256), (5, 12, 12, 1024)])
def test_conv3x3_cta_pair(B, H, W, Cin):
    from depth_estimation_b200 import ops
    from depth_estimation_b200.weights import pack_conv3x3
    dev = _dev()
    x = _rand((B, Cin, H, W), dev, seed=74).bfloat16()
    wt = _rand((256, Cin, 3, 3), dev, 0.05, seed=75).bfloat16()
    out = torch.full((B, H, W, 256), float("nan"), dtype=torch.float32, device=dev)
    ops.gemm(x.permute(0, 2, 3, 1).contiguous(), pack_conv3x3(wt), out.view(-1, 256), conv=(B, H, W, Cin), N=256,
             block_n=256, cta_group=2)
    ref = F.conv2d(x.float(), wt.float(), None, padding=1).permute(0, 2, 3, 1)
    torch.cuda.synchronize()
    assert torch.isfinite(out).all()
    assert _relerr(out, ref) < 2e-5


@pytest.mark.parametrize("M,N,K,cg,bn", [(577, 1024, 1024, 0, 0), (18464, 3072, 1024, 2, 256), (1000, 512, 320, 2, 256),
                                         (700, 128, 192, 1, 128), (333, 72, 136, 0, 0), (4617, 256, 256, 1, 256)])
@pytest.mark.parametrize("act", [0, 2])
@pytest.mark.parametrize("f32", [False, True])
def test_gemm_store_epilogue(M, N, K, cg, bn, act, f32):
    """TMA-store epilogue (bias + activation, dense output): bf16 / fp32, single CTA and CTA pair, N tails."""
    from depth_estimation_b200 import ops
    dev = _dev()
    a = _rand((M, K), dev, seed=80).bfloat16()
    w = _rand((N, K), dev, 0.05, seed=81).bfloat16()
    bias = _rand((N,), dev, seed=82)
    out = torch.full((M, N), float("nan"), dtype=torch.float32 if f32 else torch.bfloat16, device=dev)
    ops.gemm(a, w, out, bias=bias, act=act, block_n=bn, cta_group=cg)
    ref = a.float() @ w.float().t() + bias
    if act == 2:
        ref = F.gelu(ref)
    torch.cuda.synchronize()
    assert torch.isfinite(out.float()).all()
    assert _relerr(out, ref) < (2e-5 if f32 else 6e-3), _relerr(out, ref)
    # the general epilogue (epilogue=1 forces it) must agree with the store epilogue
    out2 = torch.empty_like(out)
    ops.gemm(a, w, out2, bias=bias, act=act, block_n=bn, cta_group=cg, epilogue=1)
    torch.cuda.synchronize()
    assert torch.equal(out, out2)


@pytest.mark.parametrize("M,N,K,cg,bn", [(577, 1024, 1024, 0, 0), (18464, 1024, 1024, 2, 256), (1154, 768, 3072, 0, 0),
                                         (333, 72, 136, 0, 0)])
def test_gemm_accumulate(M, N, K, cg, bn):
    """accumulate=True: x += A W^T + bias through the TMA reduce-add epilogue (the ViT residual stream update).  Exactly
    one add per element: x0 + (store-epilogue result) bit for bit, twice in a row, and row tails stay untouched."""
    from depth_estimation_b200 import ops
    dev = _dev()
    a = _rand((M, K), dev, seed=90).bfloat16()
    w = _rand((N, K), dev, 0.05, seed=91).bfloat16()
    bias = _rand((N,), dev, seed=92)
    x0 = _rand((M + 3, N), dev, seed=93)
    y = torch.empty((M, N), dtype=torch.float32, device=dev)
    ops.gemm(a, w, y, bias=bias, block_n=bn, cta_group=cg)
    x = x0.clone()
    ops.gemm(a, w, x[:M], bias=bias, block_n=bn, cta_group=cg, accumulate=True)
    torch.cuda.synchronize()
    assert torch.equal(x[:M], x0[:M] + y) and torch.equal(x[M:], x0[M:])
    ops.gemm(a, w, x[:M], bias=bias, block_n=bn, cta_group=cg, accumulate=True)
    torch.cuda.synchronize()
    assert torch.equal(x[:M], (x0[:M] + y) + y)
    with pytest.raises(RuntimeError):                      # bf16 output cannot accumulate
        ops.gemm(a, w, torch.empty((M, N), dtype=torch.bfloat16, device=dev), bias=bias, accumulate=True)


@pytest.mark.parametrize("f32", [True, False])
def test_gemm_gelu_pointwise_accuracy(f32):
    """The epilogue's exp2-polynomial GELU against torch's erf GELU on every bf16 value in [-9, 9] (identity
    weight, so the accumulator IS the input): abs <= 4e-6 and rel <= 3e-4 in fp32; <= 1 bf16 ulp when packed (where |gelu| > 1e-6)."""
    from depth_estimation_b200 import ops
    dev = _dev()
    K = 64
    vals = torch.arange(-(1 << 15), 1 << 15, dtype=torch.int32).to(torch.int16).view(torch.bfloat16).float()
    vals = vals[torch.isfinite(vals) & (vals.abs() <= 9.0)]
    M = (vals.numel() + K - 1) // K
    a = torch.zeros(M * K)
    a[: vals.numel()] = vals
    a = a.view(M, K).to(dev).bfloat16()
    w = torch.eye(K, device=dev).bfloat16()
    out = torch.empty((M, K), dtype=torch.float32 if f32 else torch.bfloat16, device=dev)
    ops.gemm(a, w, out, act=2)
    torch.cuda.synchronize()
    ref = F.gelu(a.double()).float()
    if f32:
        err = (out - ref).abs()
        assert err.max().item() <= 4e-6, err.max().item()
        m = ref.abs() > 1e-6
        assert (err[m] / ref[m].abs()).max().item() <= 3e-4
    else:
        r = ref.bfloat16()
        ulp = (out.view(torch.int16).int() - r.view(torch.int16).int()).abs()
        big = ref.abs() > 1e-6                              # below that only the absolute error means anything
        assert ulp[big].max().item() <= 1, ulp[big].max().item()
        assert (out.float() - ref)[~big].abs().max().item() <= 1e-6


@pytest.mark.parametrize("B,H,W,Cin,Cout,cg", [(3, 24, 24, 256, 256, 2), (1, 96, 96, 256, 256, 2), (2, 21, 12, 64, 128, 0),
                                               (2, 32, 40, 48, 64, 0), (1, 24, 42, 256, 256, 1)])
def test_conv3x3_store_epilogue(B, H, W, Cin, Cout, cg):
    from depth_estimation_b200 import ops
    from depth_estimation_b200.weights import pack_conv3x3
    dev = _dev()
    x = _rand((B, Cin, H, W), dev, seed=83).bfloat16()
    wt = _rand((Cout, Cin, 3, 3), dev, 0.05, seed=84).bfloat16()
    bias = _rand((Cout,), dev, seed=85)
    out = torch.full((B, H, W, Cout), float("nan"), dtype=torch.bfloat16, device=dev)
    ops.gemm(x.permute(0, 2, 3, 1).contiguous(), pack_conv3x3(wt), out.view(-1, Cout), conv=(B, H, W, Cin), N=Cout,
             bias=bias, act=1, cta_group=cg, block_n=256 if cg == 2 else 0)
    ref = F.relu(F.conv2d(x.float(), wt.float(), bias, padding=1)).permute(0, 2, 3, 1)
    torch.cuda.synchronize()
    assert torch.isfinite(out.float()).all()
    assert _relerr(out, ref) < 6e-3


CONV_CASES = [
    # B, H, W, Cin, Cout
    (2, 12, 12, 256, 256),
    (1, 24, 24, 1024, 256),
    (2, 48, 48, 512, 256),
    (1, 96, 96, 256, 256),
    (1, 24, 42, 256, 256),
    (1, 21, 12, 64, 128),
    (2, 32, 40, 48, 64),
    (1, 16, 16, 136, 256),
    (1, 64, 64, 128, 32),
]


@pytest.mark.parametrize("B,H,W,Cin,Cout", CONV_CASES)
def test_conv3x3_implicit_gemm(B, H, W, Cin, Cout):
    from depth_estimation_b200 import ops
    from depth_estimation_b200.weights import pack_conv3x3
    dev = _dev()
    x = _rand((B, Cin, H, W), dev, seed=12).bfloat16()
    wt = _rand((Cout, Cin, 3, 3), dev, 0.05, seed=13).bfloat16()
    bias = _rand((Cout,), dev, seed=14)
    x_nhwc = x.permute(0, 2, 3, 1).contiguous()
    wp = pack_conv3x3(wt)
    out = torch.full((B, H, W, Cout), float("nan"), dtype=torch.float32, device=dev)
    ops.gemm(x_nhwc, wp, out.view(-1, Cout), conv=(B, H, W, Cin), bias=bias, N=Cout)
    ref = F.conv2d(x.float(), wt.float(), bias, padding=1).permute(0, 2, 3, 1)
    torch.cuda.synchronize()
    assert torch.isfinite(out).all()
    assert _relerr(out, ref) < 2e-5, _relerr(out, ref)


def test_conv3x3_random_geometries():
    """Implicit-GEMM 3x3 convolution on seeded random geometries - images narrower than the 8-pixel halo box, prime
    extents, channel counts that are not multiples of 64 - through both epilogues (fp32 store; bf16 with residual + ReLU
    copy), against torch fp32.  Exercises the halo mode (bw % 8 == 0 boxes with zero-filled overhang) and its fallback."""
    from depth_estimation_b200 import ops
    from depth_estimation_b200.weights import pack_conv3x3
    dev = _dev()
    rng = np.random.default_rng(777)
    geos = [(1, 3, 5, 64, 64), (2, 7, 70, 72, 128), (1, 53, 9, 192, 256), (3, 8, 8, 64, 256)]
    for _ in range(8):
        geos.append((int(rng.integers(1, 4)), int(rng.integers(3, 71)), int(rng.integers(3, 71)),
                     int(rng.choice([64, 72, 128, 136, 192])), int(rng.choice([64, 128, 256]))))
    for i, (B, H, W, Cin, Cout) in enumerate(geos):
        x = _rand((B, Cin, H, W), dev, seed=400 + i).bfloat16()
        wt = _rand((Cout, Cin, 3, 3), dev, 0.05, seed=420 + i).bfloat16()
        bias = _rand((Cout,), dev, seed=440 + i)
        x_nhwc = x.permute(0, 2, 3, 1).contiguous()
        wp = pack_conv3x3(wt)
        ref = F.conv2d(x.float(), wt.float(), bias, padding=1).permute(0, 2, 3, 1)
        out = torch.full((B, H, W, Cout), float("nan"), dtype=torch.float32, device=dev)
        ops.gemm(x_nhwc, wp, out.view(-1, Cout), conv=(B, H, W, Cin), bias=bias, N=Cout)
        r = _rand((B, H, W, Cout), dev, seed=460 + i).bfloat16()
        o16 = torch.full((B, H, W, Cout), float("nan"), dtype=torch.bfloat16, device=dev)
        orl = torch.full_like(o16, float("nan"))
        ops.gemm(x_nhwc, wp, o16.view(-1, Cout), conv=(B, H, W, Cin), bias=bias, N=Cout, add1=r.view(-1, Cout), out_relu=orl.view(-1, Cout))
        torch.cuda.synchronize()
        assert torch.isfinite(out).all() and torch.isfinite(o16.float()).all(), (B, H, W, Cin, Cout)
        assert _relerr(out, ref) < 2e-5, (B, H, W, Cin, Cout, _relerr(out, ref))
        assert _relerr(o16, ref + r.float()) < 6e-3, (B, H, W, Cin, Cout)
        assert torch.equal(orl, F.relu(o16)), (B, H, W, Cin, Cout)


def test_conv3x3_rcu_epilogue():
    """RCU tail: conv2(..)+bias + x + path, with raw and relu'd copies stored."""
    from depth_estimation_b200 import ops
    from depth_estimation_b200.weights import pack_conv3x3
    dev = _dev()
    B, H, W, Cc = 2, 24, 24, 256
    x = _rand((B, H, W, Cc), dev, seed=15).bfloat16()
    skip = _rand((B, H, W, Cc), dev, seed=16).bfloat16()
    path = _rand((B, H, W, Cc), dev, seed=17).bfloat16()
    wt = _rand((Cc, Cc, 3, 3), dev, 0.03, seed=18).bfloat16()
    bias = _rand((Cc,), dev, seed=19)
    out = torch.empty((B, H, W, Cc), dtype=torch.bfloat16, device=dev)
    outr = torch.empty_like(out)
    ops.gemm(x, pack_conv3x3(wt), out.view(-1, Cc), conv=(B, H, W, Cc), bias=bias, add1=skip.view(-1, Cc),
             add2=path, out_relu=outr)
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), wt.float(), bias, padding=1).permute(0, 2, 3, 1) + skip.float() + path.float()
    torch.cuda.synchronize()
    assert _relerr(out, ref) < 6e-3
    assert torch.equal(outr, F.relu(out.float()).bfloat16())


@pytest.mark.parametrize("s,Cc", [(4, 256), (2, 512)])
def test_conv_transpose_pixel_shuffle(s, Cc):
    from depth_estimation_b200 import ops
    from depth_estimation_b200.weights import pack_conv_transpose
    dev = _dev()
    B, gh, gw = 2, 6, 7
    x = _rand((B, Cc, gh, gw), dev, seed=20).bfloat16()
    wt = _rand((Cc, Cc, s, s), dev, 0.05, seed=21).bfloat16()     # [Cin, Cout, kh, kw]
    bias = _rand((Cc,), dev, seed=22)
    a = x.permute(0, 2, 3, 1).reshape(B * gh * gw, Cc).contiguous()
    out = torch.full((B, gh * s, gw * s, Cc), float("nan"), dtype=torch.float32, device=dev)
    ops.gemm(a, pack_conv_transpose(wt), out, bias=bias, bias_mod=Cc, pixel_shuffle=(gh, gw, s, Cc))
    ref = F.conv_transpose2d(x.float(), wt.float(), bias, stride=s).permute(0, 2, 3, 1)
    torch.cuda.synchronize()
    assert torch.isfinite(out).all()
    assert _relerr(out, ref) < 2e-5


def test_head_dot():
    from depth_estimation_b200 import ops
    from depth_estimation_b200.weights import pack_conv3x3
    dev = _dev()
    B, H, W, Cin = 1, 64, 96, 128
    x = _rand((B, H, W, Cin), dev, seed=23).bfloat16()
    wt = _rand((32, Cin, 3, 3), dev, 0.05, seed=24).bfloat16()
    bias = _rand((32,), dev, seed=25)
    w2 = _rand((32,), dev, seed=26)
    out = torch.empty((B, H, W), dtype=torch.float32, device=dev)
    ops.gemm(x, pack_conv3x3(wt), out, conv=(B, H, W, Cin), bias=bias, act=1, head=(w2, 0.7), N=32)
    t = F.relu(F.conv2d(x.float().permute(0, 3, 1, 2), wt.float(), bias, padding=1))
    ref = F.relu((t * w2.view(1, 32, 1, 1)).sum(1) + 0.7)
    torch.cuda.synchronize()
    assert _relerr(out, ref) < 2e-5


@pytest.mark.parametrize("B,H,W,Cin", [(1, 64, 96, 128), (2, 20, 40, 128), (3, 8, 16, 64), (1, 37, 53, 128), (2, 96, 96, 192)])
def test_head_conv(B, H, W, Cin):
    """Halo-tiled head kernel (csrc/head_conv.cu) vs torch fp32 and vs the general kernel's head-dot epilogue, on
    tile-aligned and ragged sizes (partial tiles on both edges, TMA zero-filled halo)."""
    from depth_estimation_b200 import ops
    from depth_estimation_b200.weights import pack_conv3x3
    dev = _dev()
    x = _rand((B, H, W, Cin), dev, seed=27).bfloat16()
    wt = _rand((32, Cin, 3, 3), dev, 0.05, seed=28).bfloat16()
    bias = _rand((32,), dev, seed=29)
    w2 = _rand((32,), dev, seed=26)
    out = torch.full((B, H, W), -7.0, dtype=torch.float32, device=dev)
    ops.head_conv(x, pack_conv3x3(wt), bias, w2, 0.7, out)
    t = F.relu(F.conv2d(x.float().permute(0, 3, 1, 2), wt.float(), bias, padding=1))
    ref = F.relu((t * w2.view(1, 32, 1, 1)).sum(1) + 0.7)
    torch.cuda.synchronize()
    assert _relerr(out, ref) < 2e-5
    if Cin % 64 == 0:
        gen = torch.empty_like(out)
        ops.gemm(x, pack_conv3x3(wt), gen, conv=(B, H, W, Cin), bias=bias, act=1, head=(w2, 0.7), N=32)
        torch.cuda.synchronize()
        assert _relerr(out, gen) < 2e-6


@pytest.mark.parametrize("B,Hs,Ws,Cin,ac", [(1, 32, 48, 128, True), (2, 10, 20, 128, True), (1, 19, 27, 64, True),
                                             (2, 12, 8, 128, False), (1, 48, 48, 128, True), (2, 16, 32, 64, False),
                                             (3, 45, 61, 128, True), (1, 2, 3, 64, True), (2, 96, 96, 128, True)])
def test_head_up_conv(B, Hs, Ws, Cin, ac):
    """Fused bilinear x2 + head tail vs the two-kernel path (bit-identical interpolation; the three horizontal taps are
    summed in the epilogue instead of inside one accumulator, so fp32 rounding order differs) and vs torch."""
    from depth_estimation_b200 import ops
    from depth_estimation_b200.weights import pack_conv3x3
    dev = _dev()
    x = _rand((B, Hs, Ws, Cin), dev, seed=81).bfloat16()
    wt = _rand((32, Cin, 3, 3), dev, 0.05, seed=82).bfloat16()
    bias = _rand((32,), dev, seed=83)
    w2 = _rand((32,), dev, seed=84)
    out = torch.full((B, 2 * Hs, 2 * Ws), -7.0, dtype=torch.float32, device=dev)
    ops.head_up_conv(x, pack_conv3x3(wt), bias, w2, 0.3, out, ac)
    up = torch.empty((B, 2 * Hs, 2 * Ws, Cin), dtype=torch.bfloat16, device=dev)
    ops.upsample2x(x, up, ac)
    two = torch.empty_like(out)
    ops.head_conv(up, pack_conv3x3(wt), bias, w2, 0.3, two)
    t = F.interpolate(x.float().permute(0, 3, 1, 2), scale_factor=2, mode="bilinear", align_corners=ac)
    t = F.relu(F.conv2d(t.bfloat16().float(), wt.float(), bias, padding=1))
    ref = F.relu((t * w2.view(1, 32, 1, 1)).sum(1) + 0.3)
    torch.cuda.synchronize()
    assert _relerr(out, two) < 1e-5, _relerr(out, two)
    assert _relerr(out, ref) < 5e-3


# ------------------------------------------------------------------------------------ others
@pytest.mark.parametrize("dim", [768, 1024])
def test_layernorm(dim):
    from depth_estimation_b200 import ops
    dev = _dev()
    x = _rand((1155, dim), dev, 2.0, seed=30) + 0.5
    g, b = _rand((dim,), dev, seed=31), _rand((dim,), dev, seed=32)
    out = torch.empty((1155, dim), dtype=torch.bfloat16, device=dev)
    ops.layernorm(x, g, b, 1e-6, out)
    ref = F.layer_norm(x, (dim,), g, b, 1e-6)
    torch.cuda.synchronize()
    assert _relerr(out, ref) < 5e-3
    # fused residual add: x += y in place, out = LN(x + y); and the add-only form
    y = _rand((1155, dim), dev, seed=39)
    x0 = x.clone()
    ops.layernorm(x, g, b, 1e-6, out, y=y)
    torch.cuda.synchronize()
    assert torch.equal(x, x0 + y)
    assert _relerr(out, F.layer_norm(x0 + y, (dim,), g, b, 1e-6)) < 5e-3
    ops.layernorm(x, None, None, 1e-6, None, y=y)
    torch.cuda.synchronize()
    assert torch.equal(x, (x0 + y) + y)
    # bf16 branch output (what the ViT GEMM epilogues write): the fp32 stream adds the exact bf16 value
    yb = y.bfloat16()
    x1 = x.clone()
    ops.layernorm(x, g, b, 1e-6, out, y=yb)
    torch.cuda.synchronize()
    assert torch.equal(x, x1 + yb.float())
    assert _relerr(out, F.layer_norm(x1 + yb.float(), (dim,), g, b, 1e-6)) < 5e-3


@pytest.mark.parametrize("B,T,heads", [(2, 577, 16), (1, 1009, 12), (3, 64, 4), (1, 65, 2), (2, 128, 1), (2, 129, 3), (5, 769, 16)])
def test_attention(B, T, heads):
    from depth_estimation_b200 import ops
    dev = _dev()
    D = heads * 64
    qkv = _rand((B * T, 3 * D), dev, 1.0, seed=33).bfloat16()
    out = torch.full((B * T, D), float("nan"), dtype=torch.bfloat16, device=dev)
    ops.attention(qkv, out, B, T, heads, 0.125)
    q, k, v = qkv.float().view(B, T, 3, heads, 64).permute(2, 0, 3, 1, 4)
    ref = F.scaled_dot_product_attention(q, k, v).transpose(1, 2).reshape(B * T, D)
    torch.cuda.synchronize()
    assert torch.isfinite(out.float()).all()
    assert _relerr(out, ref) < 1.5e-2, _relerr(out, ref)


@pytest.mark.parametrize("ac", [True, False])
def test_upsample2x(ac):
    from depth_estimation_b200 import ops
    dev = _dev()
    x = _rand((2, 13, 21, 64), dev, seed=34).bfloat16()
    out = torch.empty((2, 26, 42, 64), dtype=torch.bfloat16, device=dev)
    ops.upsample2x(x, out, ac)
    ref = F.interpolate(x.float().permute(0, 3, 1, 2), scale_factor=2, mode="bilinear", align_corners=ac).permute(0, 2, 3, 1)
    torch.cuda.synchronize()
    assert _relerr(out, ref) < 5e-3


def test_im2col_cast_fill():
    from depth_estimation_b200 import ops
    dev = _dev()
    x = _rand((2, 24, 24, 64), dev, seed=35).bfloat16()
    col = torch.empty((2 * 12 * 12, 9 * 64), dtype=torch.bfloat16, device=dev)
    ops.im2col3x3(x, col, 2)
    ref = F.unfold(x.float().permute(0, 3, 1, 2), 3, padding=1, stride=2)           # [B, C*9, L], (c, tap) order
    ref = ref.view(2, 64, 9, 144).permute(0, 3, 2, 1).reshape(288, 576)
    torch.cuda.synchronize()
    assert torch.equal(col.float(), ref)
    t = _rand((3 * 10, 128), dev, seed=36)
    o = torch.empty((3 * 9, 128), dtype=torch.bfloat16, device=dev)
    ops.cast_tokens(t, o, grp=10, skip=1)
    torch.cuda.synchronize()
    assert torch.equal(o, t.view(3, 10, 128)[:, 1:].reshape(27, 128).bfloat16())
    cls, pos = _rand((128,), dev, seed=37), _rand((10, 128), dev, seed=38)
    ops.fill_cls(t, 3, 10, cls, pos)
    torch.cuda.synchronize()
    assert torch.equal(t.view(3, 10, 128)[:, 0], (cls + pos[0]).expand(3, 128))


# ------------------------------------------------------------------------------------ K1 / K8
PRE_CASES = [(640, 480, "MiDaS_small"), (1280, 720, "MiDaS_small"), (1920, 1080, "DPT_Large"),
             (384, 384, "DPT_Large"), (640, 480, "DPT_Hybrid"), (1280, 720, "DPT_Large"), (333, 517, "DPT_Hybrid")]


@pytest.mark.parametrize("W,H,mt", PRE_CASES)
@pytest.mark.parametrize("kind", ["U", "S"])
def test_preprocess_vs_oracle(W, H, mt, kind):
    from depth_estimation_b200 import ops
    from oracle import transforms as T
    from oracle.predict import synthetic_frame
    dev = _dev()
    frames = np.stack([synthetic_frame(W, H, 7, kind, i) for i in range(2)])
    name = T.transform_name_for(mt)
    _, _, mean, std = T.TRANSFORM_SPECS[name]
    ref = np.concatenate([T.midas_transform_bgr(f, mt) for f in frames])          # fp32 [2,3,h,w]
    nw, nh = ops.get_size(W, H, mt)
    assert ref.shape[2:] == (nh, nw)
    fr = torch.from_numpy(frames).to(dev)
    o32 = ops.preprocess(fr, nh, nw, mean, std, ops.LAYOUT_NCHW_F32)
    o16 = ops.preprocess(fr, nh, nw, mean, std, ops.LAYOUT_NCHW_BF16)
    op = ops.preprocess(fr, nh, nw, mean, std, ops.LAYOUT_PATCH16_BF16)
    torch.cuda.synchronize()
    reft = torch.from_numpy(ref).to(dev)
    assert (o32 - reft).abs().max().item() < 2e-5
    # bf16 output within 1 bf16 ulp of bf16(oracle fp32)
    refb = reft.bfloat16()
    # (near zero the fp32 result itself may differ by ~2e-5 absolute, many "ulps" of a tiny number)
    ulp = torch.clamp(refb.float().abs() * 2.0 ** -7, min=4e-5)
    assert ((o16.float() - refb.float()).abs() <= ulp).all()
    assert (o16 != refb).float().mean().item() < 0.01
    # patch layout is the same numbers rearranged: [B, gh, gw, c, py, px]
    gh, gw = nh // 16, nw // 16
    rearr = o16.view(2, 3, gh, 16, gw, 16).permute(0, 2, 4, 1, 3, 5).reshape(2 * gh * gw, 768)
    assert torch.equal(op, rearr)


POST_CASES = [(192, 256, 480, 640), (128, 256, 720, 1280), (384, 672, 1080, 1920), (384, 384, 384, 384), (384, 384, 100, 130)]


@pytest.mark.parametrize("h,w,H,W", POST_CASES)
def test_postprocess_vs_torch(h, w, H, W):
    from depth_estimation_b200 import ops
    dev = _dev()
    pred = _rand((2, h, w), dev, seed=40).abs() * 10
    rc = torch.tensor([[[0, 0], [H - 1, W - 1], [H // 3, W // 2]], [[5, 7], [H // 2, 3], [H - 2, W - 3]]],
                      dtype=torch.int32, device=dev)
    omap, ometric, samples = ops.postprocess(pred, H, W, want_map=True, want_metric=True, sample_rc=rc)
    ref = F.interpolate(pred.cpu().unsqueeze(1), size=(H, W), mode="bicubic", align_corners=False).squeeze(1).to(dev)
    torch.cuda.synchronize()
    assert _relerr(omap, ref) < 1e-5
    assert torch.equal(ometric, (omap * (1.0 / 2.54)).bfloat16())
    assert _relerr(ometric, ref / 2.54) < 1e-2          # north_star: bf16 metric depth within 1 %
    for b in range(2):
        for s in range(3):
            r, c = rc[b, s].tolist()
            assert samples[b, s].item() == omap[b, r, c].item()
    # video grey level (midas_mediapipe.py:141-150): uint8(map)*5+10 with numpy's wrapping casts, incl. values
    # above 255, negative overshoot and a NaN
    pred_w = pred * 40.0 - 30.0
    pred_w[0, 0, 0] = float("nan")
    vis = torch.empty((2, H, W), dtype=torch.uint8, device=dev)
    omap_w, _, _ = ops.postprocess(pred_w, H, W, out_vis=vis)
    torch.cuda.synchronize()
    host = omap_w.cpu().numpy()
    with np.errstate(invalid="ignore"):
        want = host.astype(np.uint8)
    want *= 5
    want += 10
    assert (host > 255).any() and (host < 0).any()
    assert np.array_equal(vis.cpu().numpy(), want)
    # bf16 input path
    omap2, _, _ = ops.postprocess(pred.bfloat16(), H, W)
    ref2 = F.interpolate(pred.bfloat16().float().cpu().unsqueeze(1), size=(H, W), mode="bicubic", align_corners=False).squeeze(1).to(dev)
    torch.cuda.synchronize()
    assert _relerr(omap2, ref2) < 1e-5


def test_postprocess_random_geometries():
    """K8 on seeded random geometries - up- and down-scaling, widths that are / are not multiples of 4 and 8, maps wider
    than one tile, tall and flat aspect ratios - against ATen's bicubic on the CPU; every output view from ONE launch
    agrees with the fp32 map, and the fp32-only launch (which may take the 256-column tile) equals the multi-output one."""
    from depth_estimation_b200 import ops
    dev = _dev()
    rng = np.random.default_rng(4242)
    geos = [(64, 96, 320, 480), (128, 256, 720, 1280), (96, 160, 301, 517)]      # fixed: >= 4x (wide tile), ragged output
    for _ in range(10):
        h, w = int(rng.integers(24, 200)), int(rng.integers(24, 260))
        fy, fx = rng.uniform(0.6, 6.0), rng.uniform(0.6, 6.0)
        geos.append((h, w, max(8, int(h * fy)), max(8, int(w * fx))))
    for i, (h, w, H, W) in enumerate(geos):
        pred = _rand((2, h, w), dev, seed=300 + i).abs() * 7 + 0.1
        ref = F.interpolate(pred.cpu().unsqueeze(1), size=(H, W), mode="bicubic", align_corners=False).squeeze(1).to(dev)
        vis = torch.empty((2, H, W), dtype=torch.uint8, device=dev)
        omap, ometric, _ = ops.postprocess(pred, H, W, want_map=True, want_metric=True, out_vis=vis)
        only, _, _ = ops.postprocess(pred, H, W, want_map=True)
        torch.cuda.synchronize()
        assert _relerr(omap, ref) < 1e-5, (h, w, H, W)
        assert torch.equal(only, omap), (h, w, H, W)
        assert torch.equal(ometric, (omap * (1.0 / 2.54)).bfloat16()), (h, w, H, W)
        host = omap.cpu().numpy()
        with np.errstate(invalid="ignore"):
            want = host.astype(np.uint8)
        want *= 5
        want += 10
        assert np.array_equal(vis.cpu().numpy(), want), (h, w, H, W)


# ------------------------------------------------------------------------------------ direct convs
def _tf_same_pad(x, k, s, value=0.0):
    ih, iw = x.shape[-2:]
    ph = max((math.ceil(ih / s) - 1) * s + k - ih, 0)
    pw = max((math.ceil(iw / s) - 1) * s + k - iw, 0)
    return F.pad(x, [pw // 2, pw - pw // 2, ph // 2, ph - ph // 2], value=value)


@pytest.mark.parametrize("k,s,Cout,H,W", [(3, 2, 32, 192, 256), (7, 2, 64, 96, 130), (3, 2, 32, 37, 41)])
def test_stem_conv(k, s, Cout, H, W):
    from depth_estimation_b200 import ops
    dev = _dev()
    x = _rand((2, 3, H, W), dev, seed=50).bfloat16()
    w = _rand((Cout, 3, k, k), dev, 0.1, seed=51)
    b = _rand((Cout,), dev, seed=52)
    Ho, Wo = (H + s - 1) // s, (W + s - 1) // s
    out = torch.empty((2, Ho, Wo, Cout), dtype=torch.bfloat16, device=dev)
    ops.stem_conv(x, w.permute(0, 2, 3, 1).reshape(Cout, -1).contiguous(), b, out, k, s, ops.ACT_RELU6)
    ref = F.relu6(F.conv2d(_tf_same_pad(x.float(), k, s), w, b, stride=s)).permute(0, 2, 3, 1)
    torch.cuda.synchronize()
    assert out.shape == ref.shape
    assert _relerr(out, ref) < 5e-3


@pytest.mark.parametrize("k,s,Cc,H,W,B", [(3, 1, 32, 96, 128, 2), (3, 2, 144, 96, 128, 2), (5, 2, 192, 48, 64, 2), (5, 1, 816, 12, 16, 2),
                                          (5, 2, 816, 12, 16, 2), (3, 1, 1392, 6, 8, 2), (5, 2, 48, 9, 7, 2),
                                          # frame-packed tiles: 4x8 pixels of four frames (ragged batch: the last box reads
                                          # past B and is zero-filled), 8x8 / 4x16 of two frames, batch too small to pack
                                          (5, 1, 1392, 4, 8, 7), (3, 1, 1392, 4, 8, 32), (5, 1, 232, 6, 8, 5), (3, 1, 136, 12, 16, 3),
                                          (5, 1, 96, 4, 8, 1), (3, 1, 96, 4, 8, 2), (5, 1, 72, 3, 5, 6), (3, 1, 64, 11, 7, 4)])
def test_depthwise(k, s, Cc, H, W, B):
    from depth_estimation_b200 import ops
    dev = _dev()
    x = _rand((B, Cc, H, W), dev, seed=53).bfloat16()
    w = _rand((Cc, 1, k, k), dev, 0.2, seed=54).bfloat16()
    b = _rand((Cc,), dev, seed=55)
    Ho, Wo = (H + s - 1) // s, (W + s - 1) // s
    out = torch.full((B + 1, Ho, Wo, Cc), 7.0, dtype=torch.bfloat16, device=dev)     # one guard frame behind the batch
    ops.depthwise(x.permute(0, 2, 3, 1).contiguous(), w.reshape(Cc, k * k).t().contiguous(), b, out[:B], k, s, ops.ACT_RELU6)
    ref = F.relu6(F.conv2d(_tf_same_pad(x.float(), k, s), w.float(), b, stride=s, groups=Cc)).permute(0, 2, 3, 1)
    torch.cuda.synchronize()
    assert out[:B].shape == ref.shape
    assert _relerr(out[:B], ref) < 5e-3
    assert bool((out[B] == 7.0).all()), "a packed tile wrote past the last frame"


@pytest.mark.parametrize("th_tw", [816, 808, 416, 408])
def test_depthwise_tile_shapes_agree(th_tw):
    """Every tile shape of the stride-1 kernel (8x16 of one frame, 8x8 / 4x16 of two, 4x8 of four - selected through the
    image size) gives the same bits as the 8x16 shape on the same pixels: the accumulation order per output does not
    depend on the tile."""
    from depth_estimation_b200 import ops
    dev = _dev()
    H, W, B = {816: (16, 32, 3), 808: (8, 8, 4), 416: (4, 16, 4), 408: (4, 8, 8)}[th_tw]
    Cc, k = 136, 5
    # embed the image in a larger, zero-bordered one whose tile shape is 8x16: same taps in the same (ky, kx) order on
    # the same values (zeros where the small image has its TF-same padding)
    x = _rand((B, Cc, H, W), dev, seed=60).bfloat16()
    big = torch.zeros((B, Cc, 24, 40), dtype=torch.bfloat16, device=dev)
    big[:, :, 4:4 + H, 4:4 + W] = x
    w = _rand((Cc, 1, k, k), dev, 0.2, seed=61).bfloat16()
    b = _rand((Cc,), dev, seed=62)
    wt = w.reshape(Cc, k * k).t().contiguous()
    out = torch.empty((B, H, W, Cc), dtype=torch.bfloat16, device=dev)
    out_big = torch.empty((B, 24, 40, Cc), dtype=torch.bfloat16, device=dev)
    ops.depthwise(x.permute(0, 2, 3, 1).contiguous(), wt, b, out, k, 1, ops.ACT_RELU6)
    ops.depthwise(big.permute(0, 2, 3, 1).contiguous(), wt, b, out_big, k, 1, ops.ACT_RELU6)
    torch.cuda.synchronize()
    assert torch.equal(out, out_big[:, 4:4 + H, 4:4 + W, :])


@pytest.mark.parametrize("k,Cc,cout,B,H,W,res", [
    (3, 32, 24, 2, 64, 128, False),        # DS block of stage 0: many tiles, one CTA per tile
    (3, 144, 32, 2, 33, 50, False),        # ragged image, partial last channel chunk (144 = 2.25 x 64)
    (5, 288, 48, 3, 16, 32, True),         # residual
    (3, 576, 96, 4, 8, 16, True),          # one tile per frame: chunks split over CTA groups, workspace reduction
    (5, 816, 136, 4, 8, 16, True),         # cout padded to 144 columns
    (5, 1392, 232, 5, 4, 8, True),         # tile mostly padding, cout padded to 240
    (3, 1392, 384, 2, 4, 8, False),        # two MMA column parts (192 + 192), 512 TMEM columns
])
def test_dwproject_fused_mbconv_tail(k, Cc, cout, B, H, W, res):
    """de_dwproject == de_depthwise -> de_gemm (+ residual), the pair it replaces: same depthwise arithmetic (bit for
    bit up to its bf16 rounding), the projection against torch fp32; twice in a row on one workspace (counters left
    at zero), and bitwise reproducible."""
    from depth_estimation_b200 import ops
    dev = _dev()
    x = (_rand((B, H, W, Cc), dev, 1.5, seed=120) + 0.2).bfloat16()
    dw = _rand((k * k, Cc), dev, 0.3, seed=121).bfloat16()
    db = _rand((Cc,), dev, 0.5, seed=122)
    pw = _rand((cout, Cc), dev, 0.05, seed=123).bfloat16()
    pb = _rand((cout,), dev, seed=124)
    r = _rand((B, H, W, cout), dev, seed=125).bfloat16() if res else None
    # the unfused pair
    mid = torch.empty((B, H, W, Cc), dtype=torch.bfloat16, device=dev)
    ops.depthwise(x, dw, db, mid, k, 1, ops.ACT_RELU6)
    want = torch.empty((B, H, W, cout), dtype=torch.bfloat16, device=dev)
    ops.gemm(mid.view(-1, Cc), pw, want.view(-1, cout), bias=pb, add1=r.view(-1, cout) if res else None)
    ref = F.linear(mid.float(), pw.float(), pb)
    if res:
        ref = ref + r.float()
    ws = torch.zeros(ops.dwproject_ws_bytes(B, H, W, cout), dtype=torch.uint8, device=dev)
    got = torch.full_like(want, float("nan"))
    ops.dwproject(x, dw, db, pw, pb, got, k, ws, residual=r)
    got2 = torch.full_like(want, float("nan"))
    ops.dwproject(x, dw, db, pw, pb, got2, k, ws, residual=r)
    torch.cuda.synchronize()
    assert torch.isfinite(got.float()).all()
    assert _relerr(got, ref) < 6e-3 and _relerr(want, ref) < 6e-3
    assert _relerr(got, want) < 6e-3
    assert torch.equal(got, got2)
    assert int(ws[:16384].max()) == 0


def test_maxpool():
    from depth_estimation_b200 import ops
    dev = _dev()
    for H, W in ((192, 192), (35, 22)):
        x = _rand((2, 64, H, W), dev, seed=56).bfloat16()
        out = torch.empty((2, (H + 1) // 2, (W + 1) // 2, 64), dtype=torch.bfloat16, device=dev)
        ops.maxpool3x3s2(x.permute(0, 2, 3, 1).contiguous(), out)
        ref = F.max_pool2d(_tf_same_pad(x.float(), 3, 2, float("-inf")), 3, 2).permute(0, 2, 3, 1)
        torch.cuda.synchronize()
        assert torch.equal(out.float(), ref)


@pytest.mark.parametrize("Cc,H,W,B", [(64, 96, 96, 1), (256, 48, 48, 2), (128, 24, 40, 2), (1024, 12, 12, 1), (512, 7, 9, 3)])
@pytest.mark.parametrize("with_add", [False, True])
def test_groupnorm(Cc, H, W, B, with_add):
    from depth_estimation_b200 import ops
    dev = _dev()
    x = (_rand((B, H, W, Cc), dev, 2.0, seed=60) + 0.7).bfloat16()
    g, b = _rand((Cc,), dev, seed=61), _rand((Cc,), dev, seed=62)
    add = _rand((B, H, W, Cc), dev, seed=63).bfloat16() if with_add else None
    out = torch.empty_like(x)
    ws = torch.zeros(ops.groupnorm_ws_bytes(B, H * W, Cc), dtype=torch.uint8, device=dev)    # exactly the advertised size
    o32 = torch.empty(x.shape, dtype=torch.float32, device=dev)
    ops.groupnorm(x, g, b, out, ws, act=ops.ACT_RELU, add=add, out_f32=o32)
    out_again = torch.empty_like(x)
    ops.groupnorm(x, g, b, out_again, ws, act=ops.ACT_RELU, add=add.float() if with_add else None)
    ref = F.group_norm(x.float().permute(0, 3, 1, 2), 32, g, b, 1e-5).permute(0, 2, 3, 1)
    if with_add:
        ref = ref + add.float()
    ref = F.relu(ref)
    torch.cuda.synchronize()
    assert _relerr(out, ref) < 6e-3
    assert _relerr(o32, ref) < 1e-4
    assert torch.equal(out, out_again) and torch.equal(out, o32.bfloat16())      # deterministic


def test_subsample_and_same_im2col():
    from depth_estimation_b200 import ops
    dev = _dev()
    x = _rand((2, 10, 14, 32), dev, seed=64).bfloat16()
    o = torch.empty((2, 5, 7, 32), dtype=torch.bfloat16, device=dev)
    ops.subsample(x, o, 2)
    torch.cuda.synchronize()
    assert torch.equal(o, x[:, ::2, ::2])
    col = torch.empty((2 * 5 * 7, 9 * 32), dtype=torch.bfloat16, device=dev)
    ops.im2col3x3(x, col, 2, pad=(0, 0), out_hw=(5, 7))            # TF-same, even input: pad (0 before, 1 after)
    xp = F.pad(x.float().permute(0, 3, 1, 2), [0, 1, 0, 1])
    ref = F.unfold(xp, 3, stride=2).view(2, 32, 9, 35).permute(0, 3, 2, 1).reshape(70, 288)
    torch.cuda.synchronize()
    assert torch.equal(col.float(), ref)


# ------------------------------------------------------------------ split3 (bf16x3) operand path of DPT_Hybrid's ResNetV2
def _unsplit(t, Cc):
    """[.., 3C] = [hi | hi | lo] -> fp32 hi + lo (and check the two hi copies agree)."""
    hi, hi2, lo = t[..., :Cc].float(), t[..., Cc:2 * Cc].float(), t[..., 2 * Cc:].float()
    assert torch.equal(hi, hi2)
    return hi + lo


def test_groupnorm_large_batch_shared_workspace():
    """A workspace sized by de_groupnorm_ws_bytes carries any batch (the fixed 4 MB of round 1 capped DPT_Hybrid at
    B ~ 55), and calls of different B / HW sharing it leave the arrival counters (first 16 KB) at zero."""
    from depth_estimation_b200 import ops
    dev = _dev()
    shapes = [(64, 48, 48, 64), (3, 96, 96, 64), (70, 12, 12, 256)]
    ws = torch.zeros(max(ops.groupnorm_ws_bytes(B, H * W, Cc) for B, H, W, Cc in shapes), dtype=torch.uint8, device=dev)
    for rep in range(2):
        for B, H, W, Cc in shapes:
            x = _rand((B, H, W, Cc), dev, 2.0, seed=70 + rep) + 0.3
            g, b = _rand((Cc,), dev, seed=71), _rand((Cc,), dev, seed=72)
            o32 = torch.empty_like(x)
            ops.groupnorm(x, g, b, None, ws, act=ops.ACT_NONE, out_f32=o32)
            ref = F.group_norm(x.permute(0, 3, 1, 2), 32, g, b, 1e-5).permute(0, 2, 3, 1)
            torch.cuda.synchronize()
            assert _relerr(o32, ref) < 1e-5
            assert int(ws[:16384].max()) == 0
    with pytest.raises(RuntimeError):
        ops.groupnorm(x, g, b, None, ws[:1024], act=ops.ACT_NONE, out_f32=o32)


def test_groupnorm_f32_in_split_out():
    from depth_estimation_b200 import ops
    dev = _dev()
    B, H, W, Cc = 2, 24, 40, 128
    x = _rand((B, H, W, Cc), dev, 2.0, seed=64) + 0.7
    g, b = _rand((Cc,), dev, seed=65), _rand((Cc,), dev, seed=66)
    add = _rand((B, H, W, Cc), dev, seed=67)
    ws = torch.zeros(1 << 20, dtype=torch.uint8, device=dev)
    out = torch.empty((B, H, W, 3 * Cc), dtype=torch.bfloat16, device=dev)
    o32 = torch.empty_like(x)
    opl = torch.empty((B, H, W, Cc), dtype=torch.bfloat16, device=dev)
    ops.groupnorm(x, g, b, out, ws, act=ops.ACT_RELU, add=add, out_f32=o32, split=True, out_plain=opl)
    ref = F.relu(F.group_norm(x.double().permute(0, 3, 1, 2), 32, g.double(), b.double(), 1e-5).permute(0, 2, 3, 1) + add.double())
    torch.cuda.synchronize()
    assert _relerr(o32, ref) < 2e-6
    assert torch.equal(opl, o32.bfloat16())
    assert torch.equal(out[..., :Cc], opl)
    assert (_unsplit(out, Cc) - o32).abs().max() <= o32.abs().max() * 2.0 ** -16      # hi + lo carries ~16 bits
    # fp32-only output (the stem's norm and the downsample shortcut)
    only = torch.empty_like(x)
    ops.groupnorm(x, g, b, None, ws, act=ops.ACT_RELU, add=add, out_f32=only)
    torch.cuda.synchronize()
    assert torch.equal(only, o32)


def test_maxpool_f32_in_split_out():
    from depth_estimation_b200 import ops
    dev = _dev()
    x = _rand((2, 64, 35, 22), dev, seed=68)
    out = torch.empty((2, 18, 11, 192), dtype=torch.bfloat16, device=dev)
    ops.maxpool3x3s2(x.permute(0, 2, 3, 1).contiguous(), out, split=True)
    ref = F.max_pool2d(_tf_same_pad(x, 3, 2, float("-inf")), 3, 2).permute(0, 2, 3, 1)
    torch.cuda.synchronize()
    assert (_unsplit(out, 64) - ref).abs().max() <= ref.abs().max() * 2.0 ** -16


def test_stem_conv_f32():
    from depth_estimation_b200 import ops
    dev = _dev()
    k, s, Cout, H, W = 7, 2, 64, 96, 130
    x = _rand((2, 3, H, W), dev, seed=69)
    w = _rand((Cout, 3, k, k), dev, 0.1, seed=70)
    out = torch.empty((2, H // 2, W // 2, Cout), dtype=torch.float32, device=dev)
    ops.stem_conv(x, w.permute(0, 2, 3, 1).reshape(Cout, -1).contiguous(), None, out, k, s, ops.ACT_NONE)
    ref = F.conv2d(_tf_same_pad(x.double(), k, s), w.double(), None, stride=s).permute(0, 2, 3, 1)
    torch.cuda.synchronize()
    assert _relerr(out, ref) < 1e-5


@pytest.mark.parametrize("B,H,W", [(2, 96, 130), (1, 384, 384), (3, 33, 47)])
def test_stem_im2col_split3_gemm(B, H, W):
    """DPT_Hybrid's 7x7/2 TF-same stem as split3 im2col rows x [Wh | Wl | Wh] on the bf16 GEMM: bf16x3 precision
    against fp64, identical rows to a torch unfold of the padded input, zero padding columns."""
    from depth_estimation_b200 import ops
    from depth_estimation_b200.weights import split3_conv1x1
    dev = _dev()
    k, s, Cout, kp = 7, 2, 64, 152
    x = _rand((B, 3, H, W), dev, seed=71)
    w = _rand((Cout, 3, k, k), dev, 0.1, seed=72)
    Ho, Wo = (H + 1) // 2, (W + 1) // 2
    col = torch.full((B * Ho * Wo, 3 * kp), float("nan"), dtype=torch.bfloat16, device=dev)
    ops.stem_im2col_split3(x, col, k, s, kp)
    torch.cuda.synchronize()
    # rows: (kh, kw, c) order of the padded window
    unf = F.unfold(_tf_same_pad(x, k, s), k, stride=s).view(B, 3, k * k, Ho * Wo).permute(0, 3, 2, 1).reshape(B * Ho * Wo, 147)
    hi = unf.bfloat16()
    lo = (unf - hi.float()).bfloat16()
    assert torch.equal(col[:, :147], hi) and torch.equal(col[:, kp:kp + 147], hi) and torch.equal(col[:, 2 * kp:2 * kp + 147], lo)
    assert (col[:, 147:kp] == 0).all() and (col[:, kp + 147:2 * kp] == 0).all() and (col[:, 2 * kp + 147:] == 0).all()
    wp = torch.zeros((Cout, kp), dtype=torch.float32, device=dev)
    wp[:, :147] = w.permute(0, 2, 3, 1).reshape(Cout, -1)
    out = torch.empty((B * Ho * Wo, Cout), dtype=torch.float32, device=dev)
    ops.gemm(col, split3_conv1x1(wp), out)
    ref = F.conv2d(_tf_same_pad(x.double(), k, s), w.double(), None, stride=s).permute(0, 2, 3, 1).reshape(-1, Cout)
    torch.cuda.synchronize()
    assert _relerr(out, ref) < 3e-5, _relerr(out, ref)


@pytest.mark.parametrize("conv", [False, True])
def test_split3_gemm_precision(conv):
    """Activations [Ah | Ah | Al] x weights [Wh | Wl | Wh] on the unchanged bf16 GEMM / implicit 3x3 conv: the result
    must be ~2^-16-accurate against fp64, where plain bf16 operands give ~2^-8."""
    from depth_estimation_b200 import ops
    from depth_estimation_b200.weights import split3_conv1x1, split3_conv3x3
    dev = _dev()
    B, H, W, Cin, Cout = 2, 24, 24, 64, 64
    a = F.relu(_rand((B, H, W, Cin), dev, seed=71) + 0.5)
    hi = a.bfloat16()
    lo = (a - hi.float()).bfloat16()
    a3 = torch.cat([hi, hi, lo], dim=-1).contiguous()
    out = torch.empty((B, H, W, Cout), dtype=torch.float32, device=dev)
    if conv:
        w = _rand((Cout, Cin, 3, 3), dev, 0.1, seed=72)
        ops.gemm(a3, split3_conv3x3(w), out.view(-1, Cout), conv=(B, H, W, 3 * Cin), N=Cout)
        ref = F.conv2d(a.double().permute(0, 3, 1, 2), w.double(), padding=1).permute(0, 2, 3, 1)
        plain = F.conv2d(hi.double().permute(0, 3, 1, 2), w.bfloat16().double(), padding=1).permute(0, 2, 3, 1)
    else:
        w = _rand((Cout, Cin), dev, 0.1, seed=72)
        ops.gemm(a3.view(-1, 3 * Cin), split3_conv1x1(w), out.view(-1, Cout))
        ref = a.double() @ w.double().t()
        plain = hi.double() @ w.bfloat16().double().t()
    torch.cuda.synchronize()
    e3, e1 = _relerr(out, ref), _relerr(plain, ref)
    print(f"split3 {'conv' if conv else 'gemm'}: rel err {e3:.2e} (plain bf16 operands {e1:.2e})")
    assert e3 < 3e-5 and e3 < e1 / 30


@pytest.mark.parametrize("M,N,K,ks", [(577, 768, 3072, 0), (577, 768, 768, 4), (300, 256, 2304, 3), (128, 64, 512, 16)])
def test_gemm_splitk_accumulate(M, N, K, ks):
    """Split-K of an accumulate call (ViT proj / fc2 at batch 1): the finalize pass adds the K-slices in a fixed order
    into the fp32 stream.  k_splits 0 = automatic choice.  Bitwise reproducible run to run."""
    from depth_estimation_b200 import ops
    dev = _dev()
    a = _rand((M, K), dev, seed=91).bfloat16()
    w = _rand((N, K), dev, 0.05, seed=92).bfloat16()
    bias = _rand((N,), dev, seed=93)
    x0 = _rand((M, N), dev, seed=94)
    x1, x2, x3 = x0.clone(), x0.clone(), x0.clone()
    ws = torch.empty(32 << 20, dtype=torch.uint8, device=dev).fill_(0x7f)       # no initial contents required
    ops.gemm(a, w, x1, bias=bias, accumulate=True, k_splits=1)
    ops.gemm(a, w, x2, bias=bias, accumulate=True, k_splits=ks, splitk_ws=ws)
    ws.fill_(0xff)
    ops.gemm(a, w, x3, bias=bias, accumulate=True, k_splits=ks, splitk_ws=ws)
    ref = x0.double() + a.double() @ w.double().t() + bias.double()
    torch.cuda.synchronize()
    assert _relerr(x1, ref) < 1e-5 and _relerr(x2, ref) < 1e-5
    assert _relerr(x2, x1) < 1e-5 and torch.equal(x2, x3)


@pytest.mark.parametrize("case", ["dense_bias_gelu", "conv_residual_relu", "conv_split3", "pixel_shuffle"])
def test_gemm_splitk_workspace(case):
    """Split-K through the sliced workspace + finalize launch reproduces the single-pass epilogue (bias, activation,
    fp32 / bf16 skip inputs, ReLU copy, split3 stores, pixel shuffle)."""
    from depth_estimation_b200 import ops
    from depth_estimation_b200.weights import pack_conv3x3, pack_conv_transpose
    dev = _dev()
    ws = torch.empty(32 << 20, dtype=torch.uint8, device=dev).fill_(0x7f)
    if case == "dense_bias_gelu":
        M, N, K = 577, 768, 1536
        a, w, bias = _rand((M, K), dev, seed=95).bfloat16(), _rand((N, K), dev, 0.05, seed=96).bfloat16(), _rand((N,), dev, seed=97)
        o1 = torch.empty((M, N), dtype=torch.bfloat16, device=dev)
        o2 = torch.full_like(o1, float("nan"))
        ops.gemm(a, w, o1, bias=bias, act=ops.ACT_GELU)
        ops.gemm(a, w, o2, bias=bias, act=ops.ACT_GELU, k_splits=4, splitk_ws=ws)
    elif case in ("conv_residual_relu", "conv_split3"):
        B, H, W, Cin, Cout = 2, 12, 12, 256, 64
        a = _rand((B, H, W, Cin), dev, seed=98).bfloat16()
        wp = pack_conv3x3(_rand((Cout, Cin, 3, 3), dev, 0.03, seed=99))
        bias, add1 = _rand((Cout,), dev, seed=100), _rand((B, H, W, Cout), dev, seed=101)
        add2 = _rand((B, H, W, Cout), dev, seed=102)
        if case == "conv_residual_relu":
            o1 = torch.empty((B, H, W, Cout), dtype=torch.float32, device=dev)
            o2 = torch.full_like(o1, float("nan"))
            r1 = torch.empty((B, H, W, 3 * Cout), dtype=torch.bfloat16, device=dev)
            r2 = torch.full_like(r1, float("nan"))
            for o, r, kw in ((o1, r1, {}), (o2, r2, dict(k_splits=6, splitk_ws=ws))):
                ops.gemm(a, wp, o.view(-1, Cout), conv=(B, H, W, Cin), N=Cout, bias=bias, add1=add1.view(-1, Cout), add2=add2,
                         out_relu=r, relu_split=True, **kw)
            torch.cuda.synchronize()
            assert _relerr(r2, r1) < 1e-2 and torch.isfinite(r2.float()).all()
        else:
            o1 = torch.empty((B, H, W, 3 * Cout), dtype=torch.bfloat16, device=dev)
            o2 = torch.full_like(o1, float("nan"))
            for o, kw in ((o1, {}), (o2, dict(k_splits=0, splitk_ws=ws))):           # automatic: 3 row tiles, K = 36 blocks
                ops.gemm(a, wp, o.view(-1, 3 * Cout), conv=(B, H, W, Cin), N=Cout, bias=bias, act=ops.ACT_RELU,
                         add1=add1.bfloat16().view(-1, Cout), out_split=True, **kw)
    else:
        B, gh, gw, s_, cin, cout = 2, 6, 5, 2, 512, 64
        x = _rand((B * gh * gw, cin), dev, seed=103).bfloat16()
        wt, bt = _rand((cin, cout, s_, s_), dev, 0.05, seed=104), _rand((cout,), dev, seed=105)
        o1 = torch.empty((B, gh * s_, gw * s_, cout), dtype=torch.bfloat16, device=dev)
        o2 = torch.full_like(o1, float("nan"))
        for o, kw in ((o1, {}), (o2, dict(k_splits=4, splitk_ws=ws))):
            ops.gemm(x, pack_conv_transpose(wt), o, bias=bt, bias_mod=cout, pixel_shuffle=(gh, gw, s_, cout), **kw)
    torch.cuda.synchronize()
    assert torch.isfinite(o2.float()).all()
    assert _relerr(o2, o1) < (1e-2 if o1.dtype == torch.bfloat16 else 1e-5), _relerr(o2, o1)


def _split3_ref(v):
    hi = v.bfloat16()
    lo = (v - hi.float()).bfloat16()
    return torch.cat([hi, hi, lo], dim=-1)


@pytest.mark.parametrize("pixel_shuffle", [False, True])
def test_gemm_split3_outputs(pixel_shuffle):
    """General epilogue writing split3 rows: `out` as [hi | hi | lo] of v, `out_relu` as the split3 of relu(v) next to
    an fp32 `out`, fp32 add1 + fp32 add2 skip inputs - the plumbing of the bf16x3 decoder."""
    from depth_estimation_b200 import ops
    dev = _dev()
    B, H, W, Cin, Cout = 2, 12, 20, 128, 64
    a = _rand((B, H, W, Cin), dev, seed=81).bfloat16()
    w = _rand((Cout, Cin, 3, 3), dev, 0.05, seed=82)
    bias = _rand((Cout,), dev, seed=83)
    add1 = _rand((B, H, W, Cout), dev, seed=84)
    add2 = _rand((B, H, W, Cout), dev, seed=85)
    from depth_estimation_b200.weights import pack_conv3x3
    wp = pack_conv3x3(w)
    ref = F.conv2d(a.float().permute(0, 3, 1, 2), w.bfloat16().float(), bias, padding=1).permute(0, 2, 3, 1) + add1 + add2
    if not pixel_shuffle:
        # (1) fp32 out + split3 relu copy
        o32 = torch.empty((B, H, W, Cout), dtype=torch.float32, device=dev)
        r3 = torch.full((B, H, W, 3 * Cout), float("nan"), dtype=torch.bfloat16, device=dev)
        ops.gemm(a, wp, o32.view(-1, Cout), conv=(B, H, W, Cin), N=Cout, bias=bias, add1=add1.view(-1, Cout), add2=add2,
                 out_relu=r3, relu_split=True)
        torch.cuda.synchronize()
        assert _relerr(o32, ref) < 1e-5
        assert torch.equal(r3, _split3_ref(F.relu(o32)))
        # (2) split3 out (+ plain bf16 relu copy)
        o3 = torch.full((B, H, W, 3 * Cout), float("nan"), dtype=torch.bfloat16, device=dev)
        rp = torch.empty((B, H, W, Cout), dtype=torch.bfloat16, device=dev)
        ops.gemm(a, wp, o3.view(-1, 3 * Cout), conv=(B, H, W, Cin), N=Cout, bias=bias, add1=add1.view(-1, Cout), add2=add2,
                 out_split=True, out_relu=rp)
        torch.cuda.synchronize()
        assert torch.equal(o3, _split3_ref(o32)) and torch.equal(rp, F.relu(o32).bfloat16())
        hi, lo = o3[..., :Cout].float(), o3[..., 2 * Cout:].float()
        assert _relerr(hi + lo, ref) < 3e-5                        # hi + lo carries ~16 mantissa bits
        with pytest.raises(RuntimeError):
            ops.gemm(a, wp, o32.view(-1, Cout), conv=(B, H, W, Cin), N=Cout, out_split=True)
    else:
        # ConvTranspose(k == s) pixel shuffle into split3 pixels [B, gh*s, gw*s, 3*cout]
        gh, gw, s_, cin, cout = 6, 5, 2, 64, 32
        x = _rand((B * gh * gw, cin), dev, seed=86).bfloat16()
        wt = _rand((cin, cout, s_, s_), dev, 0.1, seed=87)
        bt = _rand((cout,), dev, seed=88)
        from depth_estimation_b200.weights import pack_conv_transpose
        o3 = torch.full((B, gh * s_, gw * s_, 3 * cout), float("nan"), dtype=torch.bfloat16, device=dev)
        ops.gemm(x, pack_conv_transpose(wt), o3, bias=bt, bias_mod=cout, pixel_shuffle=(gh, gw, s_, cout), out_split=True)
        ref = F.conv_transpose2d(x.float().view(B, gh, gw, cin).permute(0, 3, 1, 2), wt.bfloat16().float(), bt, stride=s_)
        ref = ref.permute(0, 2, 3, 1)
        torch.cuda.synchronize()
        assert torch.isfinite(o3.float()).all()
        hi, lo = o3[..., :cout].float(), o3[..., 2 * cout:].float()
        assert torch.equal(o3[..., :cout], o3[..., cout:2 * cout])
        assert _relerr(hi + lo, ref) < 3e-5


def test_split3_vit_operators():
    """LayerNorm / token cast / attention / patchify with split3 output rows: hi == the plain bf16 result, hi + lo carries
    the fp32 value to ~2^-16."""
    from depth_estimation_b200 import ops
    dev = _dev()
    M, D = 300, 768
    x = _rand((M, D), dev, 2.0, seed=110)
    g, b = _rand((D,), dev, seed=111), _rand((D,), dev, seed=112)
    plain = torch.empty((M, D), dtype=torch.bfloat16, device=dev)
    ops.layernorm(x, g, b, 1e-6, plain)
    o3 = torch.full((M, 3 * D), float("nan"), dtype=torch.bfloat16, device=dev)
    ops.layernorm_split3(x, g, b, 1e-6, o3)
    ref = F.layer_norm(x, (D,), g, b, 1e-6)
    torch.cuda.synchronize()
    assert torch.equal(o3[:, :D], plain) and torch.equal(o3[:, D:2 * D], plain)
    assert _relerr(o3[:, :D].float() + o3[:, 2 * D:].float(), ref) < 2e-5
    # token cast (drop the first row of every group of 10)
    c3 = torch.full((270, 3 * D), float("nan"), dtype=torch.bfloat16, device=dev)
    ops.cast_tokens_split3(x, c3, grp=10, skip=1)
    want = x.view(30, 10, D)[:, 1:].reshape(270, D)
    torch.cuda.synchronize()
    assert torch.equal(c3, _split3_ref(want))
    # attention
    B, T, heads = 2, 197, 4
    Dh = heads * 64
    qkv = _rand((B * T, 3 * Dh), dev, 1.0, seed=113).bfloat16()
    a1 = torch.empty((B * T, Dh), dtype=torch.bfloat16, device=dev)
    a3 = torch.full((B * T, 3 * Dh), float("nan"), dtype=torch.bfloat16, device=dev)
    ops.attention(qkv, a1, B, T, heads, 0.125)
    ops.attention(qkv, a3, B, T, heads, 0.125, split=True)
    q, k, v = qkv.float().view(B, T, 3, heads, 64).permute(2, 0, 3, 1, 4)
    aref = F.scaled_dot_product_attention(q, k, v).transpose(1, 2).reshape(B * T, Dh)
    torch.cuda.synchronize()
    assert torch.equal(a3[:, :Dh], a1) and torch.equal(a3[:, Dh:2 * Dh], a1)
    e1, e3 = _relerr(a1, aref), _relerr(a3[:, :Dh].float() + a3[:, 2 * Dh:].float(), aref)
    assert torch.isfinite(a3.float()).all() and e3 <= e1 + 1e-6
    # patchify: fp32 NCHW -> split3 patch matrix (column = c*256 + py*16 + px)
    img = _rand((2, 3, 32, 48), dev, seed=114)
    p3 = torch.full((2 * 2 * 3, 3 * 768), float("nan"), dtype=torch.bfloat16, device=dev)
    ops.patchify_split3(img, p3)
    want = img.view(2, 3, 2, 16, 3, 16).permute(0, 2, 4, 1, 3, 5).reshape(12, 768)
    torch.cuda.synchronize()
    assert torch.equal(p3, _split3_ref(want))


@pytest.mark.parametrize("ac", [True, False])
def test_upsample2x_ex(ac):
    from depth_estimation_b200 import ops
    dev = _dev()
    x = _rand((2, 13, 21, 64), dev, seed=90)
    ref = F.interpolate(x.permute(0, 3, 1, 2), scale_factor=2, mode="bilinear", align_corners=ac).permute(0, 2, 3, 1).contiguous()
    o32 = torch.empty((2, 26, 42, 64), dtype=torch.float32, device=dev)
    ops.upsample2x_ex(x, o32, ac)
    o3 = torch.empty((2, 26, 42, 192), dtype=torch.bfloat16, device=dev)
    ops.upsample2x_ex(x, o3, ac, split=True)
    ob = torch.empty((2, 26, 42, 64), dtype=torch.bfloat16, device=dev)
    ops.upsample2x_ex(x.bfloat16(), ob, ac)
    torch.cuda.synchronize()
    assert _relerr(o32, ref) < 1e-6
    assert torch.equal(o3, _split3_ref(o32))
    assert _relerr(ob, ref) < 1e-2


# ------------------------------------------------------------------ NV12 ingest, calibrated metric scale
@pytest.mark.parametrize("H,W,mt,layout", [(480, 640, "MiDaS_small", 0), (360, 642, "DPT_Large", 1), (384, 384, "DPT_Large", 1),
                                           (256, 256, "MiDaS_small", 2)])
def test_preprocess_nv12_matches_bgr_path(H, W, mt, layout):
    """K1 with the NV12 -> BGR conversion fused == cv2.cvtColor(COLOR_YUV2BGR_NV12) followed by K1, bit for bit
    (resampling path, ragged width, and the identity-size copy path)."""
    import cv2
    from depth_estimation_b200 import ops
    dev = _dev()
    rng = np.random.default_rng(7)
    nv12 = rng.integers(0, 256, (2, H * 3 // 2, W), dtype=np.uint8)
    bgr = np.stack([cv2.cvtColor(f, cv2.COLOR_YUV2BGR_NV12) for f in nv12])
    nw, nh = ops.get_size(W, H, mt)
    mean, std = ((0.5,) * 3, (0.5,) * 3) if mt != "MiDaS_small" else ((0.485, 0.456, 0.406), (0.229, 0.224, 0.225))
    t = torch.from_numpy(nv12).to(dev)
    y, uv = t[:, :H], t[:, H:].reshape(2, H // 2, W // 2, 2)
    a = ops.preprocess_nv12(y, uv, nh, nw, mean, std, layout)
    b = ops.preprocess(torch.from_numpy(bgr).to(dev), nh, nw, mean, std, layout)
    torch.cuda.synchronize()
    assert a.shape == b.shape and torch.equal(a, b)


def test_postprocess_metric_affine():
    from depth_estimation_b200 import ops
    dev = _dev()
    pred = _rand((2, 96, 128), dev, seed=90).abs() * 10
    omap, ometric, _ = ops.postprocess(pred, 240, 320, want_metric=True, metric_scale=0.37, metric_shift=-1.25)
    torch.cuda.synchronize()
    assert _relerr(ometric, omap * 0.37 - 1.25) < 8e-3          # bf16 rounding of fmaf(map, scale, shift)
