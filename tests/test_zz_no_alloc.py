"""de_groupnorm_static - the allocation-free GroupNorm entry point (include/depth_b200.h; no counterpart in the
reference) - against de_groupnorm with a caller workspace.

Collected last on purpose (file name): the round's end runs `pytest -x`, and this file must not hide the others.
Works alone in a fresh process and after the other GPU tests (it creates nothing they share; the built-in scratch array
is re-zeroed by every call)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

ACT_RELU = 1


def _rand(shape, dev, scale=1.0, seed=0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(shape, generator=g) * scale).to(dev)


def _call_static(lib, x, g, b, add, out, o32, act, stream):
    B, H, W, Cc = x.shape
    dt = lambda t: 1 if t.dtype == torch.float32 else 0                 # DE_DTYPE_BF16 = 0, DE_DTYPE_F32 = 1
    return lib.de_groupnorm_static(x.data_ptr(), dt(x), g.data_ptr(), b.data_ptr(), add.data_ptr() if add is not None else None,
                                   dt(add) if add is not None else 0, out.data_ptr() if out is not None else None, 0,
                                   o32.data_ptr() if o32 is not None else None, None, B, H * W, Cc, 32, 1e-5, act, stream)


def test_groupnorm_static_matches_groupnorm_two_calls_in_a_row():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from depth_estimation_b200 import _lib, ops
    lib = _lib.load()
    assert _lib.DE_DTYPE_BF16 == 0 and _lib.DE_DTYPE_F32 == 1
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    limit = int(lib.de_groupnorm_static_max_bytes())
    assert limit == 4 << 20
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    # two shapes back to back, the larger first, so the second call finds the first one's partial sums / statistics in
    # the shared array: (B, H, W, C, bf16 input?, residual add?)  - the one-launch path (small grid) and the two-kernel path
    cases = [(6, 48, 48, 256, True, True), (2, 24, 40, 64, False, False), (6, 48, 48, 256, True, True)]
    prepared = []
    for i, (B, H, W, Cc, bf16_in, with_add) in enumerate(cases):
        x = _rand((B, H, W, Cc), dev, 2.0, seed=200 + i) + 0.7
        if bf16_in:
            x = x.bfloat16()
        g, b = _rand((Cc,), dev, seed=210 + i), _rand((Cc,), dev, seed=220 + i)
        add = _rand((B, H, W, Cc), dev, seed=230 + i).bfloat16() if with_add else None
        need = ops.groupnorm_ws_bytes(B, H * W, Cc)
        assert need <= limit
        ws = torch.zeros(need, dtype=torch.uint8, device=dev)
        want = torch.empty((B, H, W, Cc), dtype=torch.bfloat16, device=dev)
        want32 = torch.empty((B, H, W, Cc), dtype=torch.float32, device=dev)
        ops.groupnorm(x, g, b, want, ws, act=ACT_RELU, add=add, out_f32=want32)          # the existing entry point
        got = torch.full_like(want, float("nan"))
        got32 = torch.full_like(want32, float("nan"))
        prepared.append((x, g, b, add, want, want32, got, got32))
    torch.cuda.synchronize()
    # warm call (module / kernel loading), then the measured pair: no device memory may be taken by the calls
    x, g, b, add, _, _, got, got32 = prepared[0]
    assert _call_static(lib, x, g, b, add, got, got32, ACT_RELU, stream) == 0, lib.de_last_error()
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info(dev)
    alloc0 = torch.cuda.memory_allocated(dev)
    for x, g, b, add, _, _, got, got32 in prepared:                     # calls in a row on one stream, no sync between
        got.fill_(float("nan"))
        got32.fill_(float("nan"))
        assert _call_static(lib, x, g, b, add, got, got32, ACT_RELU, stream) == 0, lib.de_last_error()
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(dev)
    assert torch.cuda.memory_allocated(dev) == alloc0 and free1 >= free0, (free0, free1)
    for _, _, _, _, want, want32, got, got32 in prepared:
        # same kernels, fixed-order reductions: the tolerance of tests/test_kernels_gpu.py::test_groupnorm_* (1e-5 of
        # max on the fp32 output, bf16 output equal to its rounding) - and in fact the same bits
        assert torch.isfinite(got32).all()
        err = float((got32 - want32).abs().max() / want32.abs().max())
        assert err < 1e-5, err
        assert torch.equal(got, got32.bfloat16())
        assert torch.equal(got32, want32) and torch.equal(got, want)


def test_groupnorm_static_rejects_oversize_and_bad_shapes():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from depth_estimation_b200 import _lib, ops
    lib = _lib.load()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    B, H, W, Cc = 128, 96, 96, 64                                       # needs 4.77 MB of partial sums > 4 MiB
    assert ops.groupnorm_ws_bytes(B, H * W, Cc) > int(lib.de_groupnorm_static_max_bytes())
    x = torch.zeros((B, H, W, Cc), dtype=torch.bfloat16, device=dev)
    g = torch.ones(Cc, device=dev)
    out = torch.full_like(x, 7.0)
    rc = _call_static(lib, x, g, g, None, out, None, ACT_RELU, stream)
    msg = lib.de_last_error().decode()
    assert rc == -3 and "DE_GROUPNORM_STATIC_MAX_BYTES" in msg and "de_groupnorm" in msg, (rc, msg)
    torch.cuda.synchronize()
    assert bool((out == 7.0).all())                                     # nothing was launched
    x2 = torch.zeros((1, 4, 4, 36), dtype=torch.bfloat16, device=dev)   # 36 channels: no 8-channel vectors per group
    rc = _call_static(lib, x2, g, g, None, torch.empty_like(x2), None, ACT_RELU, stream)
    assert rc != 0
