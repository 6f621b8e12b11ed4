"""ctypes binding of libdepthb200.so (declared in include/depth_b200.h).

The product path has NO fallback: if the shared library is missing or a call fails, a RuntimeError is
raised.  torch is used only for device memory / streams; every entry point takes raw pointers.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# DEPTH_B200_LIB points at another build of the same library (kernel tuning variants); there is still no fallback
LIB_PATH = os.environ.get("DEPTH_B200_LIB") or os.path.join(_HERE, "libdepthb200.so")

DE_DTYPE_BF16, DE_DTYPE_F32 = 0, 1
ACT_NONE, ACT_RELU, ACT_GELU, ACT_RELU6 = 0, 1, 2, 3


class DeGemm(C.Structure):
    _fields_ = [
        ("a_mode", C.c_int), ("a", C.c_void_p), ("lda", C.c_int64),
        ("M", C.c_int), ("N", C.c_int), ("K", C.c_int),
        ("B", C.c_int), ("H", C.c_int), ("W", C.c_int), ("Cin", C.c_int),
        ("w", C.c_void_p), ("ldw", C.c_int64),
        ("bias", C.c_void_p), ("bias_mod", C.c_int), ("act", C.c_int),
        ("add1", C.c_void_p), ("add1_dtype", C.c_int), ("ld_add1", C.c_int64),
        ("add1_rowmode", C.c_int), ("add1_pre", C.c_int),
        ("add2", C.c_void_p),
        ("out", C.c_void_p), ("out_dtype", C.c_int), ("ldo", C.c_int64),
        ("out_relu", C.c_void_p),
        ("out_rowmode", C.c_int), ("grp", C.c_int), ("pad", C.c_int), ("off", C.c_int),
        ("ps_gh", C.c_int), ("ps_gw", C.c_int), ("ps_s", C.c_int), ("ps_cout", C.c_int),
        ("head_w", C.c_void_p), ("head_b", C.c_float),
        ("block_n", C.c_int), ("epilogue", C.c_int), ("cta_group", C.c_int),
        ("out_split", C.c_int), ("relu_split", C.c_int), ("ld_relu", C.c_int64), ("add2_dtype", C.c_int), ("ld_add2", C.c_int64),
        ("k_splits", C.c_int), ("splitk_ws", C.c_void_p), ("splitk_ws_bytes", C.c_int64),
        ("accumulate", C.c_int),
    ]


class DePredictIO(C.Structure):
    _fields_ = [
        ("frames", C.c_void_p), ("row_stride", C.c_int64), ("frame_stride", C.c_int64),
        ("uv", C.c_void_p), ("uv_row_stride", C.c_int64), ("uv_frame_stride", C.c_int64),
        ("out_map", C.c_void_p), ("out_metric", C.c_void_p), ("metric_scale", C.c_float), ("metric_shift", C.c_float),
        ("out_vis", C.c_void_p), ("sample_rc", C.c_void_p), ("n_samples", C.c_int), ("out_samples", C.c_void_p),
    ]


class DeProfileRecord(C.Structure):
    _fields_ = [("name", C.c_char * 24), ("ms", C.c_double), ("flops", C.c_double), ("dims", C.c_int64 * 4)]


DE_PRECISION_DEFAULT, DE_PRECISION_BF16, DE_PRECISION_BF16X3 = 0, 1, 2

# name -> (restype, argtypes); must list every symbol include/depth_b200.h declares
SIGNATURES = {
    "de_last_error": (C.c_char_p, []),
    "de_version": (C.c_int, []),
    "de_launch_count": (C.c_int64, []),
    "de_create": (C.c_int, [C.c_char_p, C.c_int, C.c_int, C.POINTER(C.c_void_p)]),
    "de_set_weight": (C.c_int, [C.c_void_p, C.c_char_p, C.c_void_p, C.POINTER(C.c_int64), C.c_int, C.c_int]),
    "de_finalize_weights": (C.c_int, [C.c_void_p]),
    "de_plan": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "de_predict": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(DePredictIO), C.c_int, C.c_void_p]),
    "de_forward": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                             C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_int, C.c_void_p]),
    "de_engine_info": (C.c_int, [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "de_destroy": (None, [C.c_void_p]),
    "de_profile_start": (C.c_int, []),
    "de_profile_stop": (C.c_int, [C.POINTER(DeProfileRecord), C.c_int, C.POINTER(C.c_int)]),
    "de_get_size": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "de_cubic_taps": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "de_preprocess": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int64, C.c_int64, C.c_int, C.c_int,
                                C.POINTER(C.c_float), C.POINTER(C.c_float), C.c_int, C.c_void_p, C.c_void_p]),
    "de_postprocess": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                 C.c_void_p, C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p,
                                 C.c_void_p]),
    "de_preprocess_nv12": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int64, C.c_int64, C.c_int64,
                                     C.c_int64, C.c_int, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_float), C.c_int,
                                     C.c_void_p, C.c_void_p]),
    "de_gemm": (C.c_int, [C.POINTER(DeGemm), C.c_void_p]),
    "de_layernorm": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_int, C.c_int64, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                               C.c_float, C.c_void_p, C.c_int64, C.c_void_p]),
    "de_layernorm_split3": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_void_p,
                                      C.c_int64, C.c_void_p]),
    "de_attention_split3": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_void_p]),
    "de_cast_tokens_split3": (C.c_int, [C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "de_patchify_split3": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "de_attention": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_void_p]),
    "de_cast_tokens": (C.c_int, [C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "de_fill_cls": (C.c_int, [C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "de_upsample2x": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "de_upsample2x_ex": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                   C.c_void_p]),
    "de_im2col3x3": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                               C.c_int, C.c_int, C.c_void_p]),
    "de_subsample": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "de_groupnorm_ws_bytes": (C.c_int64, [C.c_int, C.c_int, C.c_int, C.c_int]),
    "de_groupnorm": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int,
                               C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int,
                               C.c_void_p, C.c_int64, C.c_void_p]),
    "de_groupnorm_static_max_bytes": (C.c_int64, []),
    "de_groupnorm_static": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int,
                                      C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int,
                                      C.c_void_p]),
    "de_head_conv": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_float, C.c_void_p, C.c_int,
                               C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "de_head_up_conv": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_float, C.c_void_p, C.c_int,
                                  C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "de_stem_conv": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                               C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "de_stem_im2col_split3": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                        C.c_void_p]),
    "de_depthwise": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                               C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "de_dwproject_ws_bytes": (C.c_int64, [C.c_int, C.c_int, C.c_int, C.c_int]),
    "de_dwproject": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                               C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int64, C.c_void_p]),
    "de_maxpool3x3s2": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                  C.c_void_p]),
}

_lib = None


def load() -> C.CDLL:
    """Load (once) and type the shared library. Raises RuntimeError when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing - build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or depth_estimation_b200/csrc/build.sh). There is no CPU or PyTorch fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().de_last_error().decode(errors="replace")
        raise RuntimeError(f"libdepthb200 {what} failed ({rc}): {msg}")


def profile_start() -> None:
    check(load().de_profile_start(), "de_profile_start")


def profile_stop(max_records: int = 65536):
    """[(name, ms, flops, (d0, d1, d2, d3))] for every operator call since profile_start (synchronises the device)."""
    buf = (DeProfileRecord * max_records)()
    n = C.c_int()
    check(load().de_profile_stop(buf, max_records, C.byref(n)), "de_profile_stop")
    return [(buf[i].name.decode(), buf[i].ms, buf[i].flops, tuple(buf[i].dims)) for i in range(n.value)]


def launch_count() -> int:
    return int(load().de_launch_count())
