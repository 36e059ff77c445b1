"""ctypes binding of libcan_b200.so (C ABI declared in include/can_b200.h).

There is no fallback: if the library is missing or a call fails, an exception is
raised.  PyTorch is used only for device memory and streams.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# CAN_B200_LIB: another build of the same library (A/B measurements of two builds inside one GPU session)
LIB_PATH = os.environ.get("CAN_B200_LIB") or os.path.join(_HERE, "libcan_b200.so")

MODE_STORE, MODE_POOL, MODE_HEAD, MODE_UNPOOL, MODE_CTX, MODE_CTX_BWD, MODE_DFV, MODE_CTX2, MODE_CTX2_BWD = range(9)


class ConvArgs(C.Structure):
    """Mirror of ``can_conv_args`` (include/can_b200.h)."""
    _fields_ = [
        ("N", C.c_int32), ("H", C.c_int32), ("W", C.c_int32),
        ("Cin", C.c_int32), ("Cout", C.c_int32),
        ("taps", C.c_int32), ("dil", C.c_int32), ("nsplit", C.c_int32),
        ("mode", C.c_int32), ("relu", C.c_int32), ("bn", C.c_int32), ("stages", C.c_int32),
        ("x", C.c_void_p), ("x_plane", C.c_int64), ("x_cs", C.c_int32), ("x_coff", C.c_int32),
        ("w", C.c_void_p), ("bias", C.c_void_p),
        ("out", C.c_void_p), ("out_plane", C.c_int64), ("out_cs", C.c_int32), ("out_coff", C.c_int32),
        ("out_f32", C.c_void_p),
        ("mask", C.c_void_p), ("mask_cs", C.c_int32), ("mask_coff", C.c_int32), ("mask_plane", C.c_int64),
        ("add", C.c_void_p), ("add_plane", C.c_int64), ("add_cs", C.c_int32), ("add_coff", C.c_int32),
        ("arg", C.c_void_p), ("arg_cs", C.c_int32),
        ("OH", C.c_int32), ("OW", C.c_int32),
        ("head_w", C.c_void_p), ("head_b", C.c_void_p), ("head_out", C.c_void_p),
        ("tabA", C.c_void_p), ("tabB", C.c_void_p),
        ("ctx_wint", C.c_void_p), ("ctx_ttab", C.c_void_p),
        ("db", C.c_void_p),
    ]


class PackItem(C.Structure):
    """can_pack_item of include/can_b200.h"""
    _fields_ = [("w", C.c_void_p), ("dst_fwd", C.c_void_p), ("dst_dgrad", C.c_void_p), ("cout", C.c_int), ("cin", C.c_int),
                ("taps", C.c_int)]


_SIGNATURES = {
    "can_last_error_string": (C.c_char_p, []),
    "can_version": (C.c_int, []),
    "can_conv_tc": (C.c_int, [C.POINTER(ConvArgs), C.c_void_p]),
    "can_pack_conv_weight": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "can_pack_conv_weight_both": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                            C.c_void_p]),
    "can_pack_conv_weights_multi": (C.c_int, [C.POINTER(PackItem), C.c_int, C.c_int, C.c_void_p]),
    "can_pack_ctx_w2": (C.c_int, [C.c_void_p] * 5 + [C.c_int, C.c_void_p]),
    "can_transpose_f32": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "can_transpose_f32_multi": (C.c_int, [C.POINTER(C.c_void_p), C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "can_conv_first_fwd": (C.c_int, [C.c_void_p] * 4 + [C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "can_ctx_pool_workspace_bytes": (C.c_int64, [C.c_int, C.c_int]),
    "can_ctx_pool_fwd": (C.c_int, [C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
                         + [C.c_void_p] * 7),
    "can_mse_count_workspace_bytes": (C.c_int64, [C.c_int]),
    "can_mse_count_fwd_bwd": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 4
                              + [C.c_float, C.c_void_p, C.c_void_p]),
    "can_mse_count_static_max_images": (C.c_int, []),
    "can_mse_count_fwd_bwd_static": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 4
                                     + [C.c_float, C.c_void_p]),
    "can_zero_ranges": (C.c_int, [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int32), C.c_int, C.c_void_p]),
    "can_sgd_momentum_step": (C.c_int, [C.c_void_p] * 3 + [C.c_int64, C.c_float, C.c_float, C.c_float, C.c_int,
                                                          C.c_void_p]),
    "can_conv_wgrad_workspace_bytes": (C.c_int64, [C.c_int] * 7),
    "can_conv_wgrad": (C.c_int, [C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_void_p, C.c_int64, C.c_int, C.c_int]
                       + [C.c_int] * 8 + [C.c_void_p, C.c_int, C.c_float, C.c_void_p, C.c_int64, C.c_void_p]),
    "can_ctx_wint": (C.c_int, [C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "can_ctx_ttab": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    "can_ctx_pint": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "can_ctx_dtab": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    "can_ctx_pos_gemv": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p,
                                   C.c_int, C.c_void_p]),
    "can_ctx_outer": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_int, C.c_void_p,
                                C.c_float, C.c_int, C.c_void_p]),
    "can_ctx_ut_workspace_bytes": (C.c_int64, [C.c_int, C.c_int, C.c_int, C.c_int]),
    "can_ctx_ut": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                             C.c_void_p, C.c_int64, C.c_void_p]),
    "can_crowd_sample": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                   C.c_void_p]),
    "can_density_map_workspace_bytes": (C.c_int64, [C.c_int]),
    "can_density_map": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "can_im2col_first": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "can_conv_first_tc_fwd": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                        C.c_int, C.c_int, C.c_void_p]),
    "can_conv_first_wgrad_workspace_bytes": (C.c_int64, []),
    "can_conv_first_wgrad": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                       C.c_int, C.c_float, C.c_void_p, C.c_int64, C.c_void_p]),
    "can_bias_grad": (C.c_int, [C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int64, C.c_void_p,
                                C.c_void_p]),
    "can_head_bwd": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_void_p, C.c_int64,
                               C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
    "can_bilinear_taps": (C.c_int, [C.c_int, C.c_int] + [C.c_void_p] * 5),
    "can_adaptive_windows": (C.c_int, [C.c_int, C.c_int] + [C.c_void_p] * 3),
    "can_nhwc_to_nchw_f32": (C.c_int, [C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_void_p]
                             + [C.c_int] * 4 + [C.c_void_p]),
    "can_nchw_f32_to_nhwc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_int]
                             + [C.c_int] * 4 + [C.c_void_p]),
}

_lib = None


class CanB200Error(RuntimeError):
    pass


def exported_symbols():
    """Names every build of the library must export (checked by the CPU tests)."""
    return sorted(_SIGNATURES)


def lib():
    """Load the library once; raise loudly when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise CanB200Error(
                f"{LIB_PATH} not found: build it with `python __graft_entry__.py build` "
                "(or make -C can-distributed-pytorch_b200/csrc). There is no CPU/PyTorch fallback.")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().can_last_error_string().decode("utf-8", "replace")
        raise CanB200Error(f"{what or 'libcan_b200'} failed (rc={rc}): {msg}")
