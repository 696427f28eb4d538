"""ctypes binding of libfsb200.so (the C ABI declared in include/fsb200.h).

There is no fallback of any kind: if the shared library is missing the import
of an op raises, and every compute entry point needs a CUDA device.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_float, c_int, c_size_t, c_uint64, c_void_p

PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG, "libfsb200.so")

FSB_F32, FSB_F16, FSB_BF16 = 0, 1, 2
FSB_PREC_FP32, FSB_PREC_TF32, FSB_PREC_F16, FSB_PREC_BF16, FSB_PREC_TF32X3 = 0, 1, 2, 3, 4
FSB_MAX_LEVELS = 6
FSB_MAX_SRC = 4
FSB_EPI_BIAS, FSB_EPI_RELU, FSB_EPI_GATE_ZR, FSB_EPI_GATE_Q = 0, 1, 2, 3
FSB_EPI_GELU, FSB_EPI_SCALE_RES = 4, 5
FSB_CONV_LAYOUT_TAP, FSB_CONV_LAYOUT_HALO = 0, 1


class FsbError(RuntimeError):
    pass


class ConvSrc(ctypes.Structure):
    """fsb_conv_src (include/fsb200.h)"""
    _fields_ = [("ptr", c_void_p), ("C", c_int)]


class ConvSeg(ctypes.Structure):
    """fsb_conv_seg"""
    _fields_ = [("ksize", c_int), ("nsrc", c_int), ("src", c_int * FSB_MAX_SRC)]


class ConvArgs(ctypes.Structure):
    """fsb_conv_args -- field order and types follow include/fsb200.h exactly"""
    _fields_ = [
        ("src", ConvSrc * FSB_MAX_SRC),
        ("nsrc", c_int),
        ("B", c_int), ("H", c_int), ("W", c_int),
        ("seg", ConvSeg * 2),
        ("nseg", c_int),
        ("weight", c_void_p),
        ("bias", c_void_p),
        ("rows", c_int),
        ("Cout", c_int),
        ("block_n", c_int),
        ("stages", c_int),
        ("layout", c_int),
        ("epilogue", c_int),
        ("out", c_void_p),
        ("ld_out", c_int),
        ("ch_off", c_int),
        ("out_f32", c_void_p),
        ("out_scale", c_float),
        ("res", c_void_p),
        ("ld_res", c_int),
        ("gamma", c_void_p),
        ("h32", c_void_p),
        ("att", c_void_p),
        ("z_l", c_void_p), ("z_s", c_void_p), ("rh_l", c_void_p), ("rh_s", c_void_p),
        ("h32_out", c_void_p),
        ("h16_out", c_void_p),
        ("hidden", c_int),
    ]


_PROTOS = {
    "fsb_version": (c_int, []),
    "fsb_last_error": (c_char_p, []),
    "fsb_launch_count": (c_uint64, []),
    "fsb_set_pdl": (c_int, [c_int]),
    "fsb_volume_gwc_concat": (c_int, [c_void_p] * 6 + [c_int] * 8 + [c_void_p]),
    "fsb_volume_combined": (c_int, [c_void_p] * 5 + [c_int] * 8 + [c_void_p]),
    "fsb_volume_stem": (c_int, [c_void_p] * 9 + [c_int] * 9 + [c_void_p]),
    "fsb_corr_pyramid": (c_int, [c_void_p, c_void_p, POINTER(c_void_p)] + [c_int] * 7 + [c_void_p]),
    "fsb_geo_pyramid": (c_int, [c_void_p, POINTER(c_void_p)] + [c_int] * 6 + [c_void_p]),
    "fsb_geo_pyramid_in": (c_int, [c_void_p, c_int, POINTER(c_void_p)] + [c_int] * 6 + [c_void_p]),
    "fsb_geo_lookup": (c_int, [POINTER(c_void_p), POINTER(c_void_p), c_void_p, c_void_p, c_void_p]
                       + [c_int] * 9 + [c_void_p]),
    "fsb_corr_band_geometry": (c_int, [c_int] * 3 + [POINTER(c_int)] * 3),
    "fsb_corr_pyramid_band": (c_int, [c_void_p, c_void_p, POINTER(c_void_p)] + [c_int] * 8 + [c_void_p]),
    "fsb_geo_lookup_band": (c_int, [POINTER(c_void_p), POINTER(c_void_p), c_void_p, c_void_p, c_void_p]
                            + [c_int] * 9 + [c_void_p]),
    "fsb_lookup_convc1": (c_int, [POINTER(c_void_p), POINTER(c_void_p)] + [c_void_p] * 5 + [c_int] * 10 + [c_void_p]),
    "fsb_conv_args_size": (c_size_t, []),
    "fsb_conv2d_nhwc": (c_int, [POINTER(ConvArgs), c_int, c_void_p]),
    "fsb_nchw_to_nhwc": (c_int, [c_void_p] * 3 + [c_int] * 6 + [c_void_p]),
    "fsb_nhwc_to_nchw": (c_int, [c_void_p] * 2 + [c_int] * 4 + [c_void_p]),
    "fsb_pool2x_nhwc": (c_int, [c_void_p] * 2 + [c_int] * 5 + [c_void_p]),
    "fsb_resize_nhwc": (c_int, [c_void_p] * 2 + [c_int] * 7 + [c_void_p]),
    "fsb_pool2x_nhwc16": (c_int, [c_void_p] * 2 + [c_int] * 5 + [c_void_p]),
    "fsb_resize_nhwc16": (c_int, [c_void_p] * 2 + [c_int] * 7 + [c_void_p]),
    "fsb_conv7x7_c1": (c_int, [c_void_p] * 4 + [c_int] * 2 + [c_void_p] + [c_int] * 6 + [c_void_p]),
    "fsb_dwconv7x7_nhwc": (c_int, [c_void_p] * 4 + [c_int] * 5 + [c_void_p]),
    "fsb_edgenext_mlp": (c_int, [c_void_p] * 8 + [ctypes.c_longlong] + [c_int] * 3 + [c_void_p]),
    "fsb_conv3x3_cout1": (c_int, [c_void_p] * 5 + [c_int] * 5 + [c_void_p]),
    "fsb_edgenext_block": (c_int, [c_void_p] * 9 + [c_int] * 6 + [c_void_p]),
    "fsb_edgenext_block_noalloc": (c_int, [c_void_p] * 9 + [c_int] * 6 + [c_void_p]),
    "fsb_noalloc_max_pixels": (c_size_t, []),
    "fsb_axpy_f32": (c_int, [c_void_p, c_void_p, c_size_t, c_void_p]),
    "fsb_disparity_regression": (c_int, [c_void_p] * 2 + [c_int] * 6 + [c_void_p]),
    "fsb_context_upsample": (c_int, [c_void_p] * 3 + [c_int] * 4 + [c_float, c_int, c_void_p]),
}

_lib = None


def lib() -> ctypes.CDLL:
    """The loaded library; raises FsbError when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FsbError(
                f"{LIB_PATH} is missing: build it with `python -m foundationstereo_b200.build` "
                "(there is no CPU or PyTorch fallback for these ops)")
        L = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in _PROTOS.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        if L.fsb_conv_args_size() != ctypes.sizeof(ConvArgs):
            raise FsbError(f"fsb_conv_args is {L.fsb_conv_args_size()} bytes in {LIB_PATH} but "
                           f"{ctypes.sizeof(ConvArgs)} in the ctypes mirror: rebuild the library")
        _lib = L
    return _lib


def register_proto(name: str, restype, argtypes) -> None:
    """Used by modules that bind further entry points (update block)."""
    _PROTOS[name] = (restype, argtypes)
    if _lib is not None:
        fn = getattr(_lib, name)
        fn.restype = restype
        fn.argtypes = argtypes


def check(status: int) -> None:
    if status != 0:
        msg = lib().fsb_last_error()
        raise FsbError(f"libfsb200 status {status}: {msg.decode() if msg else '?'}")


def launch_count() -> int:
    return int(lib().fsb_launch_count())


def level_array(tensors):
    arr = (c_void_p * len(tensors))()
    for i, t in enumerate(tensors):
        arr[i] = t.data_ptr()
    return arr
