"""ctypes binding of `lib/librdp_b200.so` (the C ABI declared in `include/rdp_b200.h`)."""
from __future__ import annotations

import ctypes as C
import os
import threading

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(HERE), "lib", "librdp_b200.so")

OK = 0
FRAME_OK, FRAME_TOO_FEW_POINTS, FRAME_DEGENERATE_BINS, FRAME_TOO_FEW_EDGE, FRAME_FIT_INVALID, FRAME_NEEDS_KNOTS = range(6)
DEPTH_U16, DEPTH_F32, DEPTH_F32_WIDE, DEPTH_F64 = 0, 1, 2, 3
IN_RGB_U8_NHWC, IN_F32_NCHW = 0, 1
IN_FLAG_KEEP_ACTIVATIONS = 0x100
FIT_FLAG_KNOT_TIE, FIT_FLAG_BRACKET_TIE, FIT_FLAG_NPLUS_TIE = 1, 2, 4
ABI_VERSION = 3

EXPORTS = [
    "rdp_abi_version", "rdp_last_error",
    "rdp_unet_create", "rdp_unet_destroy", "rdp_unet_workspace_bytes", "rdp_unet_workspace_bytes_ex", "rdp_unet_forward", "rdp_unet_launch_count",
    "rdp_unet_read_activation", "rdp_unet_forward_profiled",
    "rdp_geom_workspace_bytes", "rdp_geom_curvature", "rdp_geom_launch_count", "rdp_geom_curvature_profiled",
    "rdp_geom_deproject_static",
    "rdp_pre_resize", "rdp_post_mask_nearest", "rdp_post_mask_png_sizes", "rdp_post_mask_png",
    "rdp_pre_decode_sizes", "rdp_pre_decode_jpeg", "rdp_pre_decode_png16",
    "rdp_unet_train_pack", "rdp_unet_train_workspace_bytes", "rdp_unet_train_forward", "rdp_unet_train_backward",
]


class ConvLayer(C.Structure):
    _fields_ = [("weight", C.c_void_p), ("shift", C.c_void_p), ("cin", C.c_int32), ("cout", C.c_int32)]


class UnetWeights(C.Structure):
    _fields_ = [("n_channels", C.c_int32), ("n_classes", C.c_int32), ("conv", ConvLayer * 18),
                ("outc_weight", C.c_void_p), ("outc_bias", C.c_void_p)]


class CurvatureOut(C.Structure):
    _fields_ = [("mean_curvature", C.c_void_p), ("max_curvature", C.c_void_p), ("spline_points", C.c_void_p),
                ("status", C.c_void_p), ("n_points", C.c_void_p), ("n_edge", C.c_void_p), ("fit_residual", C.c_void_p),
                ("points_xyz", C.c_void_p), ("edge_pixels", C.c_void_p), ("bin_counts", C.c_void_p),
                ("n_knots", C.c_void_p), ("fit_ier", C.c_void_p), ("fit_flags", C.c_void_p)]


class RdpError(RuntimeError):
    pass


_lock = threading.Lock()
_lib = None


def load() -> C.CDLL:
    """Loads the library once.  Raises (loudly) when it has not been built: there is no fallback path."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise RdpError(f"{LIB_PATH} is missing: build it with `python __graft_entry__.py` "
                           "(or rdp_b200/build.py); rdp_b200 has no CPU fallback")
        lib = C.CDLL(LIB_PATH)
        lib.rdp_abi_version.restype = C.c_int
        if lib.rdp_abi_version() != ABI_VERSION:
            raise RdpError(f"{LIB_PATH} has ABI version {lib.rdp_abi_version()}, this package needs {ABI_VERSION}: rebuild it "
                           "(python __graft_entry__.py)")
        lib.rdp_last_error.restype = C.c_char_p
        lib.rdp_unet_create.argtypes = [C.POINTER(UnetWeights), C.POINTER(C.c_void_p)]
        lib.rdp_unet_destroy.argtypes = [C.c_void_p]
        lib.rdp_unet_destroy.restype = None
        lib.rdp_unet_workspace_bytes.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_size_t)]
        lib.rdp_unet_workspace_bytes_ex.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_size_t)]
        lib.rdp_unet_workspace_bytes_ex.restype = C.c_int
        lib.rdp_unet_forward.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                         C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        lib.rdp_unet_launch_count.argtypes = [C.c_void_p]
        lib.rdp_unet_read_activation.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                                 C.c_void_p]
        lib.rdp_geom_workspace_bytes.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_size_t)]
        lib.rdp_geom_curvature.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                           C.POINTER(C.c_double), C.c_double, C.POINTER(CurvatureOut), C.c_void_p,
                                           C.c_size_t, C.c_void_p]
        lib.rdp_geom_launch_count.argtypes = []
        lib.rdp_geom_deproject_static.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                                  C.POINTER(C.c_double), C.c_double, C.c_void_p, C.c_void_p,
                                                  C.c_void_p, C.c_void_p]
        lib.rdp_geom_deproject_static.restype = C.c_int
        lib.rdp_pre_resize.argtypes = [C.c_void_p] + [C.c_int] * 7 + [C.c_void_p, C.c_void_p]
        lib.rdp_pre_resize.restype = C.c_int
        lib.rdp_post_mask_nearest.argtypes = [C.c_void_p] + [C.c_int] * 5 + [C.c_void_p, C.c_void_p, C.c_void_p]
        lib.rdp_post_mask_nearest.restype = C.c_int
        lib.rdp_post_mask_png_sizes.argtypes = [C.c_int] * 3 + [C.POINTER(C.c_size_t)] * 2
        lib.rdp_post_mask_png_sizes.restype = C.c_int
        lib.rdp_post_mask_png.argtypes = [C.c_void_p] + [C.c_int] * 3 + [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p,
                                                                        C.c_size_t, C.c_void_p]
        lib.rdp_post_mask_png.restype = C.c_int
        lib.rdp_pre_decode_sizes.argtypes = [C.c_int] * 3 + [C.POINTER(C.c_size_t)] * 2
        lib.rdp_pre_decode_sizes.restype = C.c_int
        for name in ("rdp_pre_decode_jpeg", "rdp_pre_decode_png16"):
            getattr(lib, name).argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                           C.c_void_p, C.c_size_t, C.c_void_p]
            getattr(lib, name).restype = C.c_int
        lib.rdp_unet_forward_profiled.argtypes = lib.rdp_unet_forward.argtypes + [C.POINTER(C.c_float)]
        lib.rdp_geom_curvature_profiled.argtypes = lib.rdp_geom_curvature.argtypes + [C.POINTER(C.c_float)]
        for name in ("rdp_unet_create", "rdp_unet_workspace_bytes", "rdp_unet_forward", "rdp_unet_launch_count",
                     "rdp_unet_read_activation", "rdp_geom_workspace_bytes", "rdp_geom_curvature",
                     "rdp_geom_launch_count", "rdp_unet_forward_profiled", "rdp_geom_curvature_profiled"):
            getattr(lib, name).restype = C.c_int
        _lib = lib
    return _lib


def check(rc: int, what: str) -> None:
    if rc != OK:
        msg = load().rdp_last_error().decode("utf-8", "replace")
        raise RdpError(f"{what} failed ({rc}): {msg}")
