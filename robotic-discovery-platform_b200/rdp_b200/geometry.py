"""Batched curvature profile on the GPU through the C ABI (`rdp_geom_curvature`).

Replaces `compute_curvature_profile` and its helpers (reference `pkg/geometry_utils.py:42-162`) for device tensors:
masks (B,H,W) uint8, depth (B,H,W) uint16 | float32 | float64.  No CPU fallback.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np
import torch

from . import _lib

NUM_SAMPLES = 100
NUM_BINS = 50
LAUNCH_NAMES = ["memset", "deproject", "bin_hist", "scatter", "select_sort", "fit", "fit_knots"]


@dataclass
class CurvatureBatch:
    mean_curvature: torch.Tensor   # (B,) f64
    max_curvature: torch.Tensor    # (B,) f64
    spline_points: torch.Tensor    # (B,100,3) f64
    status: torch.Tensor           # (B,) i32, rdp_b200._lib.FRAME_*
    n_points: torch.Tensor         # (B,) i32
    n_edge: torch.Tensor           # (B,) i32
    fit_residual: torch.Tensor     # (B,) f64
    n_knots: torch.Tensor = None   # (B,) i32, 8 = single cubic (FITPACK ier=-2), more = knot insertion + smoothing
    fit_ier: torch.Tensor = None   # (B,) i32, FITPACK's ier
    fit_flags: torch.Tensor = None  # (B,) i32, rdp_b200._lib.FIT_FLAG_*: the fit took a decision inside rounding noise
    points_xyz: torch.Tensor | None = None    # (B,H*W,3) f32, first n_points rows valid
    edge_pixels: torch.Tensor | None = None   # (B,ecap) u32 as int32 view, first n_edge valid
    bin_counts: torch.Tensor | None = None    # (B,50) i32
    workspace: torch.Tensor | None = None


def depth_format(depth_dtype, depth_scale) -> int:
    """NumPy promotion of `depth_image[v, u] * depth_scale` (reference `geometry_utils.py:105`)."""
    if depth_dtype in (torch.uint16, np.dtype(np.uint16)):
        return _lib.DEPTH_U16
    if depth_dtype in (torch.float32, np.dtype(np.float32)):
        # Python float (weak scalar) keeps float32; a NumPy float64 scalar promotes the product to float64
        # (so does a 0-d float64 ndarray, which is what `np.load(...)['depth_scale']` gives, server.py:94-96)
        wide = isinstance(depth_scale, (np.ndarray, np.generic)) and np.asarray(depth_scale).dtype == np.float64
        return _lib.DEPTH_F32_WIDE if wide else _lib.DEPTH_F32
    if depth_dtype in (torch.float64, np.dtype(np.float64)):
        return _lib.DEPTH_F64
    raise _lib.RdpError(f"unsupported depth dtype {depth_dtype}")


def workspace_bytes(B: int, H: int, W: int) -> int:
    n = C.c_size_t()
    _lib.check(_lib.load().rdp_geom_workspace_bytes(B, H, W, C.byref(n)), "rdp_geom_workspace_bytes")
    return n.value


def curvature_batch(masks: torch.Tensor, depth: torch.Tensor, intrinsics, depth_scale, *, want_points=False,
                    want_edge=False, want_bins=False, workspace=None, depth_fmt=None,
                    launch_ms: list | None = None) -> CurvatureBatch:
    lib = _lib.load()
    if masks.device.type != "cuda" or depth.device != masks.device:
        raise _lib.RdpError("curvature_batch needs CUDA tensors on one device; there is no CPU path")
    if masks.dtype != torch.uint8 or masks.dim() != 3 or depth.shape != masks.shape:
        raise _lib.RdpError(f"masks must be (B,H,W) uint8 and depth the same shape (got {tuple(masks.shape)} {masks.dtype}, "
                            f"{tuple(depth.shape)} {depth.dtype})")
    B, H, W = masks.shape
    dev = masks.device
    fmt = depth_format(depth.dtype, depth_scale) if depth_fmt is None else depth_fmt
    masks, depth = masks.contiguous(), depth.contiguous()
    K = np.asarray(intrinsics, dtype=np.float64)
    k4 = (C.c_double * 4)(K[0, 0], K[1, 1], K[0, 2], K[1, 2]) if K.shape == (3, 3) else (C.c_double * 4)(*K.reshape(4))
    with torch.cuda.device(dev):
        need = workspace_bytes(B, H, W)
        if workspace is None or workspace.numel() < need:
            workspace = torch.empty(need, dtype=torch.uint8, device=dev)
        res = CurvatureBatch(
            mean_curvature=torch.empty(B, dtype=torch.float64, device=dev),
            max_curvature=torch.empty(B, dtype=torch.float64, device=dev),
            spline_points=torch.empty((B, NUM_SAMPLES, 3), dtype=torch.float64, device=dev),
            status=torch.empty(B, dtype=torch.int32, device=dev),
            n_points=torch.empty(B, dtype=torch.int32, device=dev),
            n_edge=torch.empty(B, dtype=torch.int32, device=dev),
            fit_residual=torch.empty(B, dtype=torch.float64, device=dev),
            n_knots=torch.zeros(B, dtype=torch.int32, device=dev),
            fit_ier=torch.zeros(B, dtype=torch.int32, device=dev),
            fit_flags=torch.zeros(B, dtype=torch.int32, device=dev),
            workspace=workspace,
        )
        out = _lib.CurvatureOut()
        out.mean_curvature, out.max_curvature = res.mean_curvature.data_ptr(), res.max_curvature.data_ptr()
        out.spline_points, out.status = res.spline_points.data_ptr(), res.status.data_ptr()
        out.n_points, out.n_edge = res.n_points.data_ptr(), res.n_edge.data_ptr()
        out.fit_residual = res.fit_residual.data_ptr()
        out.n_knots, out.fit_ier = res.n_knots.data_ptr(), res.fit_ier.data_ptr()
        out.fit_flags = res.fit_flags.data_ptr()
        if want_points:
            res.points_xyz = torch.empty((B, H * W, 3), dtype=torch.float32, device=dev)
            out.points_xyz = res.points_xyz.data_ptr()
        if want_edge:
            res.edge_pixels = torch.empty((B, H * W // 20 + 64), dtype=torch.int32, device=dev)
            out.edge_pixels = res.edge_pixels.data_ptr()
        if want_bins:
            res.bin_counts = torch.empty((B, NUM_BINS), dtype=torch.int32, device=dev)
            out.bin_counts = res.bin_counts.data_ptr()
        stream = torch.cuda.current_stream(dev).cuda_stream
        args = (masks.data_ptr(), depth.data_ptr(), fmt, B, H, W, k4, float(depth_scale), C.byref(out),
                workspace.data_ptr(), workspace.numel(), stream)
        if launch_ms is None:
            _lib.check(lib.rdp_geom_curvature(*args), "rdp_geom_curvature")
        else:       # measurement aid: CUDA events between launches, blocks until done
            buf = (C.c_float * lib.rdp_geom_launch_count())()
            _lib.check(lib.rdp_geom_curvature_profiled(*args, buf), "rdp_geom_curvature_profiled")
            launch_ms[:] = list(buf)
    return res
