"""`pkg.geometry_utils` of the reference, served by the sm_100a geometry kernels.

`compute_curvature_profile` keeps the reference signature and keyword names (`pkg/geometry_utils.py:42-43`; the
server calls it with keywords in a different order, `server.py:127-130`) and its convention of returning an empty
`CurvatureResult()` for frames that cannot be analysed (`:64-65,69-70,95-97`).  The arrays are uploaded to the
current CUDA device, processed by `rdp_geom_curvature`, and the 2 + 300 doubles of the result come back.
There is no CPU path.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List

import numpy as np

__all__ = ["Point", "CurvatureResult", "compute_curvature_profile", "compute_curvature_profiles"]


@dataclass
class Point:
    """A 3-D point (field names are part of the wire mapping, `server.py:140`)."""
    x: float = 0.0
    y: float = 0.0
    z: float = 0.0


@dataclass
class CurvatureResult:
    mean_curvature: float = 0.0
    max_curvature: float = 0.0
    spline_points: List[Point] = field(default_factory=list)


def _to_result(status: int, mean: float, mx: float, pts: np.ndarray) -> CurvatureResult:
    from rdp_b200 import _lib
    if status == _lib.FRAME_OK:
        return CurvatureResult(mean_curvature=float(mean), max_curvature=float(mx),
                               spline_points=[Point(x, y, z) for x, y, z in np.asarray(pts, dtype=np.float64).tolist()])
    return CurvatureResult()


def compute_curvature_profiles(masks, depth_images, intrinsics, depth_scale) -> List[CurvatureResult]:
    """Batched form: (B,H,W) arrays or CUDA tensors -> list of `CurvatureResult`."""
    import torch

    from rdp_b200 import geometry

    def to_dev(a, is_mask):
        if isinstance(a, torch.Tensor):
            t = a
        else:
            a = np.ascontiguousarray(a)
            if is_mask and a.dtype != np.uint8:
                a = (a > 0).astype(np.uint8)
            t = torch.from_numpy(a)
        if is_mask and t.dtype != torch.uint8:
            t = (t > 0).to(torch.uint8)
        return t.cuda(non_blocking=True) if not t.is_cuda else t

    if not torch.cuda.is_available():
        raise RuntimeError("compute_curvature_profile (rdp_b200) has no CPU path and no CUDA device is visible")
    # `depth_image[v, u] * depth_scale` (`geometry_utils.py:105`) is computed in NumPy's promoted type: uint16 and float32
    # images have kernels of their own; every other dtype (float64, signed or wider integers) promotes to float64, so it
    # is widened to float64 on the host - exact for the integers, negative values stay negative and are dropped by z > 0.
    if not isinstance(depth_images, torch.Tensor):
        depth_images = np.asarray(depth_images)
        if depth_images.dtype not in (np.uint16, np.float32):
            depth_images = depth_images.astype(np.float64)
    m, d = to_dev(masks, True), to_dev(depth_images, False)
    if d.dtype not in (torch.uint16, torch.float32, torch.float64):
        d = d.to(torch.float64)
    fmt = geometry.depth_format(d.dtype, depth_scale)
    r = geometry.curvature_batch(m, d, np.asarray(intrinsics, dtype=np.float64), depth_scale, depth_fmt=fmt)
    status = r.status.cpu().numpy()
    mean, mx, pts = r.mean_curvature.cpu().numpy(), r.max_curvature.cpu().numpy(), r.spline_points.cpu().numpy()
    return [_to_result(int(status[i]), mean[i], mx[i], pts[i]) for i in range(status.shape[0])]


def compute_curvature_profile(mask: np.ndarray, depth_image: np.ndarray, intrinsics: np.ndarray,
                              depth_scale: float) -> CurvatureResult:
    """mask (H,W) any dtype tested `> 0`, depth (H,W) of any real dtype, K (3,3), scale -> `CurvatureResult`."""
    mask = np.asarray(mask)
    depth_image = np.asarray(depth_image)
    if mask.ndim != 2 or depth_image.shape != mask.shape:
        return CurvatureResult()
    return compute_curvature_profiles(mask[None], depth_image[None], intrinsics, depth_scale)[0]
