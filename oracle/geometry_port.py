"""ORACLE (test infrastructure, not product code): NumPy restatement of the reference geometry stage.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU-baseline / `--impl reference` legs may import
this module.  The product path never does and has no CPU fallback.

Follows `pkg/geometry_utils.py` of the reference stage by stage:

* `deproject`            <- `_get_pcd_from_mask`          (`geometry_utils.py:101-117`)
* `bin_points`           <- bin arithmetic                (`:124-132`)
* `select_edge`          <- `_find_point_cloud_edge`      (`:119-142`)
* `order_along_x`        <- `np.argsort(edge[:, 0])`      (`:74`)
* `fit_spline`           <- `splprep(..., s=0.1, k=3)`    (`:78`)  -- FITPACK `parcur`, see `oracle/fitpack_port.py`
* `curvature_of_spline`  <- `_calculate_spline_curvature` (`:144-162`) and `splev` at 100 u (`:83-84`)
* `curvature_profile`    <- `compute_curvature_profile`   (`:42-97`) incl. the 100 / 20 point guards

The one place this port is *stricter* than the reference: `argpartition` (`:139`) and the default quicksort
`argsort` (`:74`) leave the choice among equal keys unspecified.  Here ties are broken explicitly:
larger y first, then smaller row-major pixel index; the x ordering breaks ties by pixel index.  On tie-free
inputs (the float-depth fixtures; asserted in the tests) this is exactly what the unmodified reference returns.

Parity pin: `oracle/make_golden.py` runs the live reference on seeded frames and stores its outputs in
`tests/golden/`; `tests/test_oracle_geometry.py` checks this port against them.  The reference ships no tests.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

from . import fitpack_port

NUM_BINS = 50            # geometry_utils.py:119
TOP_K_PERCENT = 0.05     # geometry_utils.py:119
MIN_POINTS = 100         # geometry_utils.py:64
MIN_EDGE = 20            # geometry_utils.py:69
SMOOTH_S = 0.1           # geometry_utils.py:78
NUM_SAMPLES = 100        # geometry_utils.py:83,144

# per-frame status codes shared with the CUDA path (include/rdp_b200.h: RDP_FRAME_*)
OK, TOO_FEW_POINTS, DEGENERATE_BINS, TOO_FEW_EDGE, FIT_INVALID = 0, 1, 2, 3, 4


def deproject(mask, depth, K, depth_scale):
    """Masked pixels -> (N,3) float64 points in row-major pixel order, plus their flat pixel indices.

    NumPy promotion is part of the contract: integer depth times a Python float is float64;
    float32 depth times a Python float stays float32 and is widened afterwards.
    """
    mask = np.asarray(mask)
    depth = np.asarray(depth)
    v, u = np.nonzero(mask > 0)
    z = depth[v, u] * depth_scale
    keep = z > 0
    v, u, z = v[keep], u[keep], z[keep]
    fx, fy, cx, cy = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
    x = (u - cx) * z / fx
    y = (v - cy) * z / fy
    pcd = np.empty((z.shape[0], 3), dtype=np.float64)
    pcd[:, 0], pcd[:, 1], pcd[:, 2] = x, y, z
    return pcd, (v * mask.shape[1] + u).astype(np.int64)


def bin_points(x, num_bins=NUM_BINS):
    """Returns (bin index per point, x_min, bin_width) or None when the reference would give up."""
    if x.shape[0] < num_bins:
        return None
    x_min, x_max = x.min(), x.max()
    width = (x_max - x_min) / num_bins
    if width <= 0:
        return None
    b = np.floor((x - x_min) / width).astype(np.int64)
    return np.clip(b, 0, num_bins - 1), x_min, width


def select_edge(pcd, pix, num_bins=NUM_BINS, top_k_percent=TOP_K_PERCENT):
    """Per x-bin, the k = max(1, int(cnt*0.05)) points with the largest y.  Returns (edge rows, counts, ks)."""
    binned = bin_points(pcd[:, 0], num_bins)
    if binned is None:
        return np.zeros(0, dtype=np.int64), np.zeros(num_bins, np.int64), np.zeros(num_bins, np.int64)
    b = binned[0]
    counts = np.bincount(b, minlength=num_bins)
    ks = np.where(counts > 0, np.maximum(1, (counts * top_k_percent).astype(np.int64)), 0)
    chosen = []
    for i in range(num_bins):
        if counts[i] == 0:
            continue
        rows = np.nonzero(b == i)[0]
        order = np.lexsort((pix[rows], -pcd[rows, 1]))   # primary: y descending, secondary: pixel index ascending
        chosen.append(rows[order[: ks[i]]])
    rows = np.concatenate(chosen) if chosen else np.zeros(0, dtype=np.int64)
    return rows, counts, ks


def order_along_x(pcd, pix, rows):
    return rows[np.lexsort((pix[rows], pcd[rows, 0]))]


def fit_spline(points, s=SMOOTH_S, fit="port"):
    """(E,3) ordered points -> (t, c[3][n-4], fp, ier).  Raises ValueError like `splprep` on bad input.

    fit="port" runs the FITPACK restatement in this directory; fit="scipy" makes the very call the reference makes
    (`geometry_utils.py:78`) - used where the reference's *speed* is wanted (bench.py's CPU baseline).
    """
    if fit == "scipy":
        import warnings
        from scipy.interpolate import splprep
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            (tck, _u), fp, ier, _msg = splprep([points[:, 0], points[:, 1], points[:, 2]], s=s, k=3, full_output=1)
        if ier > 3:
            raise ValueError("Error on input data")
        return np.asarray(tck[0]), np.asarray(tck[1]), fp, ier
    return fitpack_port.parcur(points, s=s)


def _splev(u, t, c, der, fit):
    if fit == "scipy":
        from scipy.interpolate import splev
        return np.asarray(splev(u, (t, list(c), 3), der=der))
    return fitpack_port.splev(u, t, c, der)


def curvature_of_spline(t, c, num=NUM_SAMPLES, fit="port"):
    u = np.linspace(0.0, 1.0, num)
    d1 = _splev(u, t, c, 1, fit)      # (3, num)
    d2 = _splev(u, t, c, 2, fit)
    cr = np.cross(d1.T, d2.T)
    n_cr = np.sqrt((cr * cr).sum(axis=1))
    n_d1 = np.sqrt((d1 * d1).sum(axis=0))
    ok = n_d1 > 1e-6
    if not ok.any():
        return 0.0, 0.0
    kappa = n_cr[ok] / n_d1[ok] ** 3
    return float(kappa.mean()), float(kappa.max())


@dataclass
class Profile:
    status: int = OK
    mean_curvature: float = 0.0
    max_curvature: float = 0.0
    spline_points: np.ndarray = field(default_factory=lambda: np.zeros((0, 3)))
    n_points: int = 0
    n_edge: int = 0
    n_knots: int = 0
    fp: float = 0.0
    ier: int = 0
    edge_pix: np.ndarray = field(default_factory=lambda: np.zeros(0, np.int64))   # x-ordered edge pixels
    bin_counts: np.ndarray = field(default_factory=lambda: np.zeros(NUM_BINS, np.int64))
    bin_k: np.ndarray = field(default_factory=lambda: np.zeros(NUM_BINS, np.int64))
    tck: tuple | None = None


def curvature_profile(mask, depth, K, depth_scale, fit="port") -> Profile:
    pcd, pix = deproject(mask, depth, K, depth_scale)
    out = Profile(n_points=int(pcd.shape[0]))
    if pcd.shape[0] < MIN_POINTS:
        out.status = TOO_FEW_POINTS
        return out
    rows, counts, ks = select_edge(pcd, pix)
    out.bin_counts, out.bin_k, out.n_edge = counts, ks, int(rows.shape[0])
    if rows.shape[0] == 0:
        out.status = DEGENERATE_BINS
        return out
    if rows.shape[0] < MIN_EDGE:
        out.status = TOO_FEW_EDGE
        return out
    rows = order_along_x(pcd, pix, rows)
    out.edge_pix = pix[rows]
    try:
        t, c, fp, ier = fit_spline(pcd[rows], fit=fit)
    except (TypeError, ValueError):
        out.status = FIT_INVALID
        return out
    out.tck, out.n_knots, out.fp, out.ier = (t, c, 3), int(t.shape[0]), float(fp), int(ier)
    out.mean_curvature, out.max_curvature = curvature_of_spline(t, c, fit=fit)
    out.spline_points = _splev(np.linspace(0.0, 1.0, NUM_SAMPLES), t, c, 0, fit).T.copy()
    return out
