"""One-off randomised check of the geometry kernels before the spline (deprojection, binning, edge selection, ordering)
against the oracle on arbitrary masks - blobs, rings, speckle, thin lines, with holes in the depth - where every integer
quantity has to be bit-exact; the 8-knot fits that result are compared too, knot-insertion frames only up to the edge."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "robotic-discovery-platform_b200")]
from oracle import geometry_port as gp  # noqa: E402
from rdp_b200 import geometry, synth  # noqa: E402


def random_mask(rng, H, W):
    yy, xx = np.mgrid[0:H, 0:W]
    kind = int(rng.integers(0, 5))
    if kind == 0:      # ellipses
        m = np.zeros((H, W), bool)
        for _ in range(int(rng.integers(1, 5))):
            cy, cx, ry, rx = rng.uniform(0, H), rng.uniform(0, W), rng.uniform(3, H / 2), rng.uniform(3, W / 2)
            m |= ((yy - cy) / ry) ** 2 + ((xx - cx) / rx) ** 2 < 1
    elif kind == 1:    # ring
        r = np.hypot(yy - H / 2, (xx - W / 2) * H / W)
        m = (r > H * 0.2) & (r < H * rng.uniform(0.25, 0.45))
    elif kind == 2:    # speckle
        m = rng.random((H, W)) < rng.uniform(0.01, 0.6)
    elif kind == 3:    # a few thin lines
        m = np.zeros((H, W), bool)
        for _ in range(int(rng.integers(1, 6))):
            a, b = rng.uniform(-1, 1), rng.uniform(0, H)
            m |= np.abs(yy - (a * xx + b)) < rng.uniform(0.6, 3)
    else:              # everything, or a rectangle
        m = np.ones((H, W), bool) if rng.random() < 0.3 else ((yy > H * 0.3) & (yy < H * 0.8) & (xx > W * 0.1) & (xx < W * 0.7))
    return m.astype(np.uint8)


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 60
    rng = np.random.default_rng(11)
    stats = {}
    for it in range(n):
        H, W = int(rng.integers(20, 200)), int(rng.integers(20, 260))
        flav = "f32" if it % 2 else "u16"
        f = synth.make_frame(3000 + it, H, W, 0.3, flav)
        f["mask_gt"] = random_mask(rng, H, W)
        if flav == "f32":     # continuous noise keeps x and y tie-free (ties are compared through the explicit rule anyway)
            f["depth"] = (f["depth"] * (1 + 0.2 * rng.random((H, W)))).astype(np.float32)
        r = geometry.curvature_batch(torch.from_numpy(f["mask_gt"])[None].cuda(), torch.from_numpy(f["depth"])[None].cuda(),
                                     f["K"], f["depth_scale"], want_points=True, want_edge=True, want_bins=True)
        torch.cuda.synchronize()
        pcd, pix = gp.deproject(f["mask_gt"], f["depth"], f["K"], f["depth_scale"])
        npts = int(r.n_points[0])
        assert npts == pcd.shape[0], (it, "n_points")
        np.testing.assert_allclose(r.points_xyz[0, :npts].cpu().numpy().astype(np.float64), pcd, rtol=1e-5, atol=1e-7)
        st = int(r.status[0])
        if npts < gp.MIN_POINTS:
            assert st == gp.TOO_FEW_POINTS, (it, st)
            stats["few points"] = stats.get("few points", 0) + 1
            continue
        rows, counts, ks = gp.select_edge(pcd, pix)
        if rows.shape[0] == 0:
            assert st == gp.DEGENERATE_BINS, (it, st)
            continue
        assert np.array_equal(r.bin_counts[0].cpu().numpy(), counts), (it, "bins")
        assert int(r.n_edge[0]) == rows.shape[0], (it, "n_edge")
        if rows.shape[0] < gp.MIN_EDGE:
            assert st == gp.TOO_FEW_EDGE, (it, st)
            stats["few edge points"] = stats.get("few edge points", 0) + 1
            continue
        order = gp.order_along_x(pcd, pix, rows)
        E = rows.shape[0]
        assert np.array_equal(r.edge_pixels[0, :E].cpu().numpy().astype(np.int64), pix[order]), (it, "edge order")
        o = gp.curvature_profile(f["mask_gt"], f["depth"], f["K"], f["depth_scale"], fit="scipy")
        assert st == o.status, (it, st, o.status)
        key = "ok, %s" % ("8 knots" if o.n_knots == 8 else "knot insertion") if o.status == gp.OK else "status %d" % o.status
        stats[key] = stats.get(key, 0) + 1
        if o.status == gp.OK and o.n_knots == 8:
            assert int(r.n_knots[0]) == 8
            assert abs(float(r.mean_curvature[0]) - o.mean_curvature) <= 1e-8 * abs(o.mean_curvature), (it, "curvature")
    print(f"{n} random masks: every integer quantity bit-exact; outcomes {stats}")


if __name__ == "__main__":
    main()
