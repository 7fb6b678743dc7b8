"""Re-runs one frame of tools/stress_knots.py: GPU trajectory (RDP_DEBUG_KNOTS=2) or oracle-port trajectory (--oracle)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "robotic-discovery-platform_b200")]
from oracle import fitpack_port as fp  # noqa: E402
from oracle import geometry_port as gp  # noqa: E402
from oracle import service_port as sp  # noqa: E402
from rdp_b200 import synth  # noqa: E402


def frame(i, n=int(os.environ.get("N", 512))):
    rng = np.random.default_rng(123)
    fills = [float(rng.uniform(0.15, 0.6)) for _ in range(n)]
    H, W = int(os.environ.get("H", 240)), int(os.environ.get("W", 320))
    f = synth.make_frame(1000 + i, H, W, fills[i], "u16" if i % 2 else "f32")
    small = f["mask_gt"][::3, ::3] if i % 3 else f["mask_gt"][::2, ::2]
    f["mask_gt"] = np.ascontiguousarray(sp.upsize_mask_np(small, H, W))
    return f


if __name__ == "__main__":
    i = int(sys.argv[1])
    f = frame(i)
    if "--oracle" in sys.argv:
        pcd, pix = gp.deproject(f["mask_gt"], f["depth"], f["K"], f["depth_scale"])
        rows, _, _ = gp.select_edge(pcd, pix)
        pts = pcd[gp.order_along_x(pcd, pix, rows)]
        orig_l, orig_s = fp._lsq, fp._smooth

        def tl(ta, n, u, X):
            r = orig_l(ta, n, u, X)
            print("  oracle lsq n %d fp %.17g sumdiag %.17g knot checksum %.17g min diag %.3g" %
                  (n, r[3], r[0][:, 0].sum(), float(np.sum(np.asarray(ta)[:n] * np.arange(1, n + 1))), r[0][:n - 4, 0].min()))
            return r

        def ts(ta, n, u, X, a, z, q, b, p):
            c, v = orig_s(ta, n, u, X, a, z, q, b, p)
            print("  oracle smoothing p %.17g fp %.17g" % (p, v))
            return c, v

        fp._lsq, fp._smooth = tl, ts
        t, c, fpv, ier = fp.parcur(pts)
        print("oracle port: knots", len(t), "fp", fpv, "ier", ier, "m", len(pts))
        o = gp.curvature_profile(f["mask_gt"], f["depth"], f["K"], f["depth_scale"], fit="scipy")
        print("scipy: knots", o.n_knots, "fp", o.fp, "mean", o.mean_curvature, "max", o.max_curvature)
    else:
        import torch
        from rdp_b200 import geometry
        r = geometry.curvature_batch(torch.from_numpy(f["mask_gt"])[None].cuda(), torch.from_numpy(f["depth"])[None].cuda(),
                                     f["K"], f["depth_scale"])
        torch.cuda.synchronize()
        print("gpu: knots", int(r.n_knots[0]), "fp", float(r.fit_residual[0]), "mean", float(r.mean_curvature[0]), "max",
              float(r.max_curvature[0]), "m", int(r.n_edge[0]))
