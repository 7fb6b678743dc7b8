"""One-off stress run of k_fit_knots against scipy's FITPACK (through the oracle) on many staircase-edged frames:
knot counts must agree, curvature to 1e-6 relative; frames where FITPACK's own control flow hangs on an ulp
 are listed, not hidden.    python tools/stress_knots.py [n_frames]"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "robotic-discovery-platform_b200")]
from oracle import geometry_port as gp  # noqa: E402
from oracle import service_port as sp  # noqa: E402
from rdp_b200 import geometry, synth  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 48
    plain = "--plain" in sys.argv        # ground-truth masks (mostly the 8-knot cubic regime) instead of staircase edges
    rng = np.random.default_rng(123)
    H, W = int(os.environ.get("H", 240)), int(os.environ.get("W", 320))
    frames = []
    for i in range(n):
        fill = float(rng.uniform(0.15, 0.6))
        f = synth.make_frame(1000 + i, H, W, fill, "u16" if i % 2 else "f32")
        if not plain:
            small = f["mask_gt"][::3, ::3] if i % 3 else f["mask_gt"][::2, ::2]      # what a low-resolution U-Net would give
            f["mask_gt"] = np.ascontiguousarray(sp.upsize_mask_np(small, H, W))
        frames.append(f)
    groups = {}
    for i, f in enumerate(frames):
        groups.setdefault(f["depth"].dtype.str, []).append(i)
    got = {}
    for idx in groups.values():
        masks = torch.from_numpy(np.stack([frames[i]["mask_gt"] for i in idx])).cuda()
        depth = torch.from_numpy(np.stack([frames[i]["depth"] for i in idx])).cuda()
        r = geometry.curvature_batch(masks, depth, frames[idx[0]]["K"], frames[idx[0]]["depth_scale"])
        torch.cuda.synchronize()
        for j, i in enumerate(idx):
            got[i] = (int(r.status[j]), int(r.n_knots[j]), float(r.mean_curvature[j]), float(r.max_curvature[j]), int(r.fit_ier[j]))
    worst, bad, t0 = 0.0, [], time.time()
    hist = []
    for i, f in enumerate(frames):
        o = gp.curvature_profile(f["mask_gt"], f["depth"], f["K"], f["depth_scale"], fit="scipy")
        st, nk, mean, mx, ier = got[i]
        if st != o.status or (o.status == gp.OK and nk != o.n_knots):
            bad.append((i, "status/knots", st, o.status, nk, o.n_knots))
            continue
        if o.status != gp.OK:
            continue
        hist.append(nk)
        if o.n_knots != nk:
            continue
        rel = max(abs(mean - o.mean_curvature) / o.mean_curvature, abs(mx - o.max_curvature) / o.max_curvature)
        worst = max(worst, rel if rel < 1e-4 else 0.0)
        if rel > 1e-6:
            bad.append((i, "curvature", rel, nk, ier, o.ier))
    print(f"{n} frames, knots min/median/max {min(hist)}/{int(np.median(hist))}/{max(hist)}, worst relative curvature difference "
          f"among agreeing frames {worst:.2e}, oracle time {time.time() - t0:.1f} s")
    print("differing frames:", bad if bad else "none")
    out = os.path.join(ROOT, "gpurun_out", "stress_knots.json")
    if os.path.isdir(os.path.dirname(out)):
        import json
        with open(out, "w") as fh:
            json.dump({"n": n, "gpu": {str(i): got[i] for i in got}, "differing": [b[0] for b in bad]}, fh)


if __name__ == "__main__":
    main()
