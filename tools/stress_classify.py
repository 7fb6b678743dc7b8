"""Classifies the frames tools/stress_knots.py reported as differing (reads gpurun_out/stress_knots.json, CPU only):
for each, scipy's FITPACK, the oracle's FITPACK port and the GPU are compared pairwise.  Where the port itself leaves
scipy, the frame sits on one of FITPACK's rounding-decided branches and no implementation can be expected to follow."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "robotic-discovery-platform_b200"), os.path.join(ROOT, "tools")]
from oracle import geometry_port as gp  # noqa: E402
from stress_one import frame  # noqa: E402


def rel(a, b):
    return abs(a - b) / abs(b) if b else abs(a)


def main():
    d = json.load(open(os.path.join(ROOT, "gpurun_out", "stress_knots.json")))
    rows = []
    for i in d["differing"]:
        f = frame(i, d["n"])
        st, nk, mean, mx, ier = d["gpu"][str(i)]
        s = gp.curvature_profile(f["mask_gt"], f["depth"], f["K"], f["depth_scale"], fit="scipy")
        p = gp.curvature_profile(f["mask_gt"], f["depth"], f["K"], f["depth_scale"], fit="port")
        row = dict(frame=i, knots=dict(gpu=nk, port=p.n_knots, scipy=s.n_knots),
                   gpu_vs_scipy=max(rel(mean, s.mean_curvature), rel(mx, s.max_curvature)),
                   gpu_vs_port=max(rel(mean, p.mean_curvature), rel(mx, p.max_curvature)),
                   port_vs_scipy=max(rel(p.mean_curvature, s.mean_curvature), rel(p.max_curvature, s.max_curvature)))
        rows.append(row)
        print(json.dumps(row), flush=True)
    follows_port = sum(r["gpu_vs_port"] <= 1e-6 for r in rows)
    port_leaves = sum(r["port_vs_scipy"] > 1e-6 for r in rows)
    print(f"{len(rows)} differing frames of {d['n']}: the GPU follows the port in {follows_port}, the port itself leaves scipy in {port_leaves}")


if __name__ == "__main__":
    main()
