"""Runs the golden knot-insertion cases one frame at a time and prints knots / ier / fp / curvature next to the
reference values; with RDP_DEBUG_KNOTS=2 the kernel also prints its solve trajectory and phase cycle counts (dev aid)."""
import json, os, sys
import numpy as np, torch
sys.path[:0] = ['/root/repo', '/root/repo/robotic-discovery-platform_b200']
from rdp_b200 import geometry, synth
man = json.load(open('/root/repo/tests/golden/manifest.json'))
cases = [(k, v) for k, v in man["cases"].items() if v["kind"] == "geometry" and v["n_knots"] != 8]
frames = [synth.make_frame(v["seed"], v["H"], v["W"], v["fill"], v["depth"]) for _, v in cases]
for i, ((name, meta), f) in enumerate(zip(cases, frames)):
    m = torch.from_numpy(f["mask_gt"])[None].cuda(); d = torch.from_numpy(f["depth"])[None].cuda()
    r = geometry.curvature_batch(m, d, f["K"], f["depth_scale"])
    torch.cuda.synchronize()
    print(name, "knots", int(r.n_knots[0]), "ref", meta["n_knots"], "ier", int(r.fit_ier[0]), meta["ier"], "fp", float(r.fit_residual[0]),
          "n_edge", int(r.n_edge[0]), "mean", float(r.mean_curvature[0]), meta["mean_curvature"], "max", float(r.max_curvature[0]), meta["max_curvature"])
