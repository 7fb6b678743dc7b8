#!/usr/bin/env python
"""Runs the geometry chain a few times on a fixed batch (development aid for ncu captures)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "robotic-discovery-platform_b200"))
from rdp_b200 import geometry, synth
B, H, W, fill = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), float(sys.argv[4])
frames = [synth.make_frame(8000 + i, H, W, fill, "u16") for i in range(8)]
masks = torch.from_numpy(np.stack([frames[i % 8]["mask_gt"] for i in range(B)])).cuda()
depth = torch.from_numpy(np.stack([frames[i % 8]["depth"] for i in range(B)])).cuda()
ws = torch.empty(geometry.workspace_bytes(B, H, W), dtype=torch.uint8, device="cuda")
for i in range(4):
    lm = []
    r = geometry.curvature_batch(masks, depth, frames[0]["K"], 0.001, workspace=ws, launch_ms=lm)
torch.cuda.synchronize()
print(dict(zip(geometry.LAUNCH_NAMES, [round(v, 4) for v in lm])), int((r.status == 0).sum()))
