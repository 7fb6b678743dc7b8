"""One headline step (640x480 U-Net + curvature) and one server-shaped step with PNG, 64 frames each, for ncu captures."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "robotic-discovery-platform_b200")]
from pkg.segmentation_model import UNet  # noqa: E402
from rdp_b200 import geometry, service, synth  # noqa: E402

B, H, W = 64, 480, 640
frames = [synth.make_frame(9500 + i, H, W, 0.30, "u16") for i in range(8)]
rgb = torch.from_numpy(np.stack([frames[i % 8]["rgb"] for i in range(B)])).cuda()
depth = torch.from_numpy(np.stack([frames[i % 8]["depth"] for i in range(B)])).cuda()
model = UNet(3, 1).eval()
model.load_state_dict(synth.make_state_dict(2, planted=True))
model = model.cuda()
for _ in range(2):
    mask, _, _ = model.forward_mask(rgb)
    geometry.curvature_batch(mask[:, 0], depth, frames[0]["K"], frames[0]["depth_scale"])
    r = service.analyze_frames(model, rgb, depth, frames[0]["K"], frames[0]["depth_scale"])
    service.encode_mask_png(r.final_mask, to_host=False)
torch.cuda.synchronize()
print("ok")
