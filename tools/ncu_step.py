"""One U-Net forward + curvature at batch B (default 16) after one warm-up pass: the short command that the ncu
captures under profiles/ are taken on (tools/README in DESIGN 5).  Launch order per pass: 27 U-Net slots, 7 geometry."""
import os, sys
import numpy as np, torch
sys.path[:0] = ['/root/repo', '/root/repo/robotic-discovery-platform_b200']
from pkg.segmentation_model import UNet
from rdp_b200 import geometry, synth
B, H, W = int(os.environ.get('B', 16)), int(os.environ.get('H', 480)), int(os.environ.get('W', 640))
model = UNet(3, 1); model.load_state_dict(synth.make_state_dict(2, planted=True)); model = model.cuda().eval()
base = [synth.make_frame(5000 + i, H, W, 0.15, "u16") for i in range(8)]
rgb = torch.from_numpy(np.stack([base[i % 8]["rgb"] for i in range(B)])).cuda()
depth = torch.from_numpy(np.stack([base[i % 8]["depth"] for i in range(B)])).cuda()
eng = model._engine_for(rgb.device); ws = eng.new_workspace(B, H, W)
gws = torch.empty(geometry.workspace_bytes(B, H, W), dtype=torch.uint8, device="cuda")
for _ in range(2):
    mask, _, _ = model.forward_mask(rgb, workspace=ws)
    r = geometry.curvature_batch(mask[:, 0], depth, base[0]["K"], 0.001, workspace=gws)
torch.cuda.synchronize()
print("ok", int((r.status == 0).sum()))
