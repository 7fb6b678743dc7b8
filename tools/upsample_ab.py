"""Development build only: the stand-alone up4 interpolation, 2x2-block kernel against the row kernel (RDP_UPSAMPLE_ROW):
logits of one forward must agree bit for bit, and the launch time of 'upsample up4' is printed for both.
Run each mode in its own process:  python tools/upsample_ab.py quad|row|generic  (writes gpurun_out/ups_<mode>.npy)"""
import os, sys
mode = sys.argv[1]
if mode == "row":
    os.environ["RDP_UPSAMPLE_ROW"] = "1"
if mode == "generic":       # fused interpolation producers of up1..up3: per-pixel path instead of the 2x2-block one
    os.environ["RDP_INTERP_GENERIC"] = "1"
import numpy as np, torch
sys.path[:0] = ['/root/repo', '/root/repo/robotic-discovery-platform_b200']
from pkg.segmentation_model import UNet
from rdp_b200 import synth
B, H, W = 16, 480, 640
model = UNet(3, 1); model.load_state_dict(synth.make_state_dict(2, planted=False)); model = model.cuda().eval()
base = [synth.make_frame(5000 + i, H, W, 0.15, "u16") for i in range(8)]
rgb = torch.from_numpy(np.stack([base[i % 8]["rgb"] for i in range(B)])).cuda()
eng = model._engine_for(rgb.device); ws = eng.new_workspace(B, H, W)
for it in range(4):
    ms = []
    out = eng.forward(rgb, want_logits=True, want_mask=True, workspace=ws, launch_ms=ms)
torch.cuda.synchronize()
d = dict(zip(eng.LAUNCH_NAMES, ms))
print(mode, "upsample up4 ms", round(d["upsample up4"], 4), "conv up4.0", round(d["conv up4.0"], 4))
logits = out[0] if isinstance(out, (tuple, list)) else out.logits
np.save(f"/root/repo/gpurun_out/ups_{mode}.npy", logits.float().cpu().numpy())
