"""Per-launch U-Net times (CUDA events between launches) at batch 64, 640x480; run under the RDP_DEBUG_* switches
for timing experiments (dev aid)."""
import json, os, sys
import numpy as np, torch
sys.path[:0] = ['/root/repo', '/root/repo/robotic-discovery-platform_b200']
from pkg.segmentation_model import UNet
from rdp_b200 import synth

B,H,W=int(os.environ.get('B',64)),int(os.environ.get('H',480)),int(os.environ.get('W',640))
model = UNet(3,1); model.load_state_dict(synth.make_state_dict(2, planted=True)); model=model.cuda().eval()
base=[synth.make_frame(5000+i,H,W,0.15,"u16") for i in range(8)]
rgb=torch.from_numpy(np.stack([base[i%8]["rgb"] for i in range(B)])).cuda()
eng=model._engine_for(rgb.device); ws=eng.new_workspace(B,H,W)
acc=None
for it in range(6):
    ms=[]
    eng.forward(rgb, want_logits=False, want_mask=True, workspace=ws, launch_ms=ms)
    if it>=2:
        a=np.array(ms); acc = a if acc is None else acc+a
acc/=4
d=dict(zip(eng.LAUNCH_NAMES,[round(float(v),3) for v in acc]))
print(json.dumps({k:d[k] for k in d if k.startswith("conv") or "up4" in k}), "total", round(float(acc.sum()),2))
