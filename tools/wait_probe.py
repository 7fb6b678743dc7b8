#!/usr/bin/env python
"""One forward at batch 64, 640x480 with RDP_DEBUG_WAITS=1: CTA 0 of every conv prints where it waited."""
import os, sys
os.environ["RDP_DEBUG_WAITS"] = "1"
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "robotic-discovery-platform_b200"))
from pkg.segmentation_model import UNet
from rdp_b200 import synth
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
m = UNet(3, 1); m.load_state_dict(synth.make_state_dict(2, planted=True)); m = m.cuda().eval()
base = [synth.make_frame(5000 + i, 480, 640, 0.15, "u16") for i in range(8)]
rgb = torch.from_numpy(np.stack([base[i % 8]["rgb"] for i in range(B)])).cuda()
m.forward_mask(rgb); torch.cuda.synchronize()
