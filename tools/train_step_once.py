"""Two training steps (forward + backward through the drop-in UNet in train() mode) at batch B (default 32), 256x256:
the short command for the ncu launch list of the training step (profiles/r02_train_launches.csv)."""
import os, sys
import torch
sys.path[:0] = ['/root/repo', '/root/repo/robotic-discovery-platform_b200']
from oracle.make_golden_train import train_batch
from pkg.segmentation_model import UNet
from rdp_b200 import synth
B = int(os.environ.get("B", 32))
x, t = (torch.from_numpy(a).cuda() for a in train_batch(tuple(range(40, 40 + B)), 256, 256))
model = UNet(3, 1); model.load_state_dict(synth.make_state_dict(7, planted=False)); model = model.cuda().train()
crit = torch.nn.BCEWithLogitsLoss()
for _ in range(2):
    model.zero_grad()
    loss = crit(model(x), t)
    loss.backward()
torch.cuda.synchronize()
print("ok", float(loss.detach()))
