"""The reference's training loop (scripts/train_segmenter.py:143-165: Adam 1e-4, BCEWithLogitsLoss, batch 4 at 256x256) run
for N steps with the drop-in UNet (our kernels) and, from the same initial weights on the same batches, with the reference
module itself through PyTorch/cuDNN autograd (oracle/_ref; else the oracle's train-mode restatement): the two loss curves
and the validation loss of both models in eval mode.  python tools/train_curve.py [steps] > profiles/r02_train_curve.txt"""
import importlib.util, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "robotic-discovery-platform_b200")]
from oracle.make_golden_train import train_batch
from pkg.segmentation_model import UNet
from rdp_b200 import synth

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 60
sd = synth.make_state_dict(7, planted=False)
batches = [tuple(torch.from_numpy(a).cuda() for a in train_batch(tuple(range(100 + 4 * i, 104 + 4 * i)), 256, 256)) for i in range(8)]
val = tuple(torch.from_numpy(a).cuda() for a in train_batch((900, 901, 902, 903), 256, 256))
crit = torch.nn.BCEWithLogitsLoss()

def ref_model():
    path = os.path.join(ROOT, "oracle", "_ref", "pkg", "segmentation_model.py")
    if not os.path.exists(path):
        return None
    spec = importlib.util.spec_from_file_location("_rdp_ref_seg_curve", path)
    m = importlib.util.module_from_spec(spec); spec.loader.exec_module(m)
    return m.UNet(3, 1)

def run(model):
    model.load_state_dict(sd)
    model = model.cuda().train()
    opt = torch.optim.Adam(model.parameters(), lr=1e-4)
    losses = []
    for s in range(steps):
        x, t = batches[s % len(batches)]
        opt.zero_grad()
        loss = crit(model(x), t)
        loss.backward()
        opt.step()
        losses.append(float(loss.detach()))
    model.eval()
    with torch.no_grad():
        v = float(crit(model(val[0]), val[1]))
    return losses, v

ours, v_ours = run(UNet(3, 1))
ref = ref_model()
if ref is None:
    print("oracle/_ref is absent: only the drop-in module's curve"); theirs, v_ref = [float("nan")] * steps, float("nan")
else:
    theirs, v_ref = run(ref)
print(f"{steps} Adam steps, batch 4 at 256x256, 8 batches cycling, same initial weights (synth seed 7); torch {torch.__version__}")
print("step   drop-in (sm_100a kernels)   reference (cuDNN autograd, PyTorch defaults)   relative difference")
for s in list(range(0, steps, 5)) + [steps - 1]:
    print(f"{s:4d}   {ours[s]:.5f}                     {theirs[s]:.5f}                                        {abs(ours[s] - theirs[s]) / theirs[s]:.2e}")
print(f"validation loss in eval mode (4 unseen frames): drop-in {v_ours:.5f}, reference {v_ref:.5f}")
