"""Runs one training step on the small golden case with a checkpoint after every launch group (dev build + RDP_TRAIN_SYNC=1)
and prints forward / gradient agreement per tensor (dev aid)."""
import json, os, sys
import numpy as np, torch
sys.path[:0] = ['/root/repo', '/root/repo/robotic-discovery-platform_b200']
from oracle.make_golden_train import train_batch
from pkg.segmentation_model import UNet
from rdp_b200 import synth
name = sys.argv[1] if len(sys.argv) > 1 else "train_step_2x64x96"
meta = json.load(open('/root/repo/tests/golden/train_manifest.json'))["cases"][name]
g = dict(np.load(f'/root/repo/tests/golden/{name}.npz'))
model = UNet(3, 1); model.load_state_dict(synth.make_state_dict(meta["weight_seed"], planted=False)); model = model.cuda().train()
x, t = (torch.from_numpy(a).cuda() for a in train_batch(meta["frame_seeds"], meta["H"], meta["W"]))
out = model(x); torch.cuda.synchronize()
loss = torch.nn.BCEWithLogitsLoss()(out, t)
print("loss", loss.item(), "ref", float(g["loss"]))
idx = torch.from_numpy(g["logits_idx"]).cuda()
print("logit err", float((out.detach().reshape(-1)[idx].cpu() - torch.from_numpy(g["logits_val"])).abs().max()), "of", float(g["logits_absmax"]))
try:
    loss.backward(); torch.cuda.synchronize()
finally:
    import ctypes
    from rdp_b200 import _lib
    lib = _lib.load()
    if hasattr(lib, "rdp_debug_timeouts"):
        buf = (ctypes.c_int * 8)()
        print("debug timeouts rc", lib.rdp_debug_timeouts(buf), list(buf))
def rl2(a, b):
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))
print(f"{'parameter':45s} {'gpu/fp32':>9s} {'emul/fp32':>9s} {'gpu/emul':>9s} {'maxerr/max vs fp32':>9s} norm ratio")
for pname, p in model.named_parameters():
    idx = torch.from_numpy(g[f"grad_idx/{pname}"]).cuda()
    got = p.grad.reshape(-1)[idx].cpu().numpy()
    want, emu, scale = g[f"grad_val/{pname}"], g[f"grad_val_bf16/{pname}"], float(g[f"grad_absmax/{pname}"])
    print(f"{pname:45s} {rl2(got, want):9.4f} {rl2(emu, want):9.4f} {rl2(got, emu):9.4f} {np.abs(got - want).max() / scale:9.4f} "
          f"{float(p.grad.double().norm()) / float(g[f'grad_norm/{pname}']):.4f}")
for bname, b in model.named_buffers():
    if "running" in bname:
        print(f"{bname:45s} max abs err vs fp32 {np.abs(b.cpu().numpy() - g[f'buffer/{bname}']).max():.2e}  vs bf16-rounded reference "
              f"{np.abs(b.cpu().numpy() - g[f'buffer_bf16/{bname}']).max():.2e} (reference's own shift {np.abs(g[f'buffer_bf16/{bname}'] - g[f'buffer/{bname}']).max():.2e})")
