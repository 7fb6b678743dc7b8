"""Generates tests/golden/train_step_*.npz: one optimisation step of the reference's training loop
(`scripts/train_segmenter.py:143-165`: `UNet(3, 1)` in train mode, `BCEWithLogitsLoss`, `loss.backward()`, Adam 1e-4,
batch 4 at 256x256) run with the UNMODIFIED reference module under `torch.autograd` in fp32 on the CPU.

    python oracle/make_golden_train.py          (authoring container only: needs /root/reference)

Stored per case: the loss, sampled logits, for every parameter the gradient's max-abs, L2 norm and 256 sampled entries,
every BatchNorm layer's running statistics after the step, and sampled parameter values after the Adam step.  Inputs and
initial weights are regenerated from seeds by `rdp_b200.synth`.

Every quantity is stored twice: from the reference as it is (fp32 throughout, keys without suffix) and from the same
reference module with bfloat16 rounding inserted by hooks at exactly the places where the sm_100a kernels store bf16
(`*_bf16` keys): convolution weights, convolution inputs, convolution outputs before BatchNorm, activations after ReLU,
and the gradients at those same places.  A randomly initialised U-Net with batch-statistics BatchNorm amplifies
rounding: bf16 WEIGHTS alone move individual gradient entries by 10 % of their tensor's largest entry at 4 x 256 x 256
(30 % at 2 x 64 x 96, where the deepest BatchNorm sees 48 samples), although norms and the loss move by 1e-3.  So the
kernels are held tightly to the bf16-rounded reference (same algorithm, same storage precision) and loosely to the
fp32 one.  Test infrastructure only.
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "robotic-discovery-platform_b200"))
OUT = os.path.join(ROOT, "tests", "golden")
CASES = [("train_step_4x256x256", 7, (40, 41, 42, 43), 256, 256), ("train_step_2x64x96", 8, (44, 45), 64, 96)]


def train_batch(seeds, H, W):
    """(images (B,3,H,W) float32 in [0,1], masks (B,1,H,W) float32 {0,1}) as `SegmentationDataset` yields them."""
    from rdp_b200 import synth
    frames = [synth.make_frame(s, H, W, 0.2 + 0.05 * i, "u16") for i, s in enumerate(seeds)]
    x = np.stack([f["rgb"].astype(np.float32) / 255.0 for f in frames]).transpose(0, 3, 1, 2)
    t = np.stack([f["mask_gt"].astype(np.float32) for f in frames])[:, None]
    return np.ascontiguousarray(x), np.ascontiguousarray(t)


def sample_index(numel, n=256, seed=3):
    return np.sort(np.random.default_rng(seed + numel % 997).choice(numel, size=min(n, numel), replace=False))


def add_bf16_rounding(model):
    """Hooks that round to bfloat16 (forward values and the gradients flowing back through the same point) wherever the
    kernels of csrc/train.cu keep a bf16 tensor; convolution weights are rounded in place."""
    import torch

    class Round(torch.autograd.Function):
        @staticmethod
        def forward(ctx, x):
            return x.to(torch.bfloat16).float()

        @staticmethod
        def backward(ctx, g):
            return g.to(torch.bfloat16).float()

    for m in model.modules():
        if isinstance(m, torch.nn.Conv2d) and m.kernel_size == (3, 3):
            m.weight.data = m.weight.data.to(torch.bfloat16).float()
            m.register_forward_pre_hook(lambda mod, inp: (Round.apply(inp[0]),))       # X (and dX)
            m.register_forward_hook(lambda mod, inp, out: Round.apply(out))             # Z (and dZ)
        elif isinstance(m, torch.nn.ReLU):
            m.register_forward_hook(lambda mod, inp, out: Round.apply(out))             # Y (and dY)
    return model


def main():
    import importlib.util
    import torch
    from rdp_b200 import synth
    spec = importlib.util.spec_from_file_location("_rdp_reference_segmentation_model", "/root/reference/pkg/segmentation_model.py")
    seg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(seg)
    manifest = {"torch": torch.__version__, "cases": {}}
    for name, wseed, seeds, H, W in CASES:
        rec = {}
        for sfx in ("", "_bf16"):
            torch.manual_seed(0)
            model = seg.UNet(n_channels=3, n_classes=1)
            model.load_state_dict(synth.make_state_dict(wseed, planted=False))
            model.train()
            if sfx:
                add_bf16_rounding(model)
            opt = torch.optim.Adam(model.parameters(), lr=1e-4)
            crit = torch.nn.BCEWithLogitsLoss()
            x, t = (torch.from_numpy(a) for a in train_batch(seeds, H, W))
            opt.zero_grad()
            out = model(x)
            loss = crit(out, t)
            loss.backward()
            rec["loss" + sfx] = np.float64(loss.item())
            li = sample_index(out.numel())
            rec["logits_idx"], rec["logits_val" + sfx] = li, out.detach().reshape(-1)[li].numpy()
            rec["logits_absmax"] = np.float64(out.detach().abs().max())
            for pname, p in model.named_parameters():
                g = p.grad.detach().reshape(-1)
                idx = sample_index(g.numel())
                rec[f"grad_idx/{pname}"], rec[f"grad_val{sfx}/{pname}"] = idx, g[idx].numpy()
                rec[f"grad_absmax{sfx}/{pname}"] = np.float64(g.abs().max())
                rec[f"grad_norm{sfx}/{pname}"] = np.float64(g.double().norm())
            opt.step()
            for bname, b in model.named_buffers():
                rec[f"buffer{sfx}/{bname}"] = b.detach().numpy().copy()
            for pname, p in model.named_parameters():
                idx = rec[f"grad_idx/{pname}"]
                rec[f"param_after{sfx}/{pname}"] = p.detach().reshape(-1)[idx].numpy()
        worst = max(float(np.abs(rec[f"grad_val_bf16/{n}"] - rec[f"grad_val/{n}"]).max() / rec[f"grad_absmax/{n}"])
                    for n, _ in model.named_parameters())
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **rec)
        manifest["cases"][name] = dict(weight_seed=wseed, frame_seeds=list(seeds), H=H, W=W, loss=float(rec["loss"]),
                                       loss_bf16=float(rec["loss_bf16"]), worst_grad_shift_from_bf16_rounding=worst)
        print(name, manifest["cases"][name])
    with open(os.path.join(OUT, "train_manifest.json"), "w") as fh:
        json.dump(manifest, fh, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
