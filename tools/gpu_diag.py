#!/usr/bin/env python
"""Layer-by-layer diagnosis of the CUDA U-Net on a GPU box (development aid, not part of the product).

For every conv it feeds torch (on the GPU, fp32 math) the *same bf16 input the kernel saw* and compares outputs, so
a wrong descriptor / layout shows up at the first layer it affects instead of as a garbled mask.
"""
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "robotic-discovery-platform_b200"))

from rdp_b200 import synth  # noqa: E402
from rdp_b200.unet import UnetEngine, fold_and_pack  # noqa: E402


def main():
    H, W, B = (int(a) for a in (sys.argv[1:4] if len(sys.argv) >= 4 else (48, 64, 2)))
    planted = "--planted" in sys.argv
    dev = torch.device("cuda:0")
    sd = synth.make_state_dict(0, planted=planted)
    batch = synth.make_batch(range(B), H, W, 0.3, "u16")
    eng = UnetEngine(sd, 3, 1, dev)
    rgb = torch.from_numpy(batch["rgb"]).to(dev)
    logits, mask, ws = eng.forward(rgb, want_logits=True, want_mask=True, keep_activations=True)
    torch.cuda.synchronize()
    print("forward done; logits range", float(logits.min()), float(logits.max()), "coverage", float(mask.float().mean()))
    acts = [eng.read_activation(i, B, H, W, ws) for i in range(18)]
    torch.cuda.synchronize()
    table = synth.unet_layer_table(3)
    x0 = (rgb.permute(0, 3, 1, 2).float() / 255.0).to(torch.bfloat16).float()

    def ref_conv(i, x):
        conv, bn, cin, cout = table[i]
        w = sd[f"{conv}.weight"].to(dev)
        scale = sd[f"{bn}.weight"].to(dev) / torch.sqrt(sd[f"{bn}.running_var"].to(dev) + 1e-5)
        shift = sd[f"{bn}.bias"].to(dev) - sd[f"{bn}.running_mean"].to(dev) * scale
        wf = (w * scale[:, None, None, None]).to(torch.bfloat16).float()
        return F.relu(F.conv2d(x, wf, None, padding=1) + shift[None, :, None, None])

    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    inputs = {}
    inputs[0] = x0
    inputs[1] = acts[0]
    for d in range(4):
        inputs[2 * d + 2] = F.max_pool2d(acts[2 * d + 1], 2)
        inputs[2 * d + 3] = acts[2 * d + 2]
    for j in range(4):
        low, skip = acts[9 + 2 * j], acts[7 - 2 * j]
        up = F.interpolate(low, scale_factor=2, mode="bilinear", align_corners=True).to(torch.bfloat16).float()
        dy, dx = skip.shape[2] - up.shape[2], skip.shape[3] - up.shape[3]
        up = F.pad(up, [dx // 2, dx - dx // 2, dy // 2, dy - dy // 2])
        inputs[10 + 2 * j] = torch.cat([skip, up], dim=1)
        inputs[11 + 2 * j] = acts[10 + 2 * j]
    bad = 0
    for i in range(18):
        ref = ref_conv(i, inputs[i])
        got = acts[i]
        err = (got - ref).abs()
        scale = ref.abs().max().item() + 1e-9
        rel = err.max().item() / scale
        flag = "OK " if rel < 2e-2 else "BAD"
        bad += flag == "BAD"
        print(f"conv {i:2d} {table[i][0]:42s} cin {table[i][2]:4d} cout {table[i][3]:3d} shape {tuple(got.shape)} "
              f"max|err| {err.max().item():.4e} rel {rel:.3e} mean|err| {err.mean().item():.3e} refmax {scale:.3f} {flag}")
        if flag == "BAD":
            # where is it wrong?  per-channel-block and per-position summaries
            e = err.amax(dim=(0, 2, 3))
            print("   worst out-channels:", torch.topk(e, 8).indices.tolist(), "per 64-block max:",
                  [round(e[k:k + 64].max().item(), 3) for k in range(0, e.numel(), 64)])
            es = err.amax(dim=(0, 1))
            print("   err by row (first 16):", [round(v, 3) for v in es.amax(dim=1)[:16].tolist()])
            print("   err by col (first 16):", [round(v, 3) for v in es.amax(dim=0)[:16].tolist()])
            print("   got[0,0,:4,:8]:", got[0, 0, :4, :8].tolist())
            print("   ref[0,0,:4,:8]:", ref[0, 0, :4, :8].tolist())
    # head
    y = acts[17]
    ref_logits = F.conv2d(ref_conv(17, inputs[17]), sd["outc.conv.weight"].to(dev), sd["outc.conv.bias"].to(dev))
    print("logits max|err| vs fp32 head on own features:", (logits - ref_logits).abs().max().item())
    print("mask vs rule:", float(((logits > 8.940697e-8).to(torch.uint8) == mask).float().mean()))
    print("DIAG", "FAILED" if bad else "PASSED", f"({bad} bad layers)")
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
