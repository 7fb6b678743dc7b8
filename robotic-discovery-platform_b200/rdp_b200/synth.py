"""Seeded synthetic RGB-D frames and U-Net weights for parity tests and the benchmark.

There is no dataset or checkpoint in this environment, so every run fabricates its inputs:

* frames follow the camera format of the reference (`pkg/camera.py:62-63`: uint8 colour + z16 depth,
  640x480) with a curved "actuator" band in front of a flat background;
* weights follow the `state_dict` layout of the reference U-Net (`pkg/segmentation_model.py:91-107`).

Two depth flavours exist on purpose.  ``u16`` is what the service sees (millimetre-quantised, so x/y ties are
common and the reference's unstable argsort/argpartition is not a deterministic oracle).  ``f32`` carries
continuous noise in metres with ``depth_scale=1.0`` and is tie-free, which makes the unmodified reference a
deterministic oracle for the selection stages.
"""
from __future__ import annotations

import math
from collections import OrderedDict

import numpy as np

__all__ = [
    "camera_intrinsics", "make_frame", "make_batch", "band_mask",
    "unet_layer_table", "make_state_dict", "PLANTED_DEFAULT",
]


def camera_intrinsics(H: int, W: int) -> np.ndarray:
    """Pinhole K for a RealSense-like colour stream, scaled with resolution."""
    f = 600.0 * W / 640.0
    return np.array([[f, 0.0, W / 2.0 - 0.5], [0.0, f, H / 2.0 - 0.5], [0.0, 0.0, 1.0]], dtype=np.float64)


def band_mask(H: int, W: int, fill: float, bend: float = 0.15) -> np.ndarray:
    """Boolean (H, W) mask of a parabolic band covering 10-90 % of the width with ~`fill` coverage."""
    u = np.arange(W, dtype=np.float64)[None, :]
    v = np.arange(H, dtype=np.float64)[:, None]
    cx = W / 2.0 - 0.5
    s = (u - cx) / (0.4 * W)
    centre = H * (0.5 - bend / 2 + bend * s * s)
    thick = fill * H / 0.8
    inside_x = (u >= 0.1 * W) & (u < 0.9 * W)
    return inside_x & (np.abs(v - centre) <= thick / 2.0)


def make_frame(seed: int, H: int = 480, W: int = 640, fill: float = 0.15, depth: str = "u16") -> dict:
    """One synthetic RGB-D frame.

    Returns dict with ``rgb`` (H,W,3) uint8, ``depth`` (H,W) uint16 [mm] or float32 [m], ``mask_gt`` (H,W) uint8,
    ``K`` (3,3) float64 and ``depth_scale`` (Python float, as the server passes it: `server.py:94`).
    """
    rng = np.random.default_rng(seed)
    obj = band_mask(H, W, fill, bend=0.15 + 0.05 * ((seed % 5) - 2) / 2.0)
    K = camera_intrinsics(H, W)
    u = np.arange(W, dtype=np.float64)[None, :]
    s = (u - K[0, 2]) / (0.4 * W)
    z_obj = 0.40 + 0.05 * s * s + np.zeros((H, 1))
    z = np.where(obj, z_obj, 0.80) + rng.normal(0.0, 1e-3, size=(H, W))
    holes = rng.random((H, W)) < 0.02
    z = np.where(holes, 0.0, z)
    if depth == "u16":
        d = np.clip(np.rint(z / 1e-3), 0, 65535).astype(np.uint16)
        scale = 0.001
    elif depth == "f32":
        d = z.astype(np.float32)
        scale = 1.0
    else:
        raise ValueError(f"unknown depth flavour {depth!r}")
    base = np.where(obj[..., None], np.array([200.0, 60.0, 50.0]), np.array([40.0, 90.0, 140.0]))
    rgb = np.clip(np.rint(base + rng.normal(0.0, 12.0, size=(H, W, 3))), 0, 255).astype(np.uint8)
    return {"rgb": rgb, "depth": d, "mask_gt": obj.astype(np.uint8), "K": K, "depth_scale": scale,
            "fill": float(obj.mean())}


def make_batch(seeds, H: int = 480, W: int = 640, fill: float = 0.15, depth: str = "u16") -> dict:
    frames = [make_frame(int(s), H, W, fill, depth) for s in seeds]
    return {
        "rgb": np.stack([f["rgb"] for f in frames]),
        "depth": np.stack([f["depth"] for f in frames]),
        "mask_gt": np.stack([f["mask_gt"] for f in frames]),
        "K": frames[0]["K"],
        "depth_scale": frames[0]["depth_scale"],
    }


def unet_layer_table(n_channels: int = 3, n_classes: int = 1):
    """(state_dict prefix, Cin, Cout) of the 18 3x3 convs in forward order (`segmentation_model.py:97-106`)."""
    t = [("inc.double_conv", n_channels, 64, 64)]
    for i, (ci, co) in enumerate([(64, 128), (128, 256), (256, 512), (512, 512)], start=1):
        t.append((f"down{i}.maxpool_conv.1.double_conv", ci, co, co))
    for i, (ci, co) in enumerate([(1024, 256), (512, 128), (256, 64), (128, 64)], start=1):
        t.append((f"up{i}.conv.double_conv", ci, ci // 2, co))
    out = []
    for prefix, cin, mid, cout in t:
        out.append((f"{prefix}.0", f"{prefix}.1", cin, mid))
        out.append((f"{prefix}.3", f"{prefix}.4", mid, cout))
    return out


PLANTED_DEFAULT = dict(gain=4.0, colour_gain=2.0, theta=0.5, noise_w=0.05)


def make_state_dict(seed: int = 0, n_channels: int = 3, n_classes: int = 1, planted: dict | bool | None = None):
    """Seeded `state_dict` with the reference's keys, fp32, OIHW.

    Default PyTorch init gives a constant mask, so convs are He-normal, BN statistics are drawn
    away from identity (so that folding is really exercised) and, when ``planted`` is set, feature channel 0
    carries (R - B) from the input through `inc` and the `up4` skip to `outc`, which makes logits bimodal:
    object pixels sit near +gain*(colour-theta), background near -gain*theta, and every other channel adds
    zero-mean noise of relative weight ``noise_w``.
    """
    import torch

    g = torch.Generator().manual_seed(1000 + seed)
    sd = OrderedDict()
    for conv, bn, cin, cout in unet_layer_table(n_channels, n_classes):
        std = math.sqrt(2.0 / (9 * cin))
        sd[f"{conv}.weight"] = torch.randn(cout, cin, 3, 3, generator=g) * std
        sd[f"{bn}.weight"] = 0.75 + 0.5 * torch.rand(cout, generator=g)
        sd[f"{bn}.bias"] = 0.2 * torch.randn(cout, generator=g)
        sd[f"{bn}.running_mean"] = 0.2 * torch.randn(cout, generator=g)
        sd[f"{bn}.running_var"] = 0.5 + torch.rand(cout, generator=g)
        sd[f"{bn}.num_batches_tracked"] = torch.tensor(0, dtype=torch.long)
    sd["outc.conv.weight"] = torch.randn(n_classes, 64, 1, 1, generator=g) * math.sqrt(1.0 / 64)
    sd["outc.conv.bias"] = torch.zeros(n_classes)
    if planted:
        p = dict(PLANTED_DEFAULT)
        if isinstance(planted, dict):
            p.update(planted)
        eps = 1e-5

        def plant(conv, bn, src_weights):
            w = sd[f"{conv}.weight"]
            w[0].zero_()
            for ci, val in src_weights:
                w[0, ci, 1, 1] = val
            sd[f"{bn}.weight"][0] = 1.0
            sd[f"{bn}.bias"][0] = 0.0
            sd[f"{bn}.running_mean"][0] = 0.0
            sd[f"{bn}.running_var"][0] = 1.0 - eps

        a = p["colour_gain"]
        plant("inc.double_conv.0", "inc.double_conv.1", [(0, a), (2, -a)])
        plant("inc.double_conv.3", "inc.double_conv.4", [(0, 1.0)])
        plant("up4.conv.double_conv.0", "up4.conv.double_conv.1", [(0, 1.0)])   # skip channels come first
        plant("up4.conv.double_conv.3", "up4.conv.double_conv.4", [(0, 1.0)])
        wo = sd["outc.conv.weight"]
        wo.mul_(p["noise_w"] * math.sqrt(64.0))
        wo[:, 0, 0, 0] = p["gain"]
        sd["outc.conv.bias"].fill_(-p["gain"] * p["theta"])
    return sd
