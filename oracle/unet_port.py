"""ORACLE (test infrastructure, not product code): fp32 CPU restatement of the reference U-Net forward.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU-baseline / `--impl reference` legs may import
this module.  The product path (`robotic-discovery-platform_b200/`) never does and has no CPU fallback.

Restates `pkg/segmentation_model.py` of the reference as one state_dict-driven function over `torch.nn.functional`:

* `DoubleConv`  (`segmentation_model.py:24-40`)  -> `_double_conv`
* `Down`        (`:42-52`)                        -> `max_pool2d(2)` + `_double_conv`
* `Up`          (`:54-76`, bilinear branch only)  -> `_up`: bilinear x2 align_corners=True, zero pad to the skip
                                                     size, `cat([skip, up])`, `_double_conv`
* `OutConv`     (`:78-84`)                        -> 1x1 conv with bias
* `UNet.forward`(`:109-120`)                      -> `unet_forward`
* the server's mask rule `sigmoid(out) > 0.5` (`services/vision_analysis/server.py:124`) -> `mask_from_logits`

Parity pin: `oracle/make_golden.py` runs the *live* reference (`/root/reference`, torch 2.11.0) and this port on
the same seeded weights/frames; `tests/test_oracle_unet.py` checks this port against the committed vectors.
The reference itself ships no tests or golden vectors, so that is the only pin there is.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

BN_EPS = 1e-5  # nn.BatchNorm2d default, segmentation_model.py:32


def _conv_bn_relu(x, sd, conv, bn):
    x = F.conv2d(x, sd[f"{conv}.weight"], None, padding=1)
    x = F.batch_norm(x, sd[f"{bn}.running_mean"], sd[f"{bn}.running_var"], sd[f"{bn}.weight"], sd[f"{bn}.bias"],
                     training=False, eps=BN_EPS)
    return F.relu(x)


def _double_conv(x, sd, prefix, taps=None):
    x = _conv_bn_relu(x, sd, f"{prefix}.0", f"{prefix}.1")
    if taps is not None:
        taps[f"{prefix}.0"] = x
    x = _conv_bn_relu(x, sd, f"{prefix}.3", f"{prefix}.4")
    if taps is not None:
        taps[f"{prefix}.3"] = x
    return x


def _up(x_low, x_skip, sd, prefix, taps=None):
    x_low = F.interpolate(x_low, scale_factor=2, mode="bilinear", align_corners=True)
    dy = x_skip.shape[2] - x_low.shape[2]
    dx = x_skip.shape[3] - x_low.shape[3]
    x_low = F.pad(x_low, [dx // 2, dx - dx // 2, dy // 2, dy - dy // 2])
    return _double_conv(torch.cat([x_skip, x_low], dim=1), sd, prefix, taps)


@torch.no_grad()
def unet_forward(sd, x: torch.Tensor, taps: dict | None = None) -> torch.Tensor:
    """fp32 NCHW in -> fp32 logits (B, n_classes, H, W).  `taps` (optional dict) receives every conv output."""
    x = x.to(torch.float32)
    x1 = _double_conv(x, sd, "inc.double_conv", taps)
    x2 = _double_conv(F.max_pool2d(x1, 2), sd, "down1.maxpool_conv.1.double_conv", taps)
    x3 = _double_conv(F.max_pool2d(x2, 2), sd, "down2.maxpool_conv.1.double_conv", taps)
    x4 = _double_conv(F.max_pool2d(x3, 2), sd, "down3.maxpool_conv.1.double_conv", taps)
    x5 = _double_conv(F.max_pool2d(x4, 2), sd, "down4.maxpool_conv.1.double_conv", taps)
    y = _up(x5, x4, sd, "up1.conv.double_conv", taps)
    y = _up(y, x3, sd, "up2.conv.double_conv", taps)
    y = _up(y, x2, sd, "up3.conv.double_conv", taps)
    y = _up(y, x1, sd, "up4.conv.double_conv", taps)
    return F.conv2d(y, sd["outc.conv.weight"], sd["outc.conv.bias"])


def unet_forward_train(sd, x: torch.Tensor) -> torch.Tensor:
    """The same network as `UNet.forward` computes it in `train()` mode (`scripts/train_segmenter.py:151-158`): every
    BatchNorm2d normalises with the statistics of the batch (`segmentation_model.py:31-36`; running statistics are not
    touched here).  Runs under autograd: give `sd` tensors with `requires_grad` to get the reference gradients."""
    def cbr(x, conv, bn):
        x = F.conv2d(x, sd[f"{conv}.weight"], None, padding=1)
        x = F.batch_norm(x, None, None, sd[f"{bn}.weight"], sd[f"{bn}.bias"], training=True, eps=BN_EPS)
        return F.relu(x)

    def dc(x, prefix):
        return cbr(cbr(x, f"{prefix}.0", f"{prefix}.1"), f"{prefix}.3", f"{prefix}.4")

    def up(x_low, x_skip, prefix):
        x_low = F.interpolate(x_low, scale_factor=2, mode="bilinear", align_corners=True)
        dy, dx = x_skip.shape[2] - x_low.shape[2], x_skip.shape[3] - x_low.shape[3]
        x_low = F.pad(x_low, [dx // 2, dx - dx // 2, dy // 2, dy - dy // 2])
        return dc(torch.cat([x_skip, x_low], dim=1), prefix)

    x = x.to(torch.float32)
    x1 = dc(x, "inc.double_conv")
    x2 = dc(F.max_pool2d(x1, 2), "down1.maxpool_conv.1.double_conv")
    x3 = dc(F.max_pool2d(x2, 2), "down2.maxpool_conv.1.double_conv")
    x4 = dc(F.max_pool2d(x3, 2), "down3.maxpool_conv.1.double_conv")
    x5 = dc(F.max_pool2d(x4, 2), "down4.maxpool_conv.1.double_conv")
    y = up(x5, x4, "up1.conv.double_conv")
    y = up(y, x3, "up2.conv.double_conv")
    y = up(y, x2, "up3.conv.double_conv")
    y = up(y, x1, "up4.conv.double_conv")
    return F.conv2d(y, sd["outc.conv.weight"], sd["outc.conv.bias"])


def mask_from_logits(logits: torch.Tensor) -> torch.Tensor:
    """uint8 {0,1}; fp32 sigmoid then `> 0.5`, exactly as the server does (true only for logit >~ 8.9e-8)."""
    return (torch.sigmoid(logits) > 0.5).to(torch.uint8)


def rgb_u8_to_input(rgb_u8) -> torch.Tensor:
    """(B,H,W,3) uint8 -> (B,3,H,W) fp32 in [0,1]; `transforms.ToTensor` semantics (`server.py:108`)."""
    t = torch.as_tensor(rgb_u8)
    return t.permute(0, 3, 1, 2).to(torch.float32) / 255.0
