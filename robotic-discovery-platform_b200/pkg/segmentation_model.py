"""`pkg.segmentation_model` of the reference, served by the sm_100a U-Net engine.

Same class tree and parameter names as the reference (`pkg/segmentation_model.py:24-107`), so that

* `UNet(3, 1).load_state_dict(torch.load(path))` accepts reference checkpoints unchanged (`vision_utils.py:63-64`);
* `mlflow.pytorch.load_model(...)`, which unpickles a whole module by class path (`server.py:81-82`), finds
  `pkg.segmentation_model.{UNet, DoubleConv, Down, Up, OutConv}`.

The modules only *hold* parameters.  `UNet.forward` folds BatchNorm into bf16 weights (lazily, and again whenever a
parameter changes) and runs the hand-written CUDA path on a CUDA input, like the server provides
(`server.py:83,121-123`); in `train()` mode it runs the training step's forward and backward kernels
(`rdp_b200.train`, `scripts/train_segmenter.py:151-165`).  CPU tensors raise instead of silently running something else; `bilinear=False` constructs like the reference and raises in `forward` as it does there.
"""
from __future__ import annotations

import threading

import torch
from torch import nn

__all__ = ["DoubleConv", "Down", "Up", "OutConv", "UNet"]


class _ParameterHolder(nn.Module):
    def forward(self, *args, **kwargs):  # pragma: no cover - guard
        raise RuntimeError(f"{type(self).__name__} only stores parameters here; call the enclosing UNet on a CUDA tensor")


def _conv_bn_relu(cin: int, cout: int):
    return [nn.Conv2d(cin, cout, kernel_size=3, padding=1, bias=False), nn.BatchNorm2d(cout), nn.ReLU(inplace=True)]


class DoubleConv(_ParameterHolder):
    """Two (3x3 conv, BatchNorm, ReLU) stages; submodule `double_conv` with indices 0,1,3,4 carrying parameters."""

    def __init__(self, in_channels, out_channels, mid_channels=None):
        super().__init__()
        mid = mid_channels or out_channels
        self.double_conv = nn.Sequential(*_conv_bn_relu(in_channels, mid), *_conv_bn_relu(mid, out_channels))


class Down(_ParameterHolder):
    """2x2 max-pool followed by DoubleConv; parameters under `maxpool_conv.1`."""

    def __init__(self, in_channels, out_channels):
        super().__init__()
        self.maxpool_conv = nn.Sequential(nn.MaxPool2d(2), DoubleConv(in_channels, out_channels))


class Up(_ParameterHolder):
    """Bilinear x2 upsample (align_corners=True), pad, concat [skip, up], DoubleConv; parameters under `conv`."""

    def __init__(self, in_channels, out_channels, bilinear=True):
        super().__init__()
        if bilinear:
            self.up = nn.Upsample(scale_factor=2, mode="bilinear", align_corners=True)
            self.conv = DoubleConv(in_channels, out_channels, in_channels // 2)
        else:
            # Constructs like the reference (`segmentation_model.py:63-66`), so such a checkpoint still loads; its
            # forward cannot run there either (the DoubleConv expects in_channels // 2, the concat has in_channels).
            self.up = nn.ConvTranspose2d(in_channels, in_channels // 2, kernel_size=2, stride=2)
            self.conv = DoubleConv(in_channels // 2, out_channels)


class OutConv(_ParameterHolder):
    def __init__(self, in_channels, out_channels):
        super().__init__()
        self.conv = nn.Conv2d(in_channels, out_channels, kernel_size=1)


class UNet(nn.Module):
    """U-Net with the reference's constructor, attributes and `state_dict` keys, executed on sm_100a kernels."""

    def __init__(self, n_channels, n_classes, bilinear=True):
        super().__init__()
        self.n_channels = n_channels
        self.n_classes = n_classes
        self.bilinear = bilinear
        self.inc = DoubleConv(n_channels, 64)
        self.down1 = Down(64, 128)
        self.down2 = Down(128, 256)
        self.down3 = Down(256, 512)
        factor = 2 if bilinear else 1
        self.down4 = Down(512, 1024 // factor)
        self.up1 = Up(1024, 512 // factor, bilinear)
        self.up2 = Up(512, 256 // factor, bilinear)
        self.up3 = Up(256, 128 // factor, bilinear)
        self.up4 = Up(128, 64, bilinear)
        self.outc = OutConv(64, n_classes)

    # ---- engine management -------------------------------------------------------------------------------
    def _fingerprint(self):
        return tuple((t.data_ptr(), t._version, str(t.device)) for t in list(self.parameters()) + list(self.buffers()))

    def _engine_for(self, device):
        d = self.__dict__                      # objects unpickled from a reference checkpoint lack our attributes
        lock = d.setdefault("_rdp_lock", threading.Lock())
        with lock:
            fp = self._fingerprint()
            eng = d.get("_rdp_engine")
            if eng is None or d.get("_rdp_fingerprint") != fp or eng.device != device:
                from rdp_b200.unet import UnetEngine
                eng = UnetEngine(self.state_dict(), self.n_channels, self.n_classes, device)
                d["_rdp_engine"], d["_rdp_fingerprint"] = eng, fp
            return eng

    def __getstate__(self):
        state = dict(self.__dict__)
        for k in ("_rdp_engine", "_rdp_fingerprint", "_rdp_lock", "_rdp_train_buffers"):
            state.pop(k, None)
        return state

    def _check(self, x, inference_only=False):
        if self.training and inference_only:
            raise RuntimeError("this entry point is the inference fast path: call model.eval() first (server.py:83)")
        if not getattr(self, "bilinear", True):
            # the reference raises here too, from up1's first convolution (`segmentation_model.py:65,75`)
            raise RuntimeError("UNet(bilinear=False): Up.forward concatenates 1024 channels into a DoubleConv built for 512 "
                               "(expected input to have 512 channels, but got 1024 channels instead)")
        if not x.is_cuda:
            raise RuntimeError("UNet (rdp_b200) has no CPU path: move the model and the input to a CUDA device")

    # ---- reference interface -----------------------------------------------------------------------------
    def forward(self, x):
        """(B, n_channels, H, W) float32 on CUDA -> logits (B, n_classes, H, W) float32 (`segmentation_model.py:109`).

        In `eval()` mode this is the inference engine (BatchNorm folded into bf16 weights).  In `train()` mode - the
        reference's training loop, `scripts/train_segmenter.py:151-165` - BatchNorm uses the statistics of the batch and
        updates its running statistics, and the returned logits carry an autograd graph to every parameter
        (`rdp_b200.train`), so `criterion(outputs, masks).backward()` and `optimizer.step()` work unchanged."""
        self._check(x)
        if self.training and torch.is_grad_enabled():
            from rdp_b200.train import train_forward
            return train_forward(self, x)
        if self.training:
            raise RuntimeError("UNet (rdp_b200): train() mode without autograd is not a path the reference uses; "
                               "call model.eval() for inference (server.py:83)")
        with torch.no_grad():
            logits, _, _ = self._engine_for(x.device).forward(x.float(), want_logits=True, want_mask=False)
        return logits

    # ---- additive fast path ------------------------------------------------------------------------------
    @torch.no_grad()
    def forward_mask(self, rgb_u8, *, want_logits=False, workspace=None):
        """(B,H,W,C) uint8 RGB on CUDA -> (mask uint8 (B,n_classes,H,W), logits or None, workspace).

        The mask is `sigmoid(logit) > 0.5` (`server.py:124`) written straight from the last convolution's epilogue.
        """
        self._check(rgb_u8, inference_only=True)
        logits, mask, ws = self._engine_for(rgb_u8.device).forward(rgb_u8, want_logits=want_logits, want_mask=True,
                                                                   workspace=workspace)
        return mask, logits, ws
