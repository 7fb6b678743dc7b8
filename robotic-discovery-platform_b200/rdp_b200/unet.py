"""Device-side U-Net engine: weight folding/packing and the batched forward through the C ABI.

Replaces `UNet.forward` (reference `pkg/segmentation_model.py:109-120`) and, for the fast path, the server's mask
rule (`services/vision_analysis/server.py:124`).  Weights arrive as a reference `state_dict` (fp32, OIHW); eval-mode
BatchNorm is folded into the convolution (scale into the bf16 weights, shift kept in fp32 for the epilogue).
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from .synth import unet_layer_table

BN_EPS = 1e-5


def fold_and_pack(state_dict, n_channels: int, device) -> tuple[list, list]:
    """-> (weights bf16 [cout][9][cin] (first layer [cout][64], k = tap*C + c), shifts fp32 [cout]) on `device`."""
    ws, shifts = [], []
    for i, (conv, bn, cin, cout) in enumerate(unet_layer_table(n_channels)):
        w = state_dict[f"{conv}.weight"].detach().to(device=device, dtype=torch.float32)
        g = state_dict[f"{bn}.weight"].detach().to(device=device, dtype=torch.float32)
        b = state_dict[f"{bn}.bias"].detach().to(device=device, dtype=torch.float32)
        mu = state_dict[f"{bn}.running_mean"].detach().to(device=device, dtype=torch.float32)
        var = state_dict[f"{bn}.running_var"].detach().to(device=device, dtype=torch.float32)
        scale = g / torch.sqrt(var + BN_EPS)
        w = w * scale[:, None, None, None]
        packed = w.permute(0, 2, 3, 1).reshape(cout, 9, cin)          # [cout][tap = 3*ky+kx][cin]
        if i == 0:
            first = torch.zeros(cout, 64, device=device, dtype=torch.float32)
            first[:, : 9 * cin] = packed.reshape(cout, 9 * cin)
            packed = first
        ws.append(packed.contiguous().to(torch.bfloat16))
        shifts.append((b - mu * scale).contiguous())
    return ws, shifts


class UnetEngine:
    """Immutable after construction; safe to share between threads (each call brings its own workspace)."""

    def __init__(self, state_dict, n_channels: int = 3, n_classes: int = 1, device="cuda"):
        self.lib = _lib.load()
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise _lib.RdpError("UnetEngine needs a CUDA device; there is no CPU path")
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        self.n_channels, self.n_classes = n_channels, n_classes
        with torch.cuda.device(self.device):
            self._w, self._shift = fold_and_pack(state_dict, n_channels, self.device)
            self._ow = state_dict["outc.conv.weight"].detach().to(self.device, torch.float32).reshape(n_classes, 64).contiguous()
            self._ob = state_dict["outc.conv.bias"].detach().to(self.device, torch.float32).contiguous()
            desc = _lib.UnetWeights()
            desc.n_channels, desc.n_classes = n_channels, n_classes
            for i, (_, _, cin, cout) in enumerate(unet_layer_table(n_channels)):
                desc.conv[i].weight = self._w[i].data_ptr()
                desc.conv[i].shift = self._shift[i].data_ptr()
                desc.conv[i].cin, desc.conv[i].cout = cin, cout
            desc.outc_weight, desc.outc_bias = self._ow.data_ptr(), self._ob.data_ptr()
            torch.cuda.synchronize(self.device)
            handle = C.c_void_p()
            _lib.check(self.lib.rdp_unet_create(C.byref(desc), C.byref(handle)), "rdp_unet_create")
        self._handle = handle

    def __del__(self):
        h, self._handle = getattr(self, "_handle", None), None
        if h:
            self.lib.rdp_unet_destroy(h)

    @property
    def launches_per_forward(self) -> int:
        return self.lib.rdp_unet_launch_count(self._handle)

    def workspace_bytes(self, B: int, H: int, W: int, keep_activations: bool = False) -> int:
        """Bytes one forward needs: buffers with disjoint lifetimes share storage unless every activation is kept."""
        n = C.c_size_t()
        flags = _lib.IN_FLAG_KEEP_ACTIVATIONS if keep_activations else 0
        _lib.check(self.lib.rdp_unet_workspace_bytes_ex(self._handle, flags, B, H, W, C.byref(n)), "rdp_unet_workspace_bytes_ex")
        return n.value

    def new_workspace(self, B: int, H: int, W: int, keep_activations: bool = False) -> torch.Tensor:
        return torch.empty(self.workspace_bytes(B, H, W, keep_activations), dtype=torch.uint8, device=self.device)

    LAUNCH_NAMES = (["im2col", "conv inc.0", "conv inc.3"]
                    + [f"{k} down{d}{s}" for d in range(1, 5) for k, s in (("maxpool", ""), ("conv", ".0"), ("conv", ".3"))]
                    + [f"{k} up{u}{s}" for u in range(1, 5) for k, s in (("upsample", ""), ("conv", ".0"), ("conv", ".3"))])

    def forward(self, x: torch.Tensor, *, want_logits: bool = True, want_mask: bool = False, workspace=None,
                launch_ms: list | None = None, keep_activations: bool = False):
        """x: (B,H,W,C) uint8 RGB or (B,C,H,W) float32, on the engine's device.

        Returns (logits fp32 (B,n_classes,H,W) or None, mask uint8 (B,n_classes,H,W) or None, workspace).
        """
        if x.device != self.device:
            raise _lib.RdpError(f"input is on {x.device}, engine on {self.device}; there is no CPU path")
        if x.dtype == torch.uint8 and x.dim() == 4 and x.shape[3] == self.n_channels:
            fmt, (B, H, W) = _lib.IN_RGB_U8_NHWC, x.shape[:3]
        elif x.dtype == torch.float32 and x.dim() == 4 and x.shape[1] == self.n_channels:
            fmt, (B, H, W) = _lib.IN_F32_NCHW, (x.shape[0], x.shape[2], x.shape[3])
        else:
            raise _lib.RdpError(f"unsupported input {tuple(x.shape)} {x.dtype}")
        x = x.contiguous()
        if keep_activations:        # parity hook: read_activation(17) needs the last layer's output stored
            fmt |= _lib.IN_FLAG_KEEP_ACTIVATIONS
        with torch.cuda.device(self.device):
            need = self.workspace_bytes(B, H, W, keep_activations)
            if workspace is None or workspace.numel() < need:
                workspace = torch.empty(need, dtype=torch.uint8, device=self.device)
            logits = torch.empty((B, self.n_classes, H, W), dtype=torch.float32, device=self.device) if want_logits else None
            mask = torch.empty((B, self.n_classes, H, W), dtype=torch.uint8, device=self.device) if want_mask else None
            stream = torch.cuda.current_stream(self.device).cuda_stream
            args = (self._handle, x.data_ptr(), fmt, B, H, W,
                    logits.data_ptr() if logits is not None else None,
                    mask.data_ptr() if mask is not None else None,
                    workspace.data_ptr(), workspace.numel(), stream)
            if launch_ms is None:
                _lib.check(self.lib.rdp_unet_forward(*args), "rdp_unet_forward")
            else:   # measurement aid: CUDA events between launches, blocks until done
                buf = (C.c_float * self.launches_per_forward)()
                _lib.check(self.lib.rdp_unet_forward_profiled(*args, buf), "rdp_unet_forward_profiled")
                launch_ms[:] = list(buf)
        return logits, mask, workspace

    def read_activation(self, index: int, B: int, H: int, W: int, workspace: torch.Tensor) -> torch.Tensor:
        """fp32 NCHW copy of conv output `index` (0..17) of the last forward on `workspace` (parity hook)."""
        lv = [0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 3, 3, 2, 2, 1, 1, 0, 0][index]
        h, w = H, W
        for _ in range(lv):
            h, w = h // 2, w // 2
        cout = unet_layer_table(self.n_channels)[index][3]
        out = torch.empty((B, cout, h, w), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            stream = torch.cuda.current_stream(self.device).cuda_stream
            _lib.check(self.lib.rdp_unet_read_activation(self._handle, index, B, H, W, workspace.data_ptr(),
                                                         out.data_ptr(), stream), "rdp_unet_read_activation")
        return out
