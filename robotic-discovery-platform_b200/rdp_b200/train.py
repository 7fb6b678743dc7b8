"""Training-mode forward and backward of the U-Net through the C ABI (`rdp_unet_train_forward/backward`).

Replaces what `outputs = model(images)` and `loss.backward()` do in the reference's training loop
(`scripts/train_segmenter.py:151-165`) when the model is the drop-in `pkg.segmentation_model.UNet` in `train()` mode:
batch-statistics BatchNorm (with the running-statistics update, momentum 0.1, unbiased variance), every gradient of
every parameter.  The loss (`nn.BCEWithLogitsLoss`) and the optimiser (`torch.optim.Adam`) remain torch's own, as in that
loop; they see an ordinary autograd graph with one node for the whole network.  CUDA only, no fallback.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from .synth import unet_layer_table


class TrainLayer(C.Structure):
    _fields_ = [("w", C.c_void_p), ("w_t", C.c_void_p), ("gamma", C.c_void_p), ("beta", C.c_void_p),
                ("cin", C.c_int32), ("cout", C.c_int32)]


class TrainWeights(C.Structure):
    _fields_ = [("n_channels", C.c_int32), ("n_classes", C.c_int32), ("conv", TrainLayer * 18),
                ("outc_weight", C.c_void_p), ("outc_bias", C.c_void_p), ("bn_eps", C.c_float)]


class TrainBatchStats(C.Structure):
    _fields_ = [("mean", C.c_void_p * 18), ("var_biased", C.c_void_p * 18)]


class TrainGrads(C.Structure):
    _fields_ = [("dw", C.c_void_p * 18), ("dgamma", C.c_void_p * 18), ("dbeta", C.c_void_p * 18),
                ("d_outc_weight", C.c_void_p), ("d_outc_bias", C.c_void_p)]


def _bind(lib):
    if getattr(lib, "_rdp_train_bound", False):
        return
    lib.rdp_unet_train_workspace_bytes.argtypes = [C.POINTER(TrainWeights), C.c_int, C.c_int, C.c_int, C.POINTER(C.c_size_t)]
    lib.rdp_unet_train_forward.argtypes = [C.POINTER(TrainWeights), C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                           C.POINTER(TrainBatchStats), C.c_void_p, C.c_size_t, C.c_void_p]
    lib.rdp_unet_train_backward.argtypes = [C.POINTER(TrainWeights), C.c_void_p, C.c_int, C.c_int, C.c_int,
                                            C.POINTER(TrainGrads), C.c_void_p, C.c_size_t, C.c_void_p]
    lib.rdp_unet_train_pack.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    for n in ("rdp_unet_train_pack", "rdp_unet_train_workspace_bytes", "rdp_unet_train_forward", "rdp_unet_train_backward"):
        getattr(lib, n).restype = C.c_int
    lib._rdp_train_bound = True


def pack_for_training(params: dict, n_channels: int):
    """fp32 OIHW conv weights -> the two bf16 layouts the kernels read (see include/rdp_b200.h, RdpTrainLayer)."""
    ws, wts = [], []
    for i, (conv, _bn, cin, cout) in enumerate(unet_layer_table(n_channels)):
        w = params[f"{conv}.weight"].detach().float()
        fwd = w.permute(0, 2, 3, 1).reshape(cout, 9, cin)                     # [cout][tap][cin]
        if i == 0:
            first = torch.zeros(cout, 64, device=w.device)
            first[:, : 9 * cin] = fwd.reshape(cout, 9 * cin)
            ws.append(first.to(torch.bfloat16).contiguous())
            wts.append(None)
        else:
            ws.append(fwd.to(torch.bfloat16).contiguous())
            # w_t[ci][3*ky+kx][co] = W[co][ci][2-ky][2-kx]
            wts.append(w.flip(2, 3).permute(1, 2, 3, 0).reshape(cin, 9, cout).to(torch.bfloat16).contiguous())
    return ws, wts


_LEVEL = [0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 3, 3, 2, 2, 1, 1, 0, 0]


def _packed_buffers(model, dev, table):
    """bf16 operand buffers of the 18 convolutions, allocated once per (model, device) and refilled every step by
    `rdp_unet_train_pack` (the optimiser changes the fp32 parameters in place between steps)."""
    cache = model.__dict__.setdefault("_rdp_train_buffers", {})
    key = (dev.type, dev.index)
    if key not in cache:
        ws, wts = [], []
        for i, (_, _, cin, cout) in enumerate(table):
            ws.append(torch.empty((cout, 64) if i == 0 else (cout, 9, cin), dtype=torch.bfloat16, device=dev))
            wts.append(None if i == 0 else torch.empty((cin, 9, cout), dtype=torch.bfloat16, device=dev))
        cache[key] = (ws, wts)
    return cache[key]


class _TrainStep(torch.autograd.Function):
    """One autograd node for the whole network: inputs are the image batch and every parameter (in
    `unet_layer_table` order: conv weight, BN weight, BN bias per stage, then OutConv weight and bias)."""

    @staticmethod
    def forward(ctx, model, x, *params):
        lib = _lib.load()
        _bind(lib)
        dev = x.device
        n_channels, n_classes = model.n_channels, model.n_classes
        table = unet_layer_table(n_channels)
        named = {}
        k = 0
        for conv, bn, _cin, _cout in table:
            named[f"{conv}.weight"], named[f"{bn}.weight"], named[f"{bn}.bias"] = params[k], params[k + 1], params[k + 2]
            k += 3
        ow, ob = params[k], params[k + 1]
        B, _, H, W = x.shape
        with torch.cuda.device(dev):
            stream = torch.cuda.current_stream(dev).cuda_stream
            ws, wts = _packed_buffers(model, dev, table)
            for i, (conv, _, cin, cout) in enumerate(table):
                w32 = named[f"{conv}.weight"].detach().float().contiguous()
                _lib.check(lib.rdp_unet_train_pack(w32.data_ptr(), cout, cin, int(i == 0), ws[i].data_ptr(),
                                                   wts[i].data_ptr() if wts[i] is not None else None, stream), "rdp_unet_train_pack")
            gam = [named[f"{bn}.weight"].detach().float().contiguous() for _, bn, _, _ in table]
            bet = [named[f"{bn}.bias"].detach().float().contiguous() for _, bn, _, _ in table]
            owc = ow.detach().float().reshape(n_classes, 64).contiguous()
            obc = ob.detach().float().contiguous()
            desc = TrainWeights()
            desc.n_channels, desc.n_classes, desc.bn_eps = n_channels, n_classes, 1e-5
            for i, (_, _, cin, cout) in enumerate(table):
                L = desc.conv[i]
                L.w, L.w_t = ws[i].data_ptr(), (wts[i].data_ptr() if wts[i] is not None else None)
                L.gamma, L.beta, L.cin, L.cout = gam[i].data_ptr(), bet[i].data_ptr(), cin, cout
            desc.outc_weight, desc.outc_bias = owc.data_ptr(), obc.data_ptr()
            need = C.c_size_t()
            _lib.check(lib.rdp_unet_train_workspace_bytes(C.byref(desc), B, H, W, C.byref(need)), "rdp_unet_train_workspace_bytes")
            wsp = torch.empty(need.value, dtype=torch.uint8, device=dev)
            logits = torch.empty((B, n_classes, H, W), dtype=torch.float32, device=dev)
            couts = [cout for _, _, _, cout in table]
            stats = torch.empty(2 * sum(couts), dtype=torch.float32, device=dev)
            means = list(stats[: sum(couts)].split(couts))
            vars_ = list(stats[sum(couts):].split(couts))
            st = TrainBatchStats()
            for i in range(18):
                st.mean[i], st.var_biased[i] = means[i].data_ptr(), vars_[i].data_ptr()
            xin = x.detach().float().contiguous()
            _lib.check(lib.rdp_unet_train_forward(C.byref(desc), xin.data_ptr(), B, H, W, logits.data_ptr(), C.byref(st),
                                                  wsp.data_ptr(), wsp.numel(), stream), "rdp_unet_train_forward")
            # running statistics, as nn.BatchNorm2d updates them in train mode (momentum 0.1, unbiased variance)
            n = B * H * W
            with torch.no_grad():
                mods = [model.get_submodule(bn) for _, bn, _, _ in table]
                rm, rv = [m.running_mean for m in mods], [m.running_var for m in mods]
                unbias = [0.1 * ni / max(ni - 1, 1) for ni in (n >> (2 * _LEVEL[i]) for i in range(18))]
                torch._foreach_mul_(rm, 0.9)
                torch._foreach_add_(rm, means, alpha=0.1)
                torch._foreach_mul_(rv, 0.9)
                torch._foreach_add_(rv, torch._foreach_mul(vars_, unbias))
                torch._foreach_add_([m.num_batches_tracked for m in mods], 1)
        ctx.keep = (desc, ws, wts, gam, bet, owc, obc, wsp, xin)
        ctx.shape = (B, H, W, n_channels, n_classes)
        ctx.param_shapes = [p.shape for p in params]
        return logits

    @staticmethod
    def backward(ctx, dlogits):
        lib = _lib.load()
        desc, ws, wts, gam, bet, owc, obc, wsp, xin = ctx.keep
        B, H, W, n_channels, n_classes = ctx.shape
        dev = dlogits.device
        table = unet_layer_table(n_channels)
        with torch.cuda.device(dev):
            g = TrainGrads()
            couts = [cout for _, _, _, cout in table]
            small = torch.empty(2 * sum(couts), dtype=torch.float32, device=dev)
            dgs, dbs = list(small[: sum(couts)].split(couts)), list(small[sum(couts):].split(couts))
            dws = []
            for i, (_, _, cin, cout) in enumerate(table):
                dws.append(torch.empty((cout, cin, 3, 3), dtype=torch.float32, device=dev))
                g.dw[i], g.dgamma[i], g.dbeta[i] = dws[i].data_ptr(), dgs[i].data_ptr(), dbs[i].data_ptr()
            dow = torch.empty((n_classes, 64), dtype=torch.float32, device=dev)
            dob = torch.empty(n_classes, dtype=torch.float32, device=dev)
            g.d_outc_weight, g.d_outc_bias = dow.data_ptr(), dob.data_ptr()
            dl = dlogits.detach().float().contiguous()
            stream = torch.cuda.current_stream(dev).cuda_stream
            _lib.check(lib.rdp_unet_train_backward(C.byref(desc), dl.data_ptr(), B, H, W, C.byref(g), wsp.data_ptr(), wsp.numel(),
                                                   stream), "rdp_unet_train_backward")
        grads = []
        for i in range(18):
            grads += [dws[i], dgs[i], dbs[i]]
        grads += [dow.reshape(ctx.param_shapes[-2]), dob]
        return (None, None, *grads)


def train_forward(model, x: torch.Tensor) -> torch.Tensor:
    """`model(images)` in train mode: logits (B, n_classes, H, W) float32 with a graph back to every parameter."""
    if not x.is_cuda:
        raise _lib.RdpError("training runs on a CUDA device only; there is no CPU path")
    if x.shape[2] % 16 or x.shape[3] % 16:
        raise _lib.RdpError(f"training needs H and W divisible by 16, got {tuple(x.shape)} (the reference trains at 256x256)")
    table = unet_layer_table(model.n_channels)
    params = []
    for conv, bn, _, _ in table:
        params += [model.get_parameter(f"{conv}.weight"), model.get_parameter(f"{bn}.weight"), model.get_parameter(f"{bn}.bias")]
    params += [model.outc.conv.weight, model.outc.conv.bias]
    return _TrainStep.apply(model, x, *params)
