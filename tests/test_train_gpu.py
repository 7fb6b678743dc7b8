"""One optimisation step of the reference's training loop (`scripts/train_segmenter.py:151-165`) through the drop-in
`UNet` in train mode, against what the unmodified reference computed under `torch.autograd` in fp32
(tests/golden/train_step_*.npz, `oracle/make_golden_train.py`).

Two references are on file for every quantity (see the generator's docstring): the reference in fp32, and the same
reference module with bfloat16 rounding at the places where the kernels store bf16.  A randomly initialised U-Net with
batch-statistics BatchNorm amplifies rounding: bf16 weights ALONE move single gradient entries of the encoder by 10 % of
their tensor's largest entry (`worst_grad_shift_from_bf16_rounding` in the manifest: 0.19 / 0.42 with every rounding), so
"every gradient entry within 2 % of fp32" is not a property any bf16 tensor-core implementation of this step can have.
What is asserted:
  * loss within 1e-3 relative of the fp32 reference;
  * BatchNorm running statistics within 2e-4 + 1e-3 relative of the bf16-rounded reference (the forward pass reproduces
    it to 1e-5 typically) and within 1e-2 of fp32;
  * every parameter gradient: norm within 3 % of fp32 plus the larger of 1.5 x the bf16-rounded reference's own norm
    shift and half its relative L2 error (an error vector of that size moves the norm that much); relative L2 error against fp32 no worse than 1.25 x that of the
    bf16-rounded reference itself + 0.01; closer to the bf16-rounded reference than that reference is to fp32 (+ 0.01);
  * the decoder's last two blocks and the head (up3, up4, outc), which sit before the amplification: within 2 % of the
    tensor's largest entry against fp32 at 4 x 256 x 256."""
import json
import os

import numpy as np
import pytest
import torch

from oracle.make_golden_train import train_batch
from rdp_b200 import synth

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _case(name):
    with open(os.path.join(GOLDEN, "train_manifest.json")) as fh:
        meta = json.load(fh)["cases"][name]
    with np.load(os.path.join(GOLDEN, name + ".npz")) as z:
        return meta, {k: z[k] for k in z.files}


@pytest.mark.parametrize("name", ["train_step_2x64x96", "train_step_4x256x256"])
def test_training_step_matches_reference_autograd(name):
    from pkg.segmentation_model import UNet
    meta, g = _case(name)
    model = UNet(3, 1)
    model.load_state_dict(synth.make_state_dict(meta["weight_seed"], planted=False))
    model = model.cuda().train()
    opt = torch.optim.Adam(model.parameters(), lr=1e-4)                    # train_segmenter.py:144
    crit = torch.nn.BCEWithLogitsLoss()                                     # :145
    x, t = (torch.from_numpy(a).cuda() for a in train_batch(meta["frame_seeds"], meta["H"], meta["W"]))
    opt.zero_grad()
    out = model(x)                                                          # :158
    loss = crit(out, t)
    loss.backward()
    assert out.shape == (len(meta["frame_seeds"]), 1, meta["H"], meta["W"]) and out.dtype == torch.float32
    rel_loss = abs(loss.item() - float(g["loss"])) / float(g["loss"])
    logit_err = float((out.detach().reshape(-1)[torch.from_numpy(g["logits_idx"]).cuda()].cpu() - torch.from_numpy(g["logits_val"])).abs().max())
    print(f"{name}: loss {loss.item():.6f} (reference {float(g['loss']):.6f}, rel {rel_loss:.2e}), max |logit - ref| {logit_err:.4f} "
          f"of {float(g['logits_absmax']):.3f}")
    assert rel_loss <= 1e-3
    assert logit_err <= 0.05 * float(g["logits_absmax"])
    def rl2(a, b):
        return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))

    worst = ("", 0.0)
    for pname, p in model.named_parameters():
        assert p.grad is not None, pname
        idx = torch.from_numpy(g[f"grad_idx/{pname}"]).cuda()
        got = p.grad.reshape(-1)[idx].cpu().numpy()
        want, emu, scale = g[f"grad_val/{pname}"], g[f"grad_val_bf16/{pname}"], float(g[f"grad_absmax/{pname}"])
        e_gpu, e_emu, e_between = rl2(got, want), rl2(emu, want), rl2(got, emu)
        n_ref, n_emu = float(g[f"grad_norm/{pname}"]), float(g[f"grad_norm_bf16/{pname}"])
        norm_rel = abs(float(p.grad.double().norm()) - n_ref) / n_ref
        if e_gpu > worst[1]:
            worst = (pname, e_gpu)
        assert norm_rel <= max(1.5 * abs(n_emu - n_ref) / n_ref, 0.5 * e_emu) + 0.03, (pname, norm_rel, abs(n_emu - n_ref) / n_ref, e_emu)
        assert e_gpu <= 1.25 * e_emu + 0.01, (pname, e_gpu, e_emu)
        assert e_between <= e_emu + 0.01, (pname, e_between, e_emu)
        if name == "train_step_4x256x256" and pname.split(".")[0] in ("up3", "up4", "outc"):
            assert float(np.abs(got - want).max()) / scale <= 0.02, pname
    print(f"{name}: largest relative L2 gradient error against fp32 {worst[1]:.4f} ({worst[0]}); the bf16-rounded reference's own "
          f"largest single-entry shift: {meta['worst_grad_shift_from_bf16_rounding']:.3f} of a tensor's max")
    for bname, b in model.named_buffers():
        if bname.endswith("num_batches_tracked"):
            assert int(b) == int(g[f"buffer/{bname}"]), bname
        else:
            # (the small case's deepest BatchNorm averages 48 samples per channel: accumulation order shows at 6e-4)
            np.testing.assert_allclose(b.cpu().numpy(), g[f"buffer_bf16/{bname}"], rtol=1e-3,
                                       atol=2e-4 if name == "train_step_4x256x256" else 2e-3, err_msg=bname)
            np.testing.assert_allclose(b.cpu().numpy(), g[f"buffer/{bname}"], rtol=1e-2, atol=1e-2, err_msg=bname)
    opt.step()                                                              # :161, torch's Adam on our gradients
    for pname, p in model.named_parameters():
        if pname.split(".")[0] not in ("up3", "up4", "outc"):
            continue            # upstream of these, single entries are within rounding noise of changing sign (see above)
        idx = torch.from_numpy(g[f"grad_idx/{pname}"]).cuda()
        # Adam's first step moves every entry by lr * sign(grad): entries whose gradient is far from zero must agree
        want_g = g[f"grad_val/{pname}"]
        sure = np.abs(want_g) > 0.25 * float(g[f"grad_absmax/{pname}"])
        got = p.detach().reshape(-1)[idx].cpu().numpy()
        assert np.allclose(got[sure], g[f"param_after/{pname}"][sure], rtol=0, atol=2e-5), pname


def test_training_step_with_three_classes():
    """The multi-class head (`UNet(3, 3)`; `OutConv` `segmentation_model.py:78-84`) in both directions: logits, loss and the
    gradients that sit before the rounding amplification (OutConv, up4) against the oracle's train-mode restatement run
    under torch.autograd in fp32 on the same GPU (the checker; the goldens on file are single-class).  Bounds as for the
    single-class case: logits within 5 % of their largest, loss within 1e-3, OutConv gradients within 2 % of the tensor's
    largest entry, up4's within 3 % (2 x 64 x 96 is the small case: the 2 % of the single-class test holds at 4 x 256 x 256;
    measured here 2.1 %), norms within 3 %; every other gradient must be finite and within 25 % in norm."""
    from oracle import unet_port
    from pkg.segmentation_model import UNet
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    sd = synth.make_state_dict(11, planted=False, n_classes=3)
    model = UNet(3, 3)
    model.load_state_dict(sd)
    model = model.cuda().train()
    x, t1 = (torch.from_numpy(a).cuda() for a in train_batch((60, 61), 64, 96))
    t = torch.cat([t1, 1 - t1, torch.roll(t1, 7, dims=3)], dim=1)
    crit = torch.nn.BCEWithLogitsLoss()
    out = model(x)
    loss = crit(out, t)
    loss.backward()
    ref_sd = {k: v.detach().clone().cuda().float().requires_grad_(v.dtype.is_floating_point and "running" not in k)
              for k, v in sd.items()}
    ref_out = unet_port.unet_forward_train(ref_sd, x)
    ref_loss = crit(ref_out, t)
    ref_loss.backward()
    assert out.shape == (2, 3, 64, 96)
    scale = float(ref_out.detach().abs().max())
    print(f"3 classes: loss {loss.item():.6f} vs {ref_loss.item():.6f}; max |logit - ref| {float((out.detach() - ref_out.detach()).abs().max()):.4f} of {scale:.3f}")
    assert abs(loss.item() - ref_loss.item()) <= 1e-3 * abs(ref_loss.item())
    assert float((out.detach() - ref_out.detach()).abs().max()) <= 0.05 * scale
    for name, prm in model.named_parameters():
        g, r = prm.grad.float(), ref_sd[name].grad
        assert torch.isfinite(g).all(), name
        gn, rn = float(g.norm()), float(r.norm())
        if name.startswith(("outc.", "up4.")):
            err = float((g - r).abs().max()) / float(r.abs().max())
            print(f"  {name}: max error {err:.4f} of the largest entry, norm {gn:.4e} vs {rn:.4e}")
            assert err <= (0.02 if name.startswith("outc.") else 0.03) and abs(gn - rn) <= 0.03 * rn, name
        else:
            assert abs(gn - rn) <= 0.25 * rn, (name, gn, rn)


@pytest.mark.parametrize("shape", [(2, 32, 32), (1, 48, 80), (3, 16, 64)])
def test_training_step_small_and_odd_shapes(shape):
    """Shapes whose deepest levels are smaller than one tile of the kernels (2x2, 3x5, 1x4 pixels at level 4): the TMA boxes of
    the convolutions and of the weight gradient hang over the tensors on every side.  With so few samples per BatchNorm
    channel the gradients themselves are rounding-chaotic, so what is asserted is the forward (logits, loss against the
    oracle's train-mode restatement in fp32) and that every gradient is finite and of the right order of magnitude."""
    from oracle import unet_port
    from pkg.segmentation_model import UNet
    B, H, W = shape
    sd = synth.make_state_dict(21, planted=False)
    model = UNet(3, 1)
    model.load_state_dict(sd)
    model = model.cuda().train()
    g = torch.Generator().manual_seed(B * 1000 + H + W)
    x = torch.rand((B, 3, H, W), generator=g).cuda()
    t = (torch.rand((B, 1, H, W), generator=g) > 0.6).float().cuda()
    crit = torch.nn.BCEWithLogitsLoss()
    out = model(x)
    loss = crit(out, t)
    loss.backward()
    ref_sd = {k: v.detach().clone().cuda().float().requires_grad_(v.dtype.is_floating_point and "running" not in k)
              for k, v in sd.items()}
    ref_out = unet_port.unet_forward_train(ref_sd, x)
    ref_loss = crit(ref_out, t)
    ref_loss.backward()
    scale = float(ref_out.detach().abs().max())
    err = float((out.detach() - ref_out.detach()).abs().max())
    print(f"{shape}: loss {loss.item():.5f} vs {ref_loss.item():.5f}, max |logit - ref| {err:.4f} of {scale:.3f}")
    assert abs(loss.item() - ref_loss.item()) <= 1e-2 * abs(ref_loss.item())
    assert err <= 0.1 * scale
    for name, prm in model.named_parameters():
        gn, rn = float(prm.grad.float().norm()), float(ref_sd[name].grad.norm())
        assert torch.isfinite(prm.grad).all() and 0.2 * rn <= gn <= 5.0 * rn + 1e-6, (name, gn, rn)


def test_training_loop_follows_the_reference_loss_curve():
    """Fifteen Adam steps of the reference's loop (`scripts/train_segmenter.py:151-165`) from the same weights on the same
    batches: the drop-in module against the oracle's train-mode restatement under torch.autograd (fp32).  Single gradient
    entries differ at the rounding level (module docstring), the optimisation does not: the loss stays within 2 % at
    every step (measured: 0.6 % over 60 steps at 4 x 256 x 256 against the reference module, profiles/r02_train_curve.txt)."""
    from oracle import unet_port
    from pkg.segmentation_model import UNet
    sd = synth.make_state_dict(9, planted=False)
    batches = [tuple(torch.from_numpy(a).cuda() for a in train_batch((70 + 2 * i, 71 + 2 * i), 64, 96)) for i in range(3)]
    crit = torch.nn.BCEWithLogitsLoss()
    model = UNet(3, 1)
    model.load_state_dict(sd)
    model = model.cuda().train()
    opt = torch.optim.Adam(model.parameters(), lr=1e-4)
    ref = {k: v.detach().clone().cuda().float().requires_grad_(v.dtype.is_floating_point and "running" not in k) for k, v in sd.items()}
    ref_opt = torch.optim.Adam([v for v in ref.values() if v.requires_grad], lr=1e-4)
    worst = 0.0
    for step in range(15):
        x, t = batches[step % 3]
        opt.zero_grad()
        loss = crit(model(x), t)
        loss.backward()
        opt.step()
        ref_opt.zero_grad()
        ref_loss = crit(unet_port.unet_forward_train(ref, x), t)
        ref_loss.backward()
        ref_opt.step()
        worst = max(worst, abs(loss.item() - ref_loss.item()) / ref_loss.item())
    print(f"15 steps: last loss {loss.item():.5f} vs {ref_loss.item():.5f}, largest relative difference along the way {worst:.2e}")
    assert ref_loss.item() < 0.7 * 0.86 and worst <= 0.02


def test_weight_packing_kernel_matches_torch_layouts():
    """`rdp_unet_train_pack` (one kernel per layer, every step) against the same two layouts built with torch ops
    (`rdp_b200.train.pack_for_training`): bf16 [cout][tap][cin] and the flipped, transposed [cin][tap][cout], bit for bit."""
    import ctypes as C
    from rdp_b200 import _lib, train
    from rdp_b200.synth import unet_layer_table
    lib = _lib.load()
    train._bind(lib)
    sd = {k: v.cuda() for k, v in synth.make_state_dict(5, planted=False).items()}
    ws, wts = train.pack_for_training(sd, 3)
    stream = torch.cuda.current_stream().cuda_stream
    for i, (conv, _, cin, cout) in enumerate(unet_layer_table(3)):
        w32 = sd[f"{conv}.weight"].float().contiguous()
        got_w = torch.full_like(ws[i], float("nan"))
        got_t = torch.full_like(wts[i], float("nan")) if wts[i] is not None else None
        _lib.check(lib.rdp_unet_train_pack(w32.data_ptr(), cout, cin, int(i == 0), got_w.data_ptr(),
                                           got_t.data_ptr() if got_t is not None else None, stream), "rdp_unet_train_pack")
        torch.cuda.synchronize()
        assert torch.equal(got_w.view(torch.int16), ws[i].view(torch.int16)), conv
        if got_t is not None:
            assert torch.equal(got_t.view(torch.int16), wts[i].view(torch.int16)), conv
    assert lib.rdp_unet_train_pack(None, 64, 64, 0, None, None, stream) != 0       # null pointers are refused
    w32 = sd["inc.double_conv.3.weight"].float().contiguous()
    assert lib.rdp_unet_train_pack(w32.data_ptr(), 64, 64, 0, ws[1].data_ptr(), None, stream) != 0   # w_t is required


def test_train_mode_then_eval_mode_round_trip():
    """After a training step the same module serves inference again (the folded weights are rebuilt from the updated
    parameters and running statistics)."""
    from oracle import unet_port
    from pkg.segmentation_model import UNet
    model = UNet(3, 1)
    model.load_state_dict(synth.make_state_dict(8, planted=False))
    model = model.cuda().train()
    x, t = (torch.from_numpy(a).cuda() for a in train_batch((44, 45), 64, 96))
    torch.nn.BCEWithLogitsLoss()(model(x), t).backward()
    model.eval()
    with torch.no_grad():
        out = model(x)
    sd = {k: v.detach().cpu() for k, v in model.state_dict().items()}
    ref = unet_port.unet_forward(sd, x.cpu())
    assert float((out.cpu() - ref).abs().max()) <= 0.05 * float(ref.abs().max())
    # the trained module pickles as a whole (`mlflow.pytorch.log_model`, train_segmenter.py:199-204) without the engine, the
    # lock or the bf16 operand buffers of the training step, and the copy serves inference
    import io
    import pickle
    state = model.__getstate__()
    assert not any(k.startswith("_rdp_") for k in state)
    buf = io.BytesIO()
    torch.save(model, buf)
    assert buf.tell() < 80e6                                   # 17.3 M fp32 parameters + buffers, nothing else
    buf.seek(0)
    copy = torch.load(buf, weights_only=False).cuda().eval()
    with torch.no_grad():
        assert torch.equal(copy(x), out)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs in one process")
def test_training_step_on_a_second_device_of_the_process():
    """One process, two GPUs (SURVEY 8e allows threads instead of torchrun): kernel attributes, the packed operand buffers and
    the workspace belong to the device the batch lives on."""
    from pkg.segmentation_model import UNet
    x, t = (torch.from_numpy(a) for a in train_batch((44, 45), 64, 96))
    losses, norms = [], []
    for d in (0, 1):
        dev = torch.device("cuda", d)
        model = UNet(3, 1)
        model.load_state_dict(synth.make_state_dict(8, planted=False))
        model = model.to(dev).train()
        loss = torch.nn.BCEWithLogitsLoss()(model(x.to(dev)), t.to(dev))
        loss.backward()
        torch.cuda.synchronize(dev)
        losses.append(loss.item())
        norms.append(float(model.outc.conv.weight.grad.norm()))
    assert abs(losses[0] - losses[1]) <= 1e-5 * abs(losses[0]) and abs(norms[0] - norms[1]) <= 1e-3 * norms[0], (losses, norms)
