"""One optimisation step of the reference's training loop (`scripts/train_segmenter.py:151-165`) through the drop-in
`UNet` in train mode, against what the unmodified reference computed under `torch.autograd` in fp32
(tests/golden/train_step_*.npz, `oracle/make_golden_train.py`).

Tolerances (bf16 storage of activations and activation gradients, fp32 accumulation): loss 1e-3 relative; every
parameter gradient within 2 % of that tensor's largest entry; BatchNorm running statistics 1e-4 absolute + 1e-3 relative."""
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
    assert logit_err <= 0.03 * float(g["logits_absmax"])
    worst = ("", 0.0)
    for pname, p in model.named_parameters():
        assert p.grad is not None, pname
        idx = torch.from_numpy(g[f"grad_idx/{pname}"]).cuda()
        got = p.grad.reshape(-1)[idx].cpu().numpy()
        want, scale = g[f"grad_val/{pname}"], float(g[f"grad_absmax/{pname}"])
        err = float(np.abs(got - want).max()) / scale
        norm_rel = abs(float(p.grad.double().norm()) - float(g[f"grad_norm/{pname}"])) / float(g[f"grad_norm/{pname}"])
        if err > worst[1]:
            worst = (pname, err)
        assert err <= 0.02, (pname, err)
        assert norm_rel <= 0.02, (pname, norm_rel)
    print(f"{name}: worst gradient error {worst[1]:.4f} of the tensor's max-abs ({worst[0]})")
    for bname, b in model.named_buffers():
        want = g[f"buffer/{bname}"]
        if bname.endswith("num_batches_tracked"):
            assert int(b) == int(want), bname
        else:
            np.testing.assert_allclose(b.cpu().numpy(), want, rtol=1e-3, atol=1e-4, err_msg=bname)
    opt.step()                                                              # :161, torch's Adam on our gradients
    for pname, p in model.named_parameters():
        idx = torch.from_numpy(g[f"grad_idx/{pname}"]).cuda()
        # Adam's first step moves every entry by about lr * sign(grad): entries whose gradient is far from zero must agree
        want_g = g[f"grad_val/{pname}"]
        sure = np.abs(want_g) > 0.05 * float(g[f"grad_absmax/{pname}"])
        got = p.detach().reshape(-1)[idx].cpu().numpy()
        assert np.allclose(got[sure], g[f"param_after/{pname}"][sure], rtol=0, atol=2e-5), pname


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
    assert float((out.cpu() - ref).abs().max()) <= 0.03 * float(ref.abs().max())
