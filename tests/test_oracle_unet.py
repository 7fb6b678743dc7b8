"""The fp32 U-Net port in oracle/ reproduces what the unmodified reference produced (tests/golden)."""
import numpy as np
import pytest
import torch

from oracle import unet_port
from rdp_b200 import synth


def _cases(manifest):
    return [(k, v) for k, v in manifest["cases"].items() if v["kind"] == "unet"]


def test_unet_port_matches_reference_vectors(manifest, golden):
    assert _cases(manifest), "no U-Net golden cases"
    for name, meta in _cases(manifest):
        g = golden(name)
        sd = synth.make_state_dict(meta["weight_seed"], planted=meta["planted"])
        batch = synth.make_batch(meta["frame_seeds"], meta["H"], meta["W"], meta["fill"], "u16")
        taps = {}
        logits = unet_port.unet_forward(sd, unet_port.rgb_u8_to_input(batch["rgb"]), taps)
        # same library, same ops: expect (near) bit equality; allow a few ulp for thread-count dependent oneDNN paths
        np.testing.assert_allclose(logits.numpy(), g["logits"], rtol=0, atol=2e-5, err_msg=name)
        mask = unet_port.mask_from_logits(logits).numpy()
        assert (mask != g["mask"]).mean() <= 1e-4, name
        for key, val in taps.items():
            v = val.numpy()
            assert abs(v.mean(dtype=np.float64) - g[f"tap_mean/{key}"]) <= 1e-5, (name, key)
            shape = v.shape
            rng = np.random.default_rng(7)
            pos = tuple(rng.integers(0, s, size=48) for s in shape)
            np.testing.assert_allclose(v[pos], g[f"tap_samples/{key}"], rtol=0, atol=2e-5, err_msg=f"{name}:{key}")


def test_state_dict_has_reference_keys():
    """Key set a reference checkpoint has - the drop-in module must accept exactly these."""
    sd = synth.make_state_dict(0)
    keys = set(sd)
    assert "inc.double_conv.0.weight" in keys and "inc.double_conv.4.running_var" in keys
    assert "down4.maxpool_conv.1.double_conv.3.weight" in keys
    assert "up1.conv.double_conv.0.weight" in keys and "outc.conv.bias" in keys
    assert sd["up1.conv.double_conv.0.weight"].shape == (512, 1024, 3, 3)
    assert sd["up4.conv.double_conv.3.weight"].shape == (64, 64, 3, 3)
    n_params = sum(v.numel() for k, v in sd.items() if not k.endswith(("running_mean", "running_var", "num_batches_tracked")))
    assert n_params == 17_262_977      # SURVEY 2, component 1


def test_threshold_rule_is_sigmoid_gt_half():
    x = torch.tensor([[-1.0, 0.0, 5e-8, 2e-7, 3.0]])
    assert unet_port.mask_from_logits(x).tolist() == [[0, 0, 0, 1, 1]]


def test_planted_weights_give_bimodal_logits(manifest, golden):
    name = "unet_planted_96x128"
    g = golden(name)
    lg = g["logits"]
    assert (np.abs(lg) < 0.05).mean() < 1e-3
    assert 0.1 < manifest["cases"][name]["coverage"] < 0.4


def test_train_mode_port_matches_reference_autograd():
    """`unet_port.unet_forward_train` (batch-statistics BatchNorm, under autograd) against what the unmodified reference
    module computed in `train()` mode: loss and sampled gradients of tests/golden/train_step_2x64x96.npz
    (`oracle/make_golden_train.py`).  Pins the checker of the multi-class training test."""
    import json
    import os
    from oracle.make_golden_train import train_batch
    gdir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    with open(os.path.join(gdir, "train_manifest.json")) as fh:
        meta = json.load(fh)["cases"]["train_step_2x64x96"]
    with np.load(os.path.join(gdir, "train_step_2x64x96.npz")) as z:
        g = {k: z[k] for k in z.files}
    sd = {k: (v.clone().requires_grad_(True) if v.dtype.is_floating_point and "running" not in k else v)
          for k, v in synth.make_state_dict(meta["weight_seed"], planted=False).items()}
    x, t = (torch.from_numpy(a) for a in train_batch(meta["frame_seeds"], meta["H"], meta["W"]))
    loss = torch.nn.BCEWithLogitsLoss()(unet_port.unet_forward_train(sd, x), t)
    loss.backward()
    assert abs(loss.item() - float(g["loss"])) <= 1e-5 * float(g["loss"])
    for name in ("inc.double_conv.0.weight", "down4.maxpool_conv.1.double_conv.3.weight", "up1.conv.double_conv.1.bias",
                 "up4.conv.double_conv.3.weight", "outc.conv.weight", "outc.conv.bias"):
        got = sd[name].grad.reshape(-1)[torch.from_numpy(g[f"grad_idx/{name}"])].numpy()
        np.testing.assert_allclose(got, g[f"grad_val/{name}"], rtol=2e-3, atol=2e-4 * float(g[f"grad_absmax/{name}"]), err_msg=name)
