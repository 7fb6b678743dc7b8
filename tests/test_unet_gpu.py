"""GPU parity of the U-Net kernels (through the C ABI / the drop-in module) against the fp32 oracle."""
import numpy as np
import pytest
import torch

from oracle import unet_port
from rdp_b200 import synth

pytestmark = pytest.mark.gpu

BORDER = 0.05   # |reference logit| below this counts as a borderline-threshold pixel (listed, not hidden)


def _model(sd):
    from pkg.segmentation_model import UNet
    m = UNet(3, 1)
    m.load_state_dict(sd)
    return m.cuda().eval()


def _agreement(mask, ref_logits):
    ref_mask = unet_port.mask_from_logits(ref_logits).numpy()
    diff = mask != ref_mask
    border = np.abs(ref_logits.numpy()) < BORDER
    listed = np.argwhere(diff)
    return 1.0 - diff.mean(), int((diff & ~border).sum()), listed


def test_layer_activations_random_weights():
    """bf16 storage / fp32 accumulate against fp32: every conv output within 3 % of the layer's max."""
    sd = synth.make_state_dict(0, planted=False)
    batch = synth.make_batch((0, 1), 48, 64, 0.3, "u16")
    from rdp_b200.unet import UnetEngine
    eng = UnetEngine(sd, 3, 1, "cuda")
    rgb = torch.from_numpy(batch["rgb"]).cuda()
    logits, mask, ws = eng.forward(rgb, want_logits=True, want_mask=True, keep_activations=True)
    taps = {}
    ref = unet_port.unet_forward(sd, unet_port.rgb_u8_to_input(batch["rgb"]), taps)
    for i, (conv, _, _, _) in enumerate(synth.unet_layer_table(3)):
        got = eng.read_activation(i, 2, 48, 64, ws).cpu()
        want = taps[conv]
        rel = float((got - want).abs().max() / want.abs().max())
        assert rel < 3e-2, (i, conv, rel)
    rel = float((logits.cpu() - ref).abs().max() / ref.abs().max())
    assert rel < 3e-2, rel
    assert torch.equal(mask.cpu(), (logits.cpu() > 8.940697e-8).to(torch.uint8))   # mask rule applied to our own logits


@pytest.mark.parametrize("case", ["unet_planted_48x64", "unet_planted_odd_37x53", "unet_planted_96x128"])
def test_mask_agreement_with_reference_vectors(case, manifest, golden):
    meta, g = manifest["cases"][case], golden(case)
    sd = synth.make_state_dict(meta["weight_seed"], planted=meta["planted"])
    batch = synth.make_batch(meta["frame_seeds"], meta["H"], meta["W"], meta["fill"], "u16")
    model = _model(sd)
    mask, logits, _ = model.forward_mask(torch.from_numpy(batch["rgb"]).cuda(), want_logits=True)
    agree, hard_misses, listed = _agreement(mask.cpu().numpy(), torch.from_numpy(g["logits"]))
    print(f"{case}: agreement {agree:.5f}, mismatches outside |logit|<{BORDER}: {hard_misses}, listed {listed[:8].tolist()}")
    assert agree >= 0.999 and hard_misses == 0
    assert float((logits.cpu() - torch.from_numpy(g["logits"])).abs().max()) < 0.15


def test_module_interface_matches_fast_path():
    """`model(x)` with the fp32 NCHW tensor the server builds returns the logits the uint8 fast path computes."""
    sd = synth.make_state_dict(1, planted=True)
    batch = synth.make_batch((7,), 64, 80, 0.25, "u16")
    model = _model(sd)
    x = unet_port.rgb_u8_to_input(batch["rgb"]).cuda()
    out = model(x)
    assert out.shape == (1, 1, 64, 80) and out.dtype == torch.float32 and out.is_cuda
    mask, logits, _ = model.forward_mask(torch.from_numpy(batch["rgb"]).cuda(), want_logits=True)
    assert torch.equal(out, logits)
    assert torch.equal((torch.sigmoid(out) > 0.5).to(torch.uint8), mask)          # server.py:124 on our logits
    with pytest.raises(RuntimeError):
        model.train()(x)
    with pytest.raises(RuntimeError):
        model.eval()(x.cpu())


def test_reload_state_dict_repacks():
    model = _model(synth.make_state_dict(1, planted=True))
    batch = synth.make_batch((8,), 48, 64, 0.3, "u16")
    rgb = torch.from_numpy(batch["rgb"]).cuda()
    a, _, _ = model.forward_mask(rgb, want_logits=True)
    model.load_state_dict(synth.make_state_dict(3, planted=True))
    _, lb, _ = model.forward_mask(rgb, want_logits=True)
    ref = unet_port.unet_forward(synth.make_state_dict(3, planted=True), unet_port.rgb_u8_to_input(batch["rgb"]))
    assert float((lb.cpu() - ref).abs().max()) < 0.15


def test_full_resolution_frame():
    """One 640x480 frame (BASELINE config C1/C2 resolution) against the fp32 oracle run on the host."""
    sd = synth.make_state_dict(2, planted=True)
    batch = synth.make_batch((11,), 480, 640, 0.15, "u16")
    model = _model(sd)
    mask, logits, _ = model.forward_mask(torch.from_numpy(batch["rgb"]).cuda(), want_logits=True)
    ref = unet_port.unet_forward(sd, unet_port.rgb_u8_to_input(batch["rgb"]))
    agree, hard_misses, listed = _agreement(mask.cpu().numpy(), ref)
    print(f"640x480: agreement {agree:.6f}, hard misses {hard_misses}, borderline listed {len(listed)}")
    assert agree >= 0.999 and hard_misses == 0


def test_batch_is_frame_independent():
    """B=6 at 96x128: each frame's mask equals the mask of that frame run alone (tiles never mix frames)."""
    sd = synth.make_state_dict(2, planted=True)
    batch = synth.make_batch(range(20, 26), 96, 128, 0.2, "u16")
    model = _model(sd)
    rgb = torch.from_numpy(batch["rgb"]).cuda()
    _, all_logits, _ = model.forward_mask(rgb, want_logits=True)
    for i in (0, 3, 5):
        _, one, _ = model.forward_mask(rgb[i:i + 1], want_logits=True)
        assert torch.equal(one[0], all_logits[i])


def test_hd_frame_720p():
    """One 1280x720 frame (BASELINE config C3 resolution; H/16 = 45 is odd, so the Up pad path is exercised)."""
    sd = synth.make_state_dict(2, planted=True)
    batch = synth.make_batch((12,), 720, 1280, 0.15, "u16")
    model = _model(sd)
    mask, logits, _ = model.forward_mask(torch.from_numpy(batch["rgb"]).cuda(), want_logits=True)
    ref = unet_port.unet_forward(sd, unet_port.rgb_u8_to_input(batch["rgb"]))
    agree, hard_misses, listed = _agreement(mask.cpu().numpy(), ref)
    print(f"1280x720: agreement {agree:.6f}, hard misses {hard_misses}, borderline listed {len(listed)}")
    assert agree >= 0.999 and hard_misses == 0


@pytest.mark.parametrize("n_channels,n_classes", [(1, 2), (4, 1), (5, 3)])
def test_other_channel_and_class_counts(n_channels, n_classes):
    """`UNet(n_channels, n_classes)` is general in the reference; 1/3/4 input channels use the fused first-layer
    kernel, other counts the im2col fallback; several classes go through the rolled head loop."""
    from pkg.segmentation_model import UNet
    sd = synth.make_state_dict(4, n_channels=n_channels, n_classes=n_classes)
    g = torch.Generator().manual_seed(99)
    x = torch.rand(2, n_channels, 48, 64, generator=g)
    model = UNet(n_channels, n_classes)
    model.load_state_dict(sd)
    model = model.cuda().eval()
    out = model(x.cuda())
    ref = unet_port.unet_forward(sd, x)
    assert out.shape == ref.shape == (2, n_classes, 48, 64)
    rel = float((out.cpu() - ref).abs().max() / ref.abs().max())
    assert rel < 3e-2, rel
    if n_channels == 4:   # uint8 fast path with 4 channels as well
        u8 = (x * 255).round().to(torch.uint8).permute(0, 2, 3, 1).contiguous().cuda()
        mask, logits, _ = model.forward_mask(u8, want_logits=True)
        ref8 = unet_port.unet_forward(sd, u8.cpu().permute(0, 3, 1, 2).float() / 255.0)
        assert float((logits.cpu() - ref8).abs().max() / ref8.abs().max()) < 3e-2
        assert mask.shape == (2, 1, 48, 64)
