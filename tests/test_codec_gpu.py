"""GPU decoders of the wire format (`rdp_pre_decode_jpeg`, `rdp_pre_decode_png16`, through the C ABI) against OpenCV's
results: the committed golden files, OpenCV run live on camera-sized frames, and the error paths."""
import json
import os

import numpy as np
import pytest
import torch

from oracle import codec_port as cp
from rdp_b200 import synth

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def cases():
    with open(os.path.join(GOLDEN, "codec_manifest.json")) as fh:
        manifest = json.load(fh)
    with np.load(os.path.join(GOLDEN, "codec_cases.npz")) as z:
        data = {k: z[k] for k in z.files}
    return manifest["cases"], data


def test_golden_files_decode_bit_exactly(cases):
    from rdp_b200 import service
    meta, data = cases
    for name, v in meta.items():
        payload, want = data[name + "/bytes"].tobytes(), data[name + "/decoded"]
        if v["kind"] == "jpeg":
            got, st = service.decode_color_jpeg([payload, payload])       # a batch of two: both must come out the same
        else:
            got, st = service.decode_depth_png([payload, payload])
        got = got.cpu().numpy()
        assert st.cpu().tolist() == [0, 0], name
        want16 = want.astype(np.uint16) if v["kind"] == "png" else want   # 8-bit PNGs are widened
        assert got.shape[1:] == want.shape, name
        assert np.array_equal(got[0], want16) and np.array_equal(got[1], want16), (name, int((got[0] != want16).sum()))


@pytest.mark.parametrize("H,W,B", [(480, 640, 8), (720, 1280, 3), (301, 517, 4)])
def test_camera_frames_match_opencv_live(H, W, B):
    """What the client sends (`client.py:65-67`) for B different frames, decoded as one batch, against the two
    `cv2.imdecode` calls of `server.py:117-118`; differing pixels would be listed, there are none."""
    cv2 = pytest.importorskip("cv2")
    from rdp_b200 import service
    frames = [synth.make_frame(700 + i, H, W, 0.1 + 0.1 * i, "u16") for i in range(B)]
    jp = [cv2.imencode(".jpg", np.ascontiguousarray(f["rgb"][:, :, ::-1]))[1].tobytes() for f in frames]
    pn = [cv2.imencode(".png", f["depth"])[1].tobytes() for f in frames]
    color, cs = service.decode_color_jpeg(jp)
    depth, ds = service.decode_depth_png(pn)
    assert color.shape == (B, H, W, 3) and depth.shape == (B, H, W) and depth.dtype == torch.uint16
    for i in range(B):
        want = cv2.imdecode(np.frombuffer(jp[i], np.uint8), cv2.IMREAD_COLOR)
        got = color[i].cpu().numpy()
        diff = np.argwhere(got != want)
        assert diff.shape[0] == 0, (i, diff[:8].tolist())
        assert np.array_equal(depth[i].cpu().numpy(), frames[i]["depth"])
    print(f"{W}x{H}: {B} JPEGs of {np.mean([len(x) for x in jp]) / 1e3:.0f} KB and PNGs of {np.mean([len(x) for x in pn]) / 1e3:.0f} KB "
          f"(decoded frame: {H * W * 5 / 1e3:.0f} KB) bit-exact")


def test_bad_payloads_are_reported_per_image(cases):
    from rdp_b200 import _lib, service
    meta, data = cases
    good = data["jpg_client_96x128/bytes"].tobytes()
    other_size = data["jpg_odd_37x53/bytes"].tobytes()
    truncated = good[: len(good) // 2]
    not_jpeg = data["png_client_96x128/bytes"].tobytes()
    color, st = service.decode_color_jpeg([good, other_size, truncated, not_jpeg, good], check=False)
    st = st.cpu().tolist()
    assert st[0] == 0 and st[4] == 0 and st[1] == 4 and st[3] == 1 and st[2] != 0
    want = data["jpg_client_96x128/decoded"]
    assert np.array_equal(color[0].cpu().numpy(), want) and np.array_equal(color[4].cpu().numpy(), want)
    assert int(color[1].abs().sum()) == 0 and int(color[3].abs().sum()) == 0
    with pytest.raises(_lib.RdpError):
        service.decode_color_jpeg([good, truncated])
    gp = data["png_client_96x128/bytes"].tobytes()
    broken = bytearray(gp)
    broken[len(broken) // 2] ^= 0x55
    depth, st = service.decode_depth_png([gp, bytes(broken), gp[:200], good], check=False)
    st = st.cpu().tolist()
    assert st[0] == 0 and st[1] != 0 and st[2] != 0 and st[3] == 1
    assert np.array_equal(depth[0].cpu().numpy(), data["png_client_96x128/decoded"])


def test_server_loop_from_encoded_payloads_equals_the_decoded_path():
    """`analyze_encoded` (bytes in) against `analyze_frames` on what OpenCV decodes from the same bytes."""
    cv2 = pytest.importorskip("cv2")
    from pkg.segmentation_model import UNet
    from rdp_b200 import service
    model = UNet(3, 1)
    model.load_state_dict(synth.make_state_dict(2, planted=True))
    model = model.cuda().eval()
    frames = [synth.make_frame(800 + i, 240, 320, 0.3, "u16") for i in range(4)]
    jp = [cv2.imencode(".jpg", np.ascontiguousarray(f["rgb"][:, :, ::-1]))[1].tobytes() for f in frames]
    pn = [cv2.imencode(".png", f["depth"])[1].tobytes() for f in frames]
    a = service.analyze_encoded(model, jp, pn, frames[0]["K"], frames[0]["depth_scale"])
    color = torch.from_numpy(np.stack([cv2.imdecode(np.frombuffer(x, np.uint8), cv2.IMREAD_COLOR) for x in jp])).cuda()
    depth = torch.from_numpy(np.stack([f["depth"] for f in frames])).cuda()
    b = service.analyze_frames(model, color, depth, frames[0]["K"], frames[0]["depth_scale"], bgr=True)
    assert torch.equal(a.final_mask, b.final_mask) and torch.equal(a.curvature.status, b.curvature.status)
    assert torch.equal(a.curvature.mean_curvature, b.curvature.mean_curvature)
    assert torch.equal(a.mask_coverage_percent, b.mask_coverage_percent)
