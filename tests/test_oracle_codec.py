"""oracle/codec_port.py (the CPU restatement of libjpeg-turbo's / zlib+libpng's decode for the reference's wire format)
against what OpenCV returned for the committed files (tests/golden/codec_cases.npz) and against OpenCV run live."""
import json
import os

import numpy as np
import pytest

from oracle import codec_port as cp
from rdp_b200 import synth

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def cases():
    with open(os.path.join(GOLDEN, "codec_manifest.json")) as fh:
        manifest = json.load(fh)
    with np.load(os.path.join(GOLDEN, "codec_cases.npz")) as z:
        data = {k: z[k] for k in z.files}
    return manifest["cases"], data


def test_port_reproduces_opencv_on_the_golden_files(cases):
    meta, data = cases
    assert sum(v["kind"] == "jpeg" for v in meta.values()) >= 10 and sum(v["kind"] == "png" for v in meta.values()) >= 8
    for name, v in meta.items():
        payload, want = data[name + "/bytes"].tobytes(), data[name + "/decoded"]
        got = cp.decode_jpeg_bgr(payload) if v["kind"] == "jpeg" else cp.decode_png_gray(payload)
        assert got.shape == want.shape and got.dtype == want.dtype, name
        assert np.array_equal(got, want), name


def test_port_reproduces_opencv_live_on_a_camera_sized_frame():
    cv2 = pytest.importorskip("cv2")
    f = synth.make_frame(61, 240, 320, 0.2, "u16")
    bgr = np.ascontiguousarray(f["rgb"][:, :, ::-1])
    ok, buf = cv2.imencode(".jpg", bgr)                                           # client.py:65
    assert np.array_equal(cp.decode_jpeg_bgr(buf.tobytes()), cv2.imdecode(buf, cv2.IMREAD_COLOR))   # server.py:117
    ok, buf = cv2.imencode(".png", f["depth"])                                    # client.py:66
    got = cp.decode_png_gray(buf.tobytes())
    assert got.dtype == np.uint16 and np.array_equal(got, cv2.imdecode(buf, cv2.IMREAD_UNCHANGED))  # server.py:118
    assert np.array_equal(got, f["depth"])                                        # PNG is lossless


def test_port_rejects_what_it_does_not_implement(cases):
    meta, data = cases
    with pytest.raises(ValueError):
        cp.decode_jpeg_bgr(b"\xff\xd8\xff\xc2\x00\x0b\x08\x00\x10\x00\x10\x01\x01\x11\x00")     # progressive SOF2
    with pytest.raises(ValueError):
        cp.decode_png_gray(data["jpg_tiny_8x8/bytes"].tobytes())
    broken = bytearray(data["png_client_96x128/bytes"].tobytes())
    broken[len(broken) // 2] ^= 0x55
    with pytest.raises(ValueError):
        cp.decode_png_gray(bytes(broken))
