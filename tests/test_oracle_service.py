"""oracle/service_port.py (the NumPy restatements of ATen's antialiased resize and OpenCV's nearest resize) against the
vectors torchvision / OpenCV themselves produced (tests/golden/service_*.npz) and, when importable, the libraries."""
import hashlib
import json
import os

import numpy as np
import pytest

from oracle import service_port as sp
from oracle.make_golden_service import sample_positions, service_inputs

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
RESIZE_ATOL = 5e-7     # a few fp32 ulp of values in [0,1]: summation order inside ATen's vectorised passes


def _cases():
    with open(os.path.join(GOLDEN, "service_manifest.json")) as fh:
        return sorted(json.load(fh)["cases"].items())


@pytest.mark.parametrize("name,meta", _cases())
def test_resize_restatement_matches_torchvision_golden(name, meta, golden):
    g = golden(name)
    img, _ = service_inputs(meta["seed"], meta["H"], meta["W"], meta["size"])
    x = sp.preprocess_np(img, meta["size"])
    assert x.shape == (1, 3, meta["size"], meta["size"]) and x.dtype == np.float32
    pos = sample_positions(x.size)
    assert np.array_equal(pos, g["resize_pos"])
    assert np.abs(x.reshape(-1)[pos] - g["resize_val"]).max() <= RESIZE_ATOL
    assert abs(x.sum(dtype=np.float64) - float(g["resize_sum"])) <= RESIZE_ATOL * x.size * 0.05


@pytest.mark.parametrize("name,meta", _cases())
def test_nearest_restatement_matches_opencv_golden(name, meta, golden):
    g = golden(name)
    _, m = service_inputs(meta["seed"], meta["H"], meta["W"], meta["size"])
    up = sp.upsize_mask_np(m, meta["H"], meta["W"])
    digest = np.frombuffer(hashlib.sha256(np.ascontiguousarray(up).tobytes()).digest(), np.uint8)
    assert np.array_equal(digest, g["up_sha256"])
    assert np.count_nonzero(up) == int(g["up_nonzero"])
    assert sp.coverage_percent(up) == float(g["coverage"]) == meta["coverage"]


def test_restatements_against_live_libraries():
    pytest.importorskip("torchvision")
    pytest.importorskip("cv2")
    rng = np.random.default_rng(3)
    for H, W, size in [(64, 80, 32), (33, 47, 64), (96, 96, 96), (200, 50, (40, 70))]:
        img = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
        from torchvision import transforms
        oh, ow = (size, size) if isinstance(size, int) else size
        ref = transforms.Compose([transforms.ToTensor(), transforms.Resize((oh, ow), antialias=True)])(img).numpy()[None]
        assert np.abs(sp.preprocess_np(img, size) - ref).max() <= RESIZE_ATOL
        m = (rng.random((oh, ow)) > 0.5).astype(np.uint8)
        assert np.array_equal(sp.upsize_mask_np(m, H, W), sp.upsize_mask(m, H, W))
