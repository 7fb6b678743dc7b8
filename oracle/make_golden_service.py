"""Generates tests/golden/service_*.npz: outputs of the very library calls the reference server makes around the hot
path (torchvision ToTensor+Resize(antialias=True), cv2.resize INTER_NEAREST, np.count_nonzero;
`services/vision_analysis/server.py:107-110,120-125,133`), run in this container on seeded inputs.

    python oracle/make_golden_service.py

Inputs are regenerated from the seed by `service_inputs`; the outputs are stored as 4096 sampled positions plus float64
sums (resize) and a sha256 of the bytes plus the non-zero count (nearest up-size).  Test infrastructure only.
"""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import service_port as sp  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
# name, seed, H, W, model size
CASES = [
    ("service_480x640", 11, 480, 640, 256),      # the camera format of BASELINE.json configs[0..1]
    ("service_720x1280", 12, 720, 1280, 256),    # configs[2]
    ("service_odd_301x517", 13, 301, 517, 256),
    ("service_upscale_100x130", 14, 100, 130, 256),
    ("service_identity_256x256", 15, 256, 256, 256),
]


def service_inputs(seed, H, W, size):
    rng = np.random.default_rng(seed)
    img = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
    # smooth blobs + noise so that the small mask has both large regions and isolated pixels
    yy, xx = np.mgrid[0:size, 0:size]
    m = ((np.sin(yy / 17.0 + seed) * np.cos(xx / 23.0) > 0.2) ^ (rng.random((size, size)) > 0.97)).astype(np.uint8)
    return img, m


def sample_positions(n_total, n=4096, seed=5):
    return np.sort(np.random.default_rng(seed).choice(n_total, size=min(n, n_total), replace=False))


def main():
    import cv2
    import torch
    import torchvision
    manifest = {"torch": torch.__version__, "torchvision": torchvision.__version__, "opencv": cv2.__version__, "cases": {}}
    for name, seed, H, W, size in CASES:
        img, m = service_inputs(seed, H, W, size)
        x = sp.preprocess(img, size).numpy()
        pos = sample_positions(x.size)
        up = sp.upsize_mask(m, H, W)
        np.savez_compressed(os.path.join(OUT, name + ".npz"), resize_pos=pos, resize_val=x.reshape(-1)[pos],
                            resize_sum=np.float64(x.sum(dtype=np.float64)),
                            resize_abs_sum=np.float64((x.astype(np.float64) ** 2).sum()),
                            up_sha256=np.frombuffer(hashlib.sha256(np.ascontiguousarray(up).tobytes()).digest(), np.uint8),
                            up_nonzero=np.int64(np.count_nonzero(up)), coverage=np.float64(sp.coverage_percent(up)))
        manifest["cases"][name] = dict(seed=seed, H=H, W=W, size=size, coverage=float(sp.coverage_percent(up)))
        print(name, manifest["cases"][name])
    with open(os.path.join(OUT, "service_manifest.json"), "w") as fh:
        json.dump(manifest, fh, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
