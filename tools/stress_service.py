"""One-off randomised check of the service kernels (resize, nearest up-size + count, PNG) against the oracle on many
shapes, including up-scaling, extreme aspect ratios and sizes that are not multiples of anything."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "robotic-discovery-platform_b200")]
from oracle import service_port as sp  # noqa: E402
from rdp_b200 import service  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 60
    rng = np.random.default_rng(77)
    worst = 0.0
    for it in range(n):
        H, W = int(rng.integers(1, 700)), int(rng.integers(1, 900))
        oh, ow = int(rng.integers(1, 300)), int(rng.integers(1, 300))
        C = int(rng.choice([1, 3, 4]))
        B = int(rng.integers(1, 4))
        img = rng.integers(0, 256, (B, H, W, C), dtype=np.uint8)
        out = service.preprocess(torch.from_numpy(img).cuda(), (oh, ow)).cpu().numpy()
        for b in range(B):
            d = float(np.abs(out[b] - sp.preprocess_np(img[b], (oh, ow))[0]).max())
            worst = max(worst, d)
            assert d <= 2e-6, (it, H, W, oh, ow, C, d)
        m = (rng.random((B, oh, ow)) > rng.random()).astype(np.uint8)
        up, cnt = service.upsize_mask(torch.from_numpy(m).cuda(), H, W)
        pngs = service.encode_mask_png(up)
        for b in range(B):
            ref = sp.upsize_mask_np(m[b], H, W)
            assert np.array_equal(up[b].cpu().numpy(), ref) and int(cnt[b]) == int(np.count_nonzero(ref)), (it, H, W, oh, ow)
            dec = sp.decode_mask_png(pngs[b])
            assert dec is not None and np.array_equal(dec.reshape(H, W), ref * 255), (it, H, W)
    print(f"{n} random shapes: resize worst abs diff {worst:.2e}, nearest / count / PNG exact")


if __name__ == "__main__":
    main()
