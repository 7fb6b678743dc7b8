"""ORACLE (test infrastructure, not product code): the server's per-frame pre/post-processing around the hot path.

Restates `services/vision_analysis/server.py:107-110,120-125,133` of the reference (the module itself does not
import here: it needs mlflow) with the very library calls it makes - torchvision `ToTensor` + `Resize(antialias=True)`,
`cv2.resize(..., INTER_NEAREST)`, `np.count_nonzero` - so the CUDA versions can be checked against them:

* `preprocess`      <- `self.preprocess(image_rgb).unsqueeze(0)`               (`server.py:107-110,121`)
* `upsize_mask`     <- `cv2.resize(mask_resized, (W, H), INTER_NEAREST)`       (`server.py:125`)
* `coverage_percent`<- `np.count_nonzero(final_mask) / final_mask.size * 100`  (`server.py:133`)

Library versions in this image: torchvision 0.26.0, OpenCV 4.13.0 (both unpinned by the reference).
"""
from __future__ import annotations

import numpy as np

MODEL_IMG_SIZE = 256  # server.py:58


def preprocess(image_rgb_u8: np.ndarray, size: int = MODEL_IMG_SIZE):
    """(H,W,3) uint8 RGB -> (1,3,size,size) float32."""
    from torchvision import transforms
    t = transforms.Compose([transforms.ToTensor(), transforms.Resize((size, size), antialias=True)])
    return t(image_rgb_u8).unsqueeze(0)


def upsize_mask(mask_small: np.ndarray, H: int, W: int) -> np.ndarray:
    """(h,w) or (h,w,1) uint8 -> (H,W) uint8, nearest neighbour with OpenCV's index rule."""
    import cv2
    return cv2.resize(mask_small, (W, H), interpolation=cv2.INTER_NEAREST)


def coverage_percent(mask: np.ndarray) -> float:
    return (np.count_nonzero(mask) / mask.size) * 100


# ---- restatements of the two library algorithms (what the CUDA kernels implement), pinned to the libraries by
# ---- tests/test_oracle_service.py and by tests/golden/service_*.npz -----------------------------------------

def aa_weights(in_size: int, out_size: int):
    """ATen `_upsample_bilinear2d_aa` per-axis window start / weights (UpSampleKernel.cpp, `_compute_indices_weights_aa`
    with the triangle filter): float32 arithmetic, weights normalised per output index."""
    f = np.float32
    scale = f(in_size) / f(out_size)
    support = scale if scale >= 1 else f(1)
    invscale = f(1) / scale if scale >= 1 else f(1)
    starts, weights = [], []
    for i in range(out_size):
        center = scale * (f(i) + f(0.5))
        lo = max(0, int(center - support + f(0.5)))
        n = min(in_size, int(center + support + f(0.5))) - lo
        x = (np.arange(n, dtype=f) + f(lo) - center + f(0.5)) * invscale
        w = np.maximum(f(0), f(1) - np.abs(x)).astype(f)
        starts.append(lo)
        weights.append(w / w.sum(dtype=f))
    return starts, weights


def preprocess_np(image_u8: np.ndarray, size=MODEL_IMG_SIZE) -> np.ndarray:
    """(H,W,C) uint8 -> (1,C,oh,ow) float32 without torchvision: uint8/255, horizontal pass, then vertical pass."""
    oh, ow = (size, size) if isinstance(size, int) else size
    H, W, C = image_u8.shape
    x = (image_u8.astype(np.float32) / np.float32(255)).transpose(2, 0, 1)          # ToTensor
    sx, wx = aa_weights(W, ow)
    sy, wy = aa_weights(H, oh)
    tmp = np.empty((C, H, ow), np.float32)
    for j in range(ow):
        tmp[:, :, j] = (x[:, :, sx[j]:sx[j] + len(wx[j])] * wx[j]).sum(axis=2, dtype=np.float32)
    out = np.empty((C, oh, ow), np.float32)
    for i in range(oh):
        out[:, i, :] = (tmp[:, sy[i]:sy[i] + len(wy[i]), :] * wy[i][None, :, None]).sum(axis=1, dtype=np.float32)
    return out[None]


def upsize_mask_np(mask_small: np.ndarray, H: int, W: int) -> np.ndarray:
    """OpenCV `resizeNN` index rule (imgproc/src/resize.cpp): floor(dst * (1 / (dst_size / src_size))), clamped."""
    h, w = mask_small.shape[:2]
    ify, ifx = 1.0 / (H / h), 1.0 / (W / w)
    ys = np.minimum(np.floor(np.arange(H) * ify).astype(np.int64), h - 1)
    xs = np.minimum(np.floor(np.arange(W) * ifx).astype(np.int64), w - 1)
    return mask_small.reshape(h, w)[ys][:, xs]


def encode_mask_png(final_mask: np.ndarray) -> bytes:
    """`cv2.imencode('.png', final_mask * 255)` + `.tobytes()` (`server.py:142-143`)."""
    import cv2
    ok, buf = cv2.imencode(".png", final_mask * 255)
    assert ok
    return buf.tobytes()


def decode_mask_png(data: bytes) -> np.ndarray:
    """What the client does with `response.mask` (`client.py:111-112`)."""
    import cv2
    return cv2.imdecode(np.frombuffer(data, dtype=np.uint8), cv2.IMREAD_GRAYSCALE)
