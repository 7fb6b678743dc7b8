"""The body of the reference server's per-frame loop on the GPU (`services/vision_analysis/server.py:116-141`).

    color (B,H,W,3) uint8 --rdp_pre_resize--> (B,3,S,S) f32 --UNet (mask rule fused)--> (B,S,S) u8
          --rdp_post_mask_nearest--> (B,H,W) u8 + non-zero counts --rdp_geom_curvature(depth)--> curvature, coverage

`decode_color_jpeg` / `decode_depth_png` take the encoded payloads of the requests themselves (`server.py:117-118`); the
gRPC plumbing stays with the caller.  Every array between the payload bytes and the response fields stays in HBM.
No CPU fallback.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch

from . import _lib, geometry

MODEL_IMG_SIZE = 256  # server.py:58
PRE_SWAP_RB = 1


def _need_cuda_u8(t: torch.Tensor, what: str, ndim: int) -> torch.Tensor:
    if not isinstance(t, torch.Tensor) or t.device.type != "cuda":
        raise _lib.RdpError(f"{what} must be a CUDA tensor; there is no CPU path")
    if t.dtype != torch.uint8 or t.dim() != ndim:
        raise _lib.RdpError(f"{what} must be uint8 with {ndim} dimensions (got {t.dtype} {tuple(t.shape)})")
    return t.contiguous()


def preprocess(color_u8: torch.Tensor, size: int | tuple = MODEL_IMG_SIZE, *, bgr: bool = False, out=None) -> torch.Tensor:
    """(B,H,W,C) uint8 -> (B,C,size,size) float32: ToTensor + Resize(antialias=True) (`server.py:107-110,121`);
    `bgr=True` also does the `cv2.cvtColor(BGR2RGB)` of `server.py:120`."""
    x = _need_cuda_u8(color_u8, "color image batch", 4)
    B, H, W, Cn = x.shape
    oh, ow = (size, size) if isinstance(size, int) else size
    if out is None:
        out = torch.empty((B, Cn, oh, ow), dtype=torch.float32, device=x.device)
    elif out.shape != (B, Cn, oh, ow) or out.dtype != torch.float32 or not out.is_contiguous() or out.device != x.device:
        raise _lib.RdpError("preprocess: `out` must be a contiguous float32 (B,C,oh,ow) tensor on the input's device")
    with torch.cuda.device(x.device):
        _lib.check(_lib.load().rdp_pre_resize(x.data_ptr(), B, H, W, Cn, oh, ow, PRE_SWAP_RB if bgr else 0, out.data_ptr(),
                                              torch.cuda.current_stream(x.device).cuda_stream), "rdp_pre_resize")
    return out


def upsize_mask(mask_small: torch.Tensor, H: int, W: int, *, want_count: bool = True):
    """(B,h,w) uint8 -> ((B,H,W) uint8, non-zero count (B,) int32 or None): `cv2.resize(.., INTER_NEAREST)` and the
    count behind `mask_coverage_percent` (`server.py:125,133`)."""
    m = _need_cuda_u8(mask_small, "mask batch", 3)
    B, h, w = m.shape
    out = torch.empty((B, H, W), dtype=torch.uint8, device=m.device)
    cnt = torch.empty(B, dtype=torch.int32, device=m.device) if want_count else None
    with torch.cuda.device(m.device):
        _lib.check(_lib.load().rdp_post_mask_nearest(m.data_ptr(), B, h, w, H, W, out.data_ptr(),
                                                     cnt.data_ptr() if want_count else None,
                                                     torch.cuda.current_stream(m.device).cuda_stream), "rdp_post_mask_nearest")
    return out, cnt


def encode_mask_png(final_mask: torch.Tensor, *, to_host: bool = True):
    """(B,H,W) uint8 -> one PNG per frame (`cv2.imencode('.png', final_mask * 255)`, `server.py:142-143`), encoded on
    the GPU.  Returns a list of `bytes` (to_host=True: two small copies, lengths then payload) or the device buffer
    (B, stride) uint8 with the lengths (B,) int32."""
    import ctypes as C
    m = _need_cuda_u8(final_mask, "mask batch", 3)
    B, H, W = m.shape
    lib = _lib.load()
    max_png, ws_bytes = C.c_size_t(), C.c_size_t()
    _lib.check(lib.rdp_post_mask_png_sizes(B, H, W, C.byref(max_png), C.byref(ws_bytes)), "rdp_post_mask_png_sizes")
    stride = (max_png.value + 3) // 4 * 4
    png = torch.empty((B, stride), dtype=torch.uint8, device=m.device)
    nbytes = torch.empty(B, dtype=torch.int32, device=m.device)
    ws = torch.empty(ws_bytes.value, dtype=torch.uint8, device=m.device)
    with torch.cuda.device(m.device):
        _lib.check(lib.rdp_post_mask_png(m.data_ptr(), B, H, W, png.data_ptr(), stride, nbytes.data_ptr(), ws.data_ptr(),
                                         ws.numel(), torch.cuda.current_stream(m.device).cuda_stream), "rdp_post_mask_png")
    if not to_host:
        return png, nbytes
    n = nbytes.cpu().numpy()
    host = png[:, :int(n.max())].cpu().numpy()
    return [host[i, :n[i]].tobytes() for i in range(B)]


DECODE_STATUS = {1: "not an image", 2: "truncated", 3: "unsupported format", 4: "wrong size", 5: "corrupt stream",
                 6: "length mismatch", 7: "bad filter type"}


def jpeg_size(data) -> tuple:
    """(H, W) from a JPEG's first SOFn segment (host side, a few hundred bytes are looked at)."""
    d = memoryview(data)
    i = 2
    while i + 9 < len(d):
        if d[i] != 0xFF:
            break
        m = d[i + 1]
        if m == 0xFF:
            i += 1
            continue
        if 0xC0 <= m <= 0xCF and m not in (0xC4, 0xC8, 0xCC):
            return (d[i + 5] << 8) | d[i + 6], (d[i + 7] << 8) | d[i + 8]
        i += 2 + ((d[i + 2] << 8) | d[i + 3])
    raise _lib.RdpError("no frame header found in the JPEG payload")


def png_size(data) -> tuple:
    d = memoryview(data)
    if len(d) < 24 or bytes(d[:8]) != b"\x89PNG\r\n\x1a\n":
        raise _lib.RdpError("payload is not a PNG")
    return int.from_bytes(d[20:24], "big"), int.from_bytes(d[16:20], "big")


def _upload_payloads(payloads, device):
    """Encoded images of a batch -> (concatenated bytes, int64 offsets) on the device, through one pinned staging copy."""
    import numpy as np
    sizes = [len(x) for x in payloads]
    offs = np.zeros(len(sizes) + 1, dtype=np.int64)
    np.cumsum(sizes, out=offs[1:])
    host = torch.empty(int(offs[-1]) + 8, dtype=torch.uint8).pin_memory()
    hv = host.numpy()
    for x, o in zip(payloads, offs[:-1]):
        hv[o:o + len(x)] = np.frombuffer(x, dtype=np.uint8)
    return host.to(device, non_blocking=True), torch.from_numpy(offs).to(device, non_blocking=True)


def _decode(payloads, fn_name: str, shape_of, out_shape, out_dtype, which: int, device, check: bool):
    import ctypes as C
    if len(payloads) == 0:
        raise _lib.RdpError("no payloads")
    device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    H, W = shape_of(payloads[0])
    B = len(payloads)
    lib = _lib.load()
    sizes = [C.c_size_t(), C.c_size_t()]
    _lib.check(lib.rdp_pre_decode_sizes(B, H, W, C.byref(sizes[0]), C.byref(sizes[1])), "rdp_pre_decode_sizes")
    with torch.cuda.device(device):
        data, offs = _upload_payloads(payloads, device)
        out = torch.empty((B, H, W) + out_shape, dtype=out_dtype, device=device)
        status = torch.empty(B, dtype=torch.int32, device=device)
        ws = torch.empty(sizes[which].value, dtype=torch.uint8, device=device)
        _lib.check(getattr(lib, fn_name)(data.data_ptr(), offs.data_ptr(), B, H, W, out.data_ptr(), status.data_ptr(),
                                         ws.data_ptr(), ws.numel(), torch.cuda.current_stream(device).cuda_stream), fn_name)
        for t in (data, offs, ws):
            t.record_stream(torch.cuda.current_stream(device))
    if check:
        st = status.cpu().numpy()
        bad = [(i, DECODE_STATUS.get(int(v), str(int(v)))) for i, v in enumerate(st) if v]
        if bad:
            raise _lib.RdpError(f"{fn_name}: {len(bad)} of {B} payloads could not be decoded: {bad[:8]}")
    return out, status


def decode_color_jpeg(payloads, *, device=None, check: bool = True):
    """JPEG payloads of a batch of requests -> ((B,H,W,3) uint8 BGR on the GPU, status (B,) int32): what
    `cv2.imdecode(np.frombuffer(request.color_image.data, np.uint8), cv2.IMREAD_COLOR)` (`server.py:117`) returns,
    bit for bit, for the baseline JPEGs `cv2.imencode('.jpg', ...)` writes.  All images of a batch have one size.
    `check=True` reads the status back (one small synchronising copy) and raises on a failed image."""
    return _decode(payloads, "rdp_pre_decode_jpeg", jpeg_size, (3,), torch.uint8, 0, device, check)


def decode_depth_png(payloads, *, device=None, check: bool = True):
    """16-bit greyscale PNG payloads -> ((B,H,W) uint16 on the GPU, status): `cv2.imdecode(..., cv2.IMREAD_UNCHANGED)`
    of `server.py:118`, bit for bit."""
    return _decode(payloads, "rdp_pre_decode_png16", png_size, (), torch.uint16, 1, device, check)


_SIDE_STREAMS: dict = {}


def _side_stream(dev):
    s = _SIDE_STREAMS.get(dev)
    if s is None:
        s = _SIDE_STREAMS[dev] = torch.cuda.Stream(dev)
    return s


@torch.no_grad()
def analyze_encoded(model, color_payloads, depth_payloads, intrinsics, depth_scale, *, size: int = MODEL_IMG_SIZE):
    """The whole loop body of the server from the request's bytes (`server.py:117-141`): both decodes, BGR->RGB, resize,
    U-Net, mask up-size, curvature - nothing decoded ever visits the host."""
    # the two decoders are one serial bit stream per image each: run them side by side on two streams
    dev = torch.device("cuda", torch.cuda.current_device())
    main = torch.cuda.current_stream(dev)
    side = _side_stream(dev)
    side.wait_stream(main)
    with torch.cuda.stream(side):
        depth, ds = decode_depth_png(depth_payloads, device=dev, check=False)
    color, cs = decode_color_jpeg(color_payloads, device=dev, check=False)
    main.wait_stream(side)
    depth.record_stream(main)
    ds.record_stream(main)
    bad = torch.nonzero((cs != 0) | (ds != 0)).flatten().cpu().tolist()
    if bad:
        raise _lib.RdpError(f"payloads {bad[:8]} could not be decoded (JPEG status {cs.cpu().tolist()}, PNG status {ds.cpu().tolist()})")
    return analyze_frames(model, color, depth, intrinsics, depth_scale, size=size, bgr=True)


@dataclass
class FrameAnalysis:
    curvature: geometry.CurvatureBatch   # mean/max curvature, spline points, status per frame
    final_mask: torch.Tensor             # (B,H,W) uint8, camera resolution
    mask_coverage_percent: torch.Tensor  # (B,) float64
    model_input: torch.Tensor            # (B,3,S,S) float32, what the U-Net saw
    mask_small: torch.Tensor             # (B,S,S) uint8, the U-Net's own resolution


@torch.no_grad()
def analyze_frames(model, color_u8: torch.Tensor, depth: torch.Tensor, intrinsics, depth_scale, *,
                   size: int = MODEL_IMG_SIZE, bgr: bool = False) -> FrameAnalysis:
    """One pass of the server loop body over a batch of decoded frames (`server.py:120-141`).

    `model` is the drop-in `pkg.segmentation_model.UNet` (eval mode, on the frames' device); `depth` is the decoded
    depth image batch (B,H,W) uint16 | float32 at camera resolution."""
    B, H, W, _ = color_u8.shape
    x = preprocess(color_u8, size, bgr=bgr)
    model._check(x, inference_only=True)
    _, mask, _ = model._engine_for(x.device).forward(x, want_logits=False, want_mask=True)   # sigmoid(logit) > 0.5
    if mask.shape[1] != 1:
        raise _lib.RdpError("the server's mask path is defined for n_classes == 1 (server.py:124-125)")
    mask_small = mask[:, 0]
    final_mask, cnt = upsize_mask(mask_small, H, W)
    curv = geometry.curvature_batch(final_mask, depth, intrinsics, depth_scale)
    # count / size * 100 in that order (server.py:133); tensor / tensor, because torch divides by a Python scalar
    # through its reciprocal, which rounds differently
    coverage = cnt.to(torch.float64) / torch.full_like(cnt, H * W, dtype=torch.float64) * 100.0
    return FrameAnalysis(curv, final_mask, coverage, x, mask_small)


class StreamRing:
    """Round-robin CUDA streams for back-to-back batches of the server-shaped step.

    The spline fits of a batch are a latency chain on one CTA per frame (64 frames leave more than half of the SMs idle
    for about half of the step); issued on alternating streams, that tail overlaps the next batch's U-Net:
    640x480, 64 frames: 9.5 -> 7.3 ms per batch with two streams; 32 frames: 7.3 -> 4.0 ms with three.

        ring = StreamRing(2)
        for rgb, depth in batches:
            with ring.next():
                r = analyze_frames(model, rgb, depth, K, scale)     # inputs must already be visible to that stream
        ring.synchronize()
    """

    def __init__(self, n: int = 2, device=None):
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        with torch.cuda.device(self.device):
            self.streams = [torch.cuda.Stream(self.device) for _ in range(max(1, int(n)))]
        self._i = 0

    def next(self):
        s = self.streams[self._i % len(self.streams)]
        self._i += 1
        s.wait_stream(torch.cuda.current_stream(self.device))      # what the caller queued so far (its inputs) comes first
        return torch.cuda.stream(s)

    def synchronize(self) -> None:
        for s in self.streams:
            torch.cuda.current_stream(self.device).wait_stream(s)
            s.synchronize()
