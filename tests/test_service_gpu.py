"""CUDA pre/post-processing (rdp_pre_resize, rdp_post_mask_nearest) and the whole server loop body against the oracle
restatement of `services/vision_analysis/server.py:107-141`, through the C ABI."""
import hashlib
import json
import os

import numpy as np
import pytest
import torch

from oracle import geometry_port as gp
from oracle import service_port as sp
from oracle import unet_port
from oracle.make_golden_service import sample_positions, service_inputs
from rdp_b200 import _lib, service, synth

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
RESIZE_ATOL = 1e-6     # fp32, values in [0,1]; the kernel sums the taps in a different order from ATen (6e-7 seen over 150 random shapes)


def _cases():
    with open(os.path.join(GOLDEN, "service_manifest.json")) as fh:
        return sorted(json.load(fh)["cases"].items())


@pytest.mark.parametrize("name,meta", _cases())
def test_resize_matches_torchvision_golden_and_oracle(name, meta, golden):
    g = golden(name)
    img, _ = service_inputs(meta["seed"], meta["H"], meta["W"], meta["size"])
    x = service.preprocess(torch.from_numpy(img)[None].cuda(), meta["size"]).cpu().numpy()
    assert np.abs(x.reshape(-1)[sample_positions(x.size)] - g["resize_val"]).max() <= RESIZE_ATOL
    assert np.abs(x - sp.preprocess_np(img, meta["size"])).max() <= RESIZE_ATOL


@pytest.mark.parametrize("name,meta", _cases())
def test_nearest_upsize_and_coverage_are_exact(name, meta, golden):
    g = golden(name)
    _, m = service_inputs(meta["seed"], meta["H"], meta["W"], meta["size"])
    up, cnt = service.upsize_mask(torch.from_numpy(m)[None].cuda(), meta["H"], meta["W"])
    up = up.cpu().numpy()[0]
    digest = np.frombuffer(hashlib.sha256(np.ascontiguousarray(up).tobytes()).digest(), np.uint8)
    assert np.array_equal(digest, g["up_sha256"])
    assert int(cnt[0]) == int(g["up_nonzero"])
    assert np.array_equal(up, sp.upsize_mask_np(m, meta["H"], meta["W"]))


def test_batched_bgr_other_sizes_and_repeat_calls():
    rng = np.random.default_rng(8)
    img = rng.integers(0, 256, (5, 120, 200, 3), dtype=np.uint8)
    d = torch.from_numpy(img).cuda()
    out = service.preprocess(d, (48, 80), bgr=True).cpu().numpy()
    for i in range(5):
        assert np.abs(out[i] - sp.preprocess_np(img[i][:, :, ::-1], (48, 80))[0]).max() <= RESIZE_ATOL
    gray = service.preprocess(d[..., :1].contiguous(), 64).cpu().numpy()
    assert np.abs(gray[2] - sp.preprocess_np(img[2][:, :, :1], 64)[0]).max() <= RESIZE_ATOL
    m = (rng.random((5, 48, 80)) > 0.6).astype(np.uint8)
    dm = torch.from_numpy(m).cuda()
    for _ in range(2):                     # the counts are cleared by every call
        up, cnt = service.upsize_mask(dm, 120, 200)
        for i in range(5):
            ref = sp.upsize_mask_np(m[i], 120, 200)
            assert np.array_equal(up[i].cpu().numpy(), ref) and int(cnt[i]) == np.count_nonzero(ref)
    down, cnt = service.upsize_mask(dm, 20, 33)      # INTER_NEAREST also shrinks
    assert np.array_equal(down[1].cpu().numpy(), sp.upsize_mask_np(m[1], 20, 33))
    empty, cnt = service.upsize_mask(torch.zeros((2, 16, 16), dtype=torch.uint8, device="cuda"), 40, 40)
    assert int(cnt.sum()) == 0 and int(empty.sum()) == 0


def test_bad_arguments_are_refused():
    with pytest.raises(_lib.RdpError):
        service.preprocess(torch.zeros((1, 8, 8, 3), dtype=torch.uint8), 4)                    # host tensor
    with pytest.raises(_lib.RdpError):
        service.preprocess(torch.zeros((1, 8, 8, 9), dtype=torch.uint8, device="cuda"), 4)     # 9 channels
    with pytest.raises(_lib.RdpError):
        service.preprocess(torch.zeros((1, 8, 8, 1), dtype=torch.uint8, device="cuda"), 4, bgr=True)
    with pytest.raises(_lib.RdpError):
        service.upsize_mask(torch.zeros((1, 8, 8), dtype=torch.float32, device="cuda"), 16, 16)


def test_server_loop_body_matches_the_oracle_end_to_end():
    """decode -> [resize -> U-Net -> threshold -> nearest up-size -> curvature, coverage] as server.py:120-141 does it."""
    from pkg.segmentation_model import UNet
    H, W, B = 480, 640, 3
    sd = synth.make_state_dict(5, planted=True)
    model = UNet(3, 1).eval()
    model.load_state_dict(sd)
    model = model.cuda()
    frames = [synth.make_frame(200 + i, H, W, 0.30, "u16") for i in range(B)]
    rgb = np.stack([f["rgb"] for f in frames])
    depth = np.stack([f["depth"] for f in frames])
    K, scale = frames[0]["K"], frames[0]["depth_scale"]
    res = service.analyze_frames(model, torch.from_numpy(rgb).cuda(), torch.from_numpy(depth).cuda(), K, scale)
    for i in range(B):
        x = sp.preprocess_np(rgb[i])
        assert np.abs(res.model_input[i].cpu().numpy() - x[0]).max() <= RESIZE_ATOL
        # the U-Net leg has its own parity tests; here the oracle continues from the device's small mask so that a
        # borderline logit (bf16 vs fp32) cannot hide a post-processing difference ...
        small = res.mask_small[i].cpu().numpy()
        final = sp.upsize_mask_np(small, H, W)
        assert np.array_equal(res.final_mask[i].cpu().numpy(), final)
        assert float(res.mask_coverage_percent[i]) == sp.coverage_percent(final)
        o = gp.curvature_profile(final, depth[i], K, scale)
        assert int(res.curvature.status[i]) == o.status
        if o.status == _lib.FRAME_OK:
            assert np.isclose(float(res.curvature.mean_curvature[i]), o.mean_curvature, rtol=1e-6)
            assert np.isclose(float(res.curvature.max_curvature[i]), o.max_curvature, rtol=1e-6)
        # ... and the oracle's own fp32 mask from the same input agrees with the device's except on borderline pixels
        logits = unet_port.unet_forward(sd, torch.from_numpy(x)).numpy()[0, 0]
        borderline = np.abs(logits) < 0.05
        assert np.array_equal((logits > 0)[~borderline], small.astype(bool)[~borderline])
        assert small.mean() > 0.02          # the planted weights segment something at 256x256


def test_micro_batcher_serves_ten_request_threads_like_the_direct_call():
    """server.py:172 runs 10 worker threads; each gets the same numbers as one direct batched pass (frames are
    independent, so batch composition must not matter)."""
    import threading

    from pkg.segmentation_model import UNet
    from rdp_b200.batching import MicroBatcher
    H, W, n_threads, per_thread = 240, 320, 10, 3
    model = UNet(3, 1).eval()
    model.load_state_dict(synth.make_state_dict(5, planted=True))
    model = model.cuda()
    frames = [synth.make_frame(300 + i, H, W, 0.30, "u16") for i in range(n_threads * per_thread)]
    K, scale = frames[0]["K"], frames[0]["depth_scale"]
    bgr = [np.ascontiguousarray(f["rgb"][:, :, ::-1]) for f in frames]          # what cv2.imdecode hands the server
    direct = service.analyze_frames(model, torch.from_numpy(np.stack(bgr)).cuda(),
                                    torch.from_numpy(np.stack([f["depth"] for f in frames])).cuda(), K, scale, bgr=True)
    got = {}
    with MicroBatcher(model, K, scale, max_batch=16, max_wait_ms=5.0, want_png=True) as mb:
        def worker(t):
            for j in range(per_thread):
                i = t * per_thread + j
                got[i] = mb.analyze(bgr[i], frames[i]["depth"])
        threads = [threading.Thread(target=worker, args=(t,)) for t in range(n_threads)]
        for t in threads:
            t.start()
        for t in threads:
            t.join(120)
        assert mb.frames_run == len(frames) and mb.batches_run < len(frames)
    assert sorted(got) == list(range(len(frames)))
    for i, r in got.items():
        assert np.array_equal(r.final_mask, direct.final_mask[i].cpu().numpy())
        assert r.mask_coverage == float(direct.mask_coverage_percent[i])
        assert np.array_equal(sp.decode_mask_png(r.mask_png), r.final_mask * 255)     # what the client would see
        ok = int(direct.curvature.status[i]) == _lib.FRAME_OK
        assert len(r.curvature_results.spline_points) == (100 if ok else 0)
        if ok:
            assert r.curvature_results.mean_curvature == float(direct.curvature.mean_curvature[i])
            assert r.curvature_results.max_curvature == float(direct.curvature.max_curvature[i])


def _png_chunks(data):
    import struct
    import zlib
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    pos, out = 8, []
    while pos < len(data):
        n, kind = struct.unpack(">I4s", data[pos:pos + 8])
        body = data[pos + 8:pos + 8 + n]
        assert zlib.crc32(kind + body) == struct.unpack(">I", data[pos + 8 + n:pos + 12 + n])[0], kind
        out.append((kind, body))
        pos += 12 + n
    return out


def test_mask_png_from_the_gpu_decodes_to_the_reference_picture():
    """server.py:142-143 encodes with OpenCV, client.py:111-112 decodes with OpenCV: the decoded picture is the contract."""
    import zlib
    rng = np.random.default_rng(4)
    yy, xx = np.mgrid[0:480, 0:640]
    band = ((np.sin(xx / 90.0) * 60 + 200 < yy) & (yy < np.sin(xx / 90.0) * 60 + 330)).astype(np.uint8)
    masks = [band, np.zeros((480, 640), np.uint8), np.ones((480, 640), np.uint8), (rng.random((480, 640)) > 0.5).astype(np.uint8),
             (rng.random((480, 640)) > 0.999).astype(np.uint8), band * 7]                 # any non-zero value is "set"
    pngs = service.encode_mask_png(torch.from_numpy(np.stack(masks)).cuda())
    for m, data in zip(masks, pngs):
        chunks = _png_chunks(data)                                                        # CRC-32 of every chunk
        assert [k for k, _ in chunks] == [b"IHDR", b"IDAT", b"IEND"]
        raw = zlib.decompress(chunks[1][1])                                               # Adler-32 checked by zlib
        assert len(raw) == 480 * 81
        want = (m != 0).astype(np.uint8) * 255
        assert np.array_equal(sp.decode_mask_png(data), want)
        assert np.array_equal(sp.decode_mask_png(sp.encode_mask_png((m != 0).astype(np.uint8))), want)   # the reference's round trip
    assert len(pngs[0]) < 4 * len(sp.encode_mask_png(band))                               # run-length coding is effective
    print("png bytes:", [len(p) for p in pngs], "cv2:", [len(sp.encode_mask_png((m != 0).astype(np.uint8))) for m in masks])


@pytest.mark.parametrize("H,W,B", [(1, 1, 1), (3, 5, 2), (97, 131, 3), (33, 2600, 2), (720, 1280, 2), (700, 9, 1)])
def test_mask_png_other_shapes(H, W, B):
    rng = np.random.default_rng(H * W)
    m = (rng.random((B, H, W)) > 0.7).astype(np.uint8)
    m[0, :, : W // 2] = 1
    for _ in range(2):                                                                    # buffers are reused across calls
        pngs = service.encode_mask_png(torch.from_numpy(m).cuda())
        for i in range(B):
            _png_chunks(pngs[i])
            assert np.array_equal(sp.decode_mask_png(pngs[i]).reshape(H, W), m[i] * 255)


def test_png_entry_point_refuses_bad_buffers():
    import ctypes as C
    lib = _lib.load()
    B, H, W = 1, 16, 24
    max_png, ws_bytes = C.c_size_t(), C.c_size_t()
    _lib.check(lib.rdp_post_mask_png_sizes(B, H, W, C.byref(max_png), C.byref(ws_bytes)), "sizes")
    assert max_png.value >= 43 + 20 and ws_bytes.value > 0
    m = torch.zeros((B, H, W), dtype=torch.uint8, device="cuda")
    png = torch.empty((B, max_png.value + 8), dtype=torch.uint8, device="cuda")
    n = torch.empty(B, dtype=torch.int32, device="cuda")
    ws = torch.empty(ws_bytes.value, dtype=torch.uint8, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    ok_stride = (max_png.value + 3) // 4 * 4
    assert lib.rdp_post_mask_png(m.data_ptr(), B, H, W, png.data_ptr(), 32, n.data_ptr(), ws.data_ptr(), ws.numel(), stream) == -1
    assert b"png_stride" in lib.rdp_last_error()
    assert lib.rdp_post_mask_png(m.data_ptr(), B, H, W, png.data_ptr(), ok_stride, n.data_ptr(), ws.data_ptr(), 16, stream) == -4
    assert lib.rdp_post_mask_png(m.data_ptr(), B, H, W, None, ok_stride, n.data_ptr(), ws.data_ptr(), ws.numel(), stream) == -1
    assert lib.rdp_post_mask_png_sizes(0, H, W, C.byref(max_png), C.byref(ws_bytes)) == -1
    assert lib.rdp_pre_resize(m.data_ptr(), 1, 4, 4, 9, 2, 2, 0, png.data_ptr(), stream) == -2           # 9 channels
    assert lib.rdp_post_mask_nearest(None, 1, 4, 4, 8, 8, png.data_ptr(), None, stream) == -1
    assert lib.rdp_post_mask_png(m.data_ptr(), B, H, W, png.data_ptr(), ok_stride, n.data_ptr(), ws.data_ptr(), ws.numel(), stream) == 0
    torch.cuda.synchronize()
    assert np.array_equal(sp.decode_mask_png(png[0, :int(n[0])].cpu().numpy().tobytes()), np.zeros((H, W), np.uint8))


def test_stream_ring_gives_the_same_results_as_one_stream():
    from pkg.segmentation_model import UNet
    H, W, B = 240, 320, 6
    model = UNet(3, 1).eval()
    model.load_state_dict(synth.make_state_dict(5, planted=True))
    model = model.cuda()
    batches = []
    for k in range(5):
        frames = [synth.make_frame(400 + 10 * k + i, H, W, 0.30, "u16") for i in range(B)]
        batches.append((torch.from_numpy(np.stack([f["rgb"] for f in frames])).cuda(),
                        torch.from_numpy(np.stack([f["depth"] for f in frames])).cuda(), frames[0]["K"], frames[0]["depth_scale"]))
    single = [service.analyze_frames(model, *b) for b in batches]
    torch.cuda.synchronize()
    ring = service.StreamRing(3)
    ringed = []
    for b in batches:
        with ring.next():
            ringed.append(service.analyze_frames(model, *b))
    ring.synchronize()
    for a, c in zip(single, ringed):
        assert torch.equal(a.final_mask, c.final_mask) and torch.equal(a.mask_coverage_percent, c.mask_coverage_percent)
        assert torch.equal(a.curvature.status, c.curvature.status) and torch.equal(a.curvature.n_knots, c.curvature.n_knots)
        assert torch.equal(a.curvature.mean_curvature, c.curvature.mean_curvature)
        assert torch.equal(a.curvature.spline_points, c.curvature.spline_points)
