"""`rdp_geom_deproject_static`: the deprojection entry point that brings its own scratch memory.

Some hosts of this library cannot allocate on the GPU once they are running - a serving process whose memory pool
owns the whole device, a control loop with no room for an allocation's latency or failure - so they cannot give
`rdp_geom_curvature` the workspace it asks for.  This entry point keeps the kernel's bookkeeping in a fixed array
inside the library instead.  The test calls it through the C ABI, twice in a row on different frames (the array
keeps the previous call's contents, so a missing clear would show on the second call), and checks points, pixel
ids and counts against what the workspace-based entry point returns for the same frames.

Collected last on purpose (file name), so that a run stopping at the first failure has already run everything else.
"""
import ctypes as C

import numpy as np
import pytest
import torch

from rdp_b200 import _lib, synth


def _static_deproject(masks, depth, K, scale, fmt):
    lib = _lib.load()
    B, H, W = masks.shape
    pts = torch.full((B, H * W, 3), float("nan"), dtype=torch.float32, device=masks.device)
    ids = torch.zeros((B, H * W), dtype=torch.int32, device=masks.device)
    cnt = torch.full((B,), -1, dtype=torch.int32, device=masks.device)
    k4 = (C.c_double * 4)(K[0, 0], K[1, 1], K[0, 2], K[1, 2])
    stream = torch.cuda.current_stream(masks.device).cuda_stream
    rc = lib.rdp_geom_deproject_static(masks.data_ptr(), depth.data_ptr(), fmt, B, H, W, k4, float(scale),
                                       pts.data_ptr(), ids.data_ptr(), cnt.data_ptr(), stream)
    _lib.check(rc, "rdp_geom_deproject_static")
    return pts, ids, cnt


@pytest.mark.gpu
def test_static_scratch_matches_workspace_path_twice():
    from rdp_b200 import geometry
    dev = torch.device("cuda", torch.cuda.current_device())
    calls = [
        ([synth.make_frame(60 + i, 240, 320, 0.10 + 0.1 * i, "u16") for i in range(3)], _lib.DEPTH_U16),
        ([synth.make_frame(70 + i, 123, 187, 0.35, "f32") for i in range(5)], _lib.DEPTH_F32),   # ragged size, more frames
    ]
    for frames, fmt in calls:
        masks = torch.from_numpy(np.stack([f["mask_gt"] for f in frames])).to(dev)
        depth = torch.from_numpy(np.stack([f["depth"] for f in frames])).to(dev)
        K, scale = frames[0]["K"], frames[0]["depth_scale"]
        pts, ids, cnt = _static_deproject(masks, depth, K, scale, fmt)
        ref = geometry.curvature_batch(masks, depth, K, scale, want_points=True, depth_fmt=fmt)
        torch.cuda.synchronize()
        assert torch.equal(cnt, ref.n_points)
        for b in range(len(frames)):
            n = int(cnt[b])
            assert n > 0
            # same tolerance as tests/test_geometry_gpu.py uses for the point cloud (the kernels are the same, so
            # in practice the values are identical)
            np.testing.assert_allclose(pts[b, :n].cpu().numpy(), ref.points_xyz[b, :n].cpu().numpy(), rtol=1e-5, atol=1e-7)
            want_ids = np.flatnonzero((frames[b]["mask_gt"] > 0) & (frames[b]["depth"] > 0))
            assert np.array_equal(ids[b, :n].cpu().numpy().astype(np.int64), want_ids)


def test_static_scratch_rejects_oversized_batches():
    """The limit check comes before any CUDA call, so it also runs on a machine without a GPU."""
    lib = _lib.load()
    k4 = (C.c_double * 4)(600.0, 600.0, 319.5, 239.5)
    dummy = C.c_void_p(256)         # never dereferenced: the call is refused first
    rc = lib.rdp_geom_deproject_static(dummy, dummy, _lib.DEPTH_U16, 65, 480, 640, k4, 0.001, dummy, dummy, dummy, None)
    assert rc == -2 and b"at most 64 frames" in lib.rdp_last_error()
    rc = lib.rdp_geom_deproject_static(dummy, dummy, _lib.DEPTH_U16, 64, 1080, 1920, k4, 0.001, dummy, dummy, dummy, None)
    assert rc == -2 and b"4096 tiles" in lib.rdp_last_error()
    rc = lib.rdp_geom_deproject_static(None, dummy, _lib.DEPTH_U16, 1, 480, 640, k4, 0.001, dummy, dummy, dummy, None)
    assert rc == -1
