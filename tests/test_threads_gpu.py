"""The reference server runs 10 worker threads that share one model object (`server.py:104,172-175`).

Handles are immutable and workspaces are per call, so concurrent callers must get the same answers as a single one.
"""
import threading

import numpy as np
import pytest
import torch

from rdp_b200 import synth

pytestmark = pytest.mark.gpu


def test_concurrent_callers_share_one_model():
    from pkg.geometry_utils import compute_curvature_profile
    from pkg.segmentation_model import UNet
    sd = synth.make_state_dict(2, planted=True)
    model = UNet(3, 1)
    model.load_state_dict(sd)
    model = model.cuda().eval()
    frames = [synth.make_frame(200 + i, 96, 128, 0.2, "u16") for i in range(6)]

    def run(frame, stream=None):
        x = torch.from_numpy(frame["rgb"][None]).permute(0, 3, 1, 2).float().div(255.0).cuda()
        with torch.cuda.stream(stream) if stream is not None else torch.cuda.stream(torch.cuda.current_stream()):
            logits = model(x)                                               # the call the server makes (server.py:123)
            mask = (torch.sigmoid(logits) > 0.5).squeeze(0).cpu().numpy().transpose(1, 2, 0).astype(np.uint8)[..., 0]
            res = compute_curvature_profile(mask=mask, depth_image=frame["depth"], intrinsics=frame["K"],
                                            depth_scale=frame["depth_scale"])
        return mask, res

    want = [run(f) for f in frames]
    got = [None] * len(frames)
    errors = []

    def worker(i):
        try:
            s = torch.cuda.Stream()
            for _ in range(5):
                got[i] = run(frames[i], s)
        except Exception as e:  # pragma: no cover
            errors.append(e)

    threads = [threading.Thread(target=worker, args=(i,)) for i in range(len(frames))]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    for (m0, r0), (m1, r1) in zip(want, got):
        assert np.array_equal(m0, m1)
        assert r0.mean_curvature == r1.mean_curvature and r0.max_curvature == r1.max_curvature
        assert len(r1.spline_points) == 100


def test_frame_pipeline_returns_what_the_direct_call_returns():
    """Three streams, two slots, shared workspaces: every micro-batch's host-side results, in submission order, equal the
    direct forward_mask + curvature_batch call on the same frames - also when slots are reused many times."""
    import numpy as np
    import torch

    from pkg.segmentation_model import UNet
    from rdp_b200 import geometry, synth
    from rdp_b200.pipeline import FramePipeline
    B, H, W, n_batches = 4, 96, 128, 7
    model = UNet(3, 1).eval()
    model.load_state_dict(synth.make_state_dict(2, planted=True))
    model = model.cuda()
    frames = [synth.make_frame(700 + i, H, W, 0.25 + 0.05 * (i % 4), "u16") for i in range(B * n_batches)]
    K, scale = frames[0]["K"], frames[0]["depth_scale"]
    rgb = [torch.from_numpy(np.stack([f["rgb"] for f in frames[b * B:(b + 1) * B]])).pin_memory() for b in range(n_batches)]
    dep = [torch.from_numpy(np.stack([f["depth"] for f in frames[b * B:(b + 1) * B]])).pin_memory() for b in range(n_batches)]
    pipe = FramePipeline(model, K, scale, B, H, W, slots=2)
    got = []
    for b in range(n_batches):
        if pipe.in_flight == 2:
            got.append(pipe.collect())
        pipe.submit(rgb[b], dep[b])
    while pipe.in_flight:
        got.append(pipe.collect())
    assert len(got) == n_batches
    for b in range(n_batches):
        mask, _, _ = model.forward_mask(rgb[b].cuda())
        r = geometry.curvature_batch(mask[:, 0], dep[b].cuda(), K, scale)
        assert np.array_equal(got[b]["status"], r.status.cpu().numpy())
        assert np.array_equal(got[b]["mean"], r.mean_curvature.cpu().numpy())
        assert np.array_equal(got[b]["max"], r.max_curvature.cpu().numpy())
        assert np.array_equal(got[b]["points"], r.spline_points.cpu().numpy())
    assert any((g["status"] == 0).any() for g in got)
    with pytest.raises(RuntimeError):            # a third submit without a collect would overwrite uncollected results
        pipe.submit(rgb[0], dep[0]); pipe.submit(rgb[1], dep[1]); pipe.submit(rgb[2], dep[2])
