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
