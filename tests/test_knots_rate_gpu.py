"""How often the knot-insertion regime of `splprep` (FITPACK, reference `pkg/geometry_utils.py:74-78`) is reproduced on
service-shaped frames, and whether the caller is told when it is not.

Every frame of the server-shaped workload is in this regime: the mask comes back from a 256x256 U-Net by nearest
up-sizing (`services/vision_analysis/server.py:124-125`), its edge is a staircase, the single cubic no longer fits
and FITPACK inserts hundreds of knots, choosing each from the largest interval residual.  Residuals that are equal in
exact arithmetic differ in the last bit in floating point, so a different (equally orthogonal) order of rotations can
put one knot elsewhere and move the curvature by far more than the 1e-3 bar.  This test pins

  * the fraction of frames within 1e-3 of scipy run live on the same edge points;
  * that every frame where the kernel leaves the sequential FITPACK port carries `fit_flags != 0`
    (`RDP_FIT_FLAG_*`, include/rdp_b200.h), with the three-way classification scipy / port / GPU printed for each.
"""
import concurrent.futures
import os

import numpy as np
import pytest
import torch

from oracle import geometry_port as gp
from oracle import service_port as sp
from rdp_b200 import synth

pytestmark = pytest.mark.gpu

RATE_BAR = 0.97        # fraction of frames within 1e-3 of scipy (measured: 0.99)
FOLLOW = 1e-6          # "the same result": knot sequence reproduced, curvature equal to solver rounding


def _frames(n, H, W, seed0, rng_seed):
    rng = np.random.default_rng(rng_seed)
    out = []
    for i in range(n):
        fill = float(rng.uniform(0.15, 0.6))
        f = synth.make_frame(seed0 + i, H, W, fill, "u16" if i % 2 else "f32")
        step = 3 if i % 3 else 2                      # what a low-resolution U-Net's mask looks like after up-sizing
        f["mask_gt"] = np.ascontiguousarray(sp.upsize_mask_np(np.ascontiguousarray(f["mask_gt"][::step, ::step]), H, W))
        out.append(f)
    return out


def _gpu(frames):
    from rdp_b200 import geometry
    got = [None] * len(frames)
    for flavour in ("<u2", "<f4"):
        idx = [i for i, f in enumerate(frames) if f["depth"].dtype.str == flavour]
        if not idx:
            continue
        masks = torch.from_numpy(np.stack([frames[i]["mask_gt"] for i in idx])).cuda()
        depth = torch.from_numpy(np.stack([frames[i]["depth"] for i in idx])).cuda()
        r = geometry.curvature_batch(masks, depth, frames[idx[0]]["K"], frames[idx[0]]["depth_scale"])
        torch.cuda.synchronize()
        st, nk, mean, mx, fl = (t.cpu().numpy() for t in (r.status, r.n_knots, r.mean_curvature, r.max_curvature, r.fit_flags))
        for j, i in enumerate(idx):
            got[i] = dict(status=int(st[j]), n_knots=int(nk[j]), mean=float(mean[j]), max=float(mx[j]), flags=int(fl[j]))
    return got


def _rel(a_mean, a_max, b_mean, b_max):
    return max(abs(a_mean - b_mean) / abs(b_mean), abs(a_max - b_max) / abs(b_max))


def _port(args):
    mask, depth, K, scale = args
    o = gp.curvature_profile(mask, depth, K, scale, fit="port")
    return o.n_knots, o.mean_curvature, o.max_curvature


@pytest.mark.parametrize("n,H,W,seed0", [(256, 240, 320, 1000), (32, 480, 640, 3000)])
def test_knot_regime_rate_and_flags(n, H, W, seed0):
    frames = _frames(n, H, W, seed0, 123)
    got = _gpu(frames)
    ref = [gp.curvature_profile(f["mask_gt"], f["depth"], f["K"], f["depth_scale"], fit="scipy") for f in frames]
    in_regime = [i for i in range(n) if ref[i].status == gp.OK and ref[i].n_knots > 8]
    assert len(in_regime) >= 0.9 * n, "the workload is meant to sit in the knot-insertion regime"
    assert all(got[i]["status"] == ref[i].status for i in range(n))
    rel = {i: _rel(got[i]["mean"], got[i]["max"], ref[i].mean_curvature, ref[i].max_curvature) for i in in_regime}
    within = sum(1 for i in in_regime if rel[i] <= 1e-3)
    differing = [i for i in in_regime if rel[i] > FOLLOW]
    flagged = sum(1 for i in in_regime if got[i]["flags"])
    knots = [ref[i].n_knots for i in in_regime]
    print(f"{W}x{H}: {len(in_regime)} frames in the regime, knots {min(knots)}..{max(knots)}; within 1e-3 of scipy: {within} "
          f"({within / len(in_regime):.4f}); differing by more than {FOLLOW:g}: {len(differing)}; flagged: {flagged} "
          f"({sum(1 for i in in_regime if got[i]['flags'] and rel[i] <= FOLLOW)} of them agree anyway)")
    # three-way classification of every differing frame: does the sequential port follow scipy, does the GPU follow the port
    with concurrent.futures.ProcessPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        ports = list(ex.map(_port, [(frames[i]["mask_gt"], frames[i]["depth"], frames[i]["K"], frames[i]["depth_scale"])
                                    for i in differing]))
    unflagged_departures = []
    for i, (pk, pmean, pmax) in zip(differing, ports):
        g, o = got[i], ref[i]
        port_vs_scipy = _rel(pmean, pmax, o.mean_curvature, o.max_curvature)
        gpu_vs_port = _rel(g["mean"], g["max"], pmean, pmax)
        kind = ("FITPACK itself is not reproducible here (the port leaves scipy, the GPU follows the port)"
                if port_vs_scipy > FOLLOW and gpu_vs_port <= FOLLOW else
                "the GPU's solve order moved a tie-decided knot" if port_vs_scipy <= FOLLOW else
                "port and GPU both leave scipy, differently")
        print(f"  frame {i} ({frames[i]['depth'].dtype}): knots scipy {o.n_knots} port {pk} gpu {g['n_knots']}, gpu_vs_scipy {rel[i]:.2e} "
              f"port_vs_scipy {port_vs_scipy:.2e} gpu_vs_port {gpu_vs_port:.2e}, fit_flags {g['flags']}: {kind}")
        if gpu_vs_port > FOLLOW and not g["flags"]:
            unflagged_departures.append(i)
    assert within / len(in_regime) >= RATE_BAR
    assert not unflagged_departures, f"frames that leave the sequential port without a flag: {unflagged_departures}"
