"""GPU parity of the geometry kernels (through the C ABI) against the oracle and the reference's golden vectors."""
import numpy as np
import pytest
import torch

from oracle import geometry_port as gp
from rdp_b200 import synth

pytestmark = pytest.mark.gpu


def _run(frames, **kw):
    from rdp_b200 import geometry
    masks = torch.from_numpy(np.stack([f["mask_gt"] for f in frames])).cuda()
    depth = torch.from_numpy(np.stack([f["depth"] for f in frames])).cuda()
    r = geometry.curvature_batch(masks, depth, frames[0]["K"], frames[0]["depth_scale"], want_points=True,
                                 want_edge=True, want_bins=True, **kw)
    torch.cuda.synchronize()
    return r


def _golden_frame(meta):
    """The frame a golden case was generated from (round-2 cases may carry a staircase-edged mask, `stair`)."""
    f = synth.make_frame(meta["seed"], meta["H"], meta["W"], meta["fill"], meta["depth"])
    if meta.get("stair"):
        from oracle import service_port as sp
        n = meta["stair"]
        f["mask_gt"] = np.ascontiguousarray(sp.upsize_mask_np(np.ascontiguousarray(f["mask_gt"][::n, ::n]), meta["H"], meta["W"]))
    return f


def _check_frame(r, i, f, curv_rtol=1e-9):
    o = gp.curvature_profile(f["mask_gt"], f["depth"], f["K"], f["depth_scale"])
    n = int(r.n_points[i])
    assert n == o.n_points
    pcd, pix = gp.deproject(f["mask_gt"], f["depth"], f["K"], f["depth_scale"])
    got = r.points_xyz[i, :n].cpu().numpy().astype(np.float64)
    # fp32 output: within 1e-5 relative of the fp64 reference points (north_star), in the same row-major order
    np.testing.assert_allclose(got, pcd, rtol=1e-5, atol=1e-7)
    if o.n_points >= gp.MIN_POINTS and o.status != gp.DEGENERATE_BINS:
        assert np.array_equal(r.bin_counts[i].cpu().numpy(), o.bin_counts)        # bit-exact bin indices
    st = int(r.status[i])
    assert st == o.status, (st, o.status)
    if o.status != gp.OK:
        assert float(r.mean_curvature[i]) == 0.0 and float(r.max_curvature[i]) == 0.0
        return o
    E = int(r.n_edge[i])
    assert E == o.n_edge
    edge = r.edge_pixels[i, :E].cpu().numpy().astype(np.int64)
    assert np.array_equal(edge, o.edge_pix)                                        # bit-exact selection and order
    assert int(r.n_knots[i]) == o.n_knots and int(r.fit_ier[i]) == o.ier
    if o.n_knots > 8:       # knot insertion + smoothing: same knots, looser (still far inside the 1e-3 bar) numerics
        curv_rtol = max(curv_rtol, 1e-6)
    assert abs(float(r.fit_residual[i]) - o.fp) <= 1e-9 * max(o.fp, 1e-12)
    assert abs(float(r.mean_curvature[i]) - o.mean_curvature) <= curv_rtol * abs(o.mean_curvature)
    assert abs(float(r.max_curvature[i]) - o.max_curvature) <= curv_rtol * abs(o.max_curvature)
    np.testing.assert_allclose(r.spline_points[i].cpu().numpy(), o.spline_points, rtol=0, atol=1e-10)
    return o


@pytest.mark.parametrize("flavour", ["f32", "u16"])
def test_batch_matches_oracle(flavour):
    frames = [synth.make_frame(30 + s, 240, 320, fill, flavour) for s, fill in enumerate([0.05, 0.15, 0.30, 0.45])]
    r = _run(frames)
    for i, f in enumerate(frames):
        o = _check_frame(r, i, f)
        assert o.status == gp.OK


def test_matches_reference_golden_vectors(manifest, golden):
    """Curvature within 1e-3 relative of what the unmodified reference returned (north_star bar); tie-free frames."""
    for name, meta in manifest["cases"].items():
        if meta["kind"] != "geometry" or meta["depth"] != "f32" or meta["n_knots"] != 8:
            continue
        g = golden(name)
        f = synth.make_frame(meta["seed"], meta["H"], meta["W"], meta["fill"], meta["depth"])
        r = _run([f])
        assert int(r.status[0]) == 0 and int(r.n_points[0]) == meta["n_points"] and int(r.n_edge[0]) == meta["n_edge"]
        assert abs(float(r.mean_curvature[0]) - meta["mean_curvature"]) <= 1e-3 * meta["mean_curvature"]
        assert abs(float(r.max_curvature[0]) - meta["max_curvature"]) <= 1e-3 * meta["max_curvature"]
        # in fact the agreement is far tighter than the bar
        assert abs(float(r.mean_curvature[0]) - meta["mean_curvature"]) <= 1e-8 * meta["mean_curvature"]
        np.testing.assert_allclose(r.spline_points[0].cpu().numpy(), g["spline_points"], rtol=0, atol=1e-9)
        W = meta["W"]
        edge = r.edge_pixels[0, : meta["n_edge"]].cpu().numpy().astype(np.int64)
        pcd, pix = gp.deproject(f["mask_gt"], f["depth"], f["K"], f["depth_scale"])
        lut = {int(p): j for j, p in enumerate(pix)}
        pts = pcd[[lut[int(e)] for e in edge]]
        assert np.array_equal(pts[np.lexsort((pts[:, 1], pts[:, 0]))], g["edge_sorted"]), name


def test_knot_insertion_regime_matches_reference(manifest, golden):
    """FITPACK's general branch (fp of the 8-knot cubic > s): knot count, ier and curvature vs the reference's values.

    The bar is 1e-3 relative (north_star); the kernel reproduces FITPACK's knot sequence exactly, so a far tighter
    bound holds - on the thousands-of-knots frames that SURVEY 0.6 expected to be chaotic (640x480 with 2039 knots,
    1280x720 with 3942 and 5409), on staircase-edged (service-shaped) masks, and on uint16 depth, where the value on
    file is the reference's own code run with its tie-dependent NumPy calls made deterministic (`ties: stable`,
    `oracle/make_golden.py`).  A frame whose fit took a decision inside rounding noise says so in `fit_flags`; only
    such a frame may leave the reference's knot sequence, and it is reported.
    """
    cases = [(k, v) for k, v in manifest["cases"].items() if v["kind"] == "geometry" and v["n_knots"] != 8]
    assert len(cases) >= 10
    assert any(v["H"] == 720 for _, v in cases) and any(v["depth"] == "u16" for _, v in cases)
    flagged = []
    for shape_key in sorted({(v["H"], v["W"], v["depth"]) for _, v in cases}):
        group = [(k, v) for k, v in cases if (v["H"], v["W"], v["depth"]) == shape_key]
        frames = [_golden_frame(v) for _, v in group]
        r = _run(frames)
        for i, (name, meta) in enumerate(group):
            g = golden(name)
            assert int(r.status[i]) == 0, name
            assert int(r.n_points[i]) == meta["n_points"] and int(r.n_edge[i]) == meta["n_edge"], name
            rel_mean = abs(float(r.mean_curvature[i]) - meta["mean_curvature"]) / meta["mean_curvature"]
            rel_max = abs(float(r.max_curvature[i]) - meta["max_curvature"]) / meta["max_curvature"]
            flags = int(r.fit_flags[i])
            print(f"{name}: knots {int(r.n_knots[i])} (ref {meta['n_knots']}), ier {int(r.fit_ier[i])}, fit_flags {flags}, "
                  f"rel err mean {rel_mean:.2e} max {rel_max:.2e}")
            tight = 1e-8 if meta["n_knots"] <= 500 else 1e-6
            exact = (int(r.n_knots[i]) == meta["n_knots"] and int(r.fit_ier[i]) == meta["ier"]
                     and rel_mean <= tight and rel_max <= tight)
            if flags == 0:
                assert exact, name
                np.testing.assert_allclose(r.spline_points[i].cpu().numpy(), g["spline_points"], rtol=0, atol=1e-8)
            elif not exact:
                flagged.append(name)
    # the tie-free float-depth frames of round 1 must stay exact whatever their flags say
    for name in ("geom_f32_480x640_f552_knots136", "geom_f32_480x640_f555_knots167", "geom_f32_480x640_f60_knots2039"):
        assert name not in flagged, name
    print("frames that left the reference's knot sequence with fit_flags set:", flagged if flagged else "none")
    assert len(flagged) <= 2, flagged


def test_degenerate_and_mixed_batch():
    H, W = 120, 160
    base = synth.make_frame(20, H, W, 0.3, "f32")
    K = base["K"]

    def fr(mask, depth):
        return {"mask_gt": mask, "depth": depth, "K": K, "depth_scale": 1.0}

    m99 = np.zeros((H, W), np.uint8); m99.reshape(-1)[np.flatnonzero(base["depth"].reshape(-1) > 0)[:99]] = 1
    mcol = np.zeros((H, W), np.uint8); mcol[:, 80] = 1
    dcol = base["depth"].copy(); dcol[:, 80] = 0.5
    mrow = np.zeros((H, W), np.uint8); mrow[60, 10:150] = 1
    frames = [
        fr(np.zeros((H, W), np.uint8), base["depth"]),          # empty mask
        fr(base["mask_gt"], np.zeros((H, W), np.float32)),      # all-zero depth
        fr(m99, base["depth"]),                                 # 99 points
        fr(mcol, dcol),                                         # single column, constant depth: bin width 0
        fr(mrow, np.abs(base["depth"]) + 0.1),                  # 140 points -> tiny but valid fit
        fr(base["mask_gt"], base["depth"]),                     # a normal frame in the same batch
        fr(base["mask_gt"], np.where(base["depth"] > 0, np.float32(np.nan), base["depth"]).astype(np.float32)),  # NaN depth
    ]
    r = _run(frames)
    want = [gp.TOO_FEW_POINTS, gp.TOO_FEW_POINTS, gp.TOO_FEW_POINTS, gp.DEGENERATE_BINS, gp.OK, gp.OK, gp.TOO_FEW_POINTS]
    assert r.status.cpu().tolist() == want
    for i, f in enumerate(frames[:6]):
        _check_frame(r, i, f)


def test_ragged_shape_and_scalar_path():
    """H*W not a multiple of 16 takes the scalar load path; odd sizes must still be exact."""
    f = synth.make_frame(41, 123, 187, 0.3, "u16")
    r = _run([f, synth.make_frame(42, 123, 187, 0.2, "u16")])
    _check_frame(r, 0, f)


def test_numpy_promotion_modes():
    f = synth.make_frame(43, 120, 160, 0.3, "f32")
    f["depth_scale"] = 0.5
    f["depth"] = (f["depth"] * 2).astype(np.float32)
    _check_frame(_run([f]), 0, f)                       # float32 * python float -> float32 product
    f2 = dict(f); f2["depth_scale"] = np.float64(0.5)
    _check_frame(_run([f2]), 0, f2)                     # float32 * numpy float64 -> float64 product


def test_depth_dtypes_follow_numpy_promotion():
    """`depth_image[v, u] * depth_scale` (`geometry_utils.py:105`) in the image's own dtype: float64 stays float64,
    signed and wider integers promote to float64 and negative readings are dropped by `z > 0` (they must not wrap
    into large positive depths); a 0-d float64 ndarray scale (`np.load(...)['depth_scale']`) promotes float32 depth."""
    from pkg.geometry_utils import compute_curvature_profile
    f = synth.make_frame(45, 120, 160, 0.3, "u16")
    mask, K = f["mask_gt"], f["K"]
    d16 = f["depth"]
    want = gp.curvature_profile(mask, d16, K, 0.001)
    assert want.status == gp.OK
    for dt in (np.int16, np.int32, np.uint32, np.int64, np.float64):
        d = d16.astype(dt)
        if np.issubdtype(dt, np.signedinteger):
            d[::7, ::5] = -5                      # invalid readings: dropped, like the zeros
        o = gp.curvature_profile(mask, d, K, 0.001)
        got = compute_curvature_profile(mask=mask, depth_image=d, intrinsics=K, depth_scale=0.001)
        assert o.status == gp.OK and len(got.spline_points) == 100, dt
        assert abs(got.mean_curvature - o.mean_curvature) <= 1e-9 * o.mean_curvature, dt
        assert abs(got.max_curvature - o.max_curvature) <= 1e-9 * o.max_curvature, dt
    f32 = synth.make_frame(46, 120, 160, 0.3, "f32")
    f32["depth"] = (f32["depth"] * 2).astype(np.float32)
    d64 = f32["depth"].astype(np.float64) * 1.0000001        # values float32 cannot hold
    o = gp.curvature_profile(f32["mask_gt"], d64, K, 0.5)
    got = compute_curvature_profile(mask=f32["mask_gt"], depth_image=d64, intrinsics=K, depth_scale=0.5)
    assert abs(got.mean_curvature - o.mean_curvature) <= 1e-9 * o.mean_curvature
    f3 = dict(f32); f3["depth_scale"] = np.asarray(0.5)        # 0-d float64 array
    _check_frame(_run([f3]), 0, f3)


def test_large_bin_uses_global_paths():
    """All points but one fall into a single x-bin: exercises a 66k-point select and the global-memory sort."""
    H, W = 480, 640
    f = synth.make_frame(44, H, W, 0.5, "f32")
    m = np.zeros((H, W), np.uint8); m[:, 200:420] = f["mask_gt"][:, 200:420]; m[240, 0] = 1
    d = f["depth"].copy(); d[240, 0] = 30.0      # far outlier stretches the x range: the band collapses into bin 49
    f2 = {"mask_gt": m, "depth": d, "K": f["K"], "depth_scale": 1.0}
    r = _run([f2])
    o = gp.curvature_profile(m, d, f["K"], 1.0)
    assert o.bin_counts.max() > 6144 and o.bin_k.max() > 2048
    assert np.array_equal(r.bin_counts[0].cpu().numpy(), o.bin_counts)
    E = int(r.n_edge[0])
    assert E == o.n_edge and np.array_equal(r.edge_pixels[0, :E].cpu().numpy().astype(np.int64), o.edge_pix)


def test_full_size_batch_properties():
    """BASELINE config C2/C4 size (640x480, B=64): properties that need no oracle at this size."""
    B, H, W = 64, 480, 640
    frames = [synth.make_frame(100 + i, H, W, 0.15, "u16") for i in range(4)]
    masks = torch.from_numpy(np.stack([f["mask_gt"] for f in frames])).cuda().repeat(B // 4, 1, 1)
    depth = torch.from_numpy(np.stack([f["depth"] for f in frames])).cuda().repeat(B // 4, 1, 1)
    from rdp_b200 import geometry
    r = geometry.curvature_batch(masks, depth, frames[0]["K"], 0.001, want_points=True, want_edge=True, want_bins=True)
    torch.cuda.synchronize()
    assert (r.status == 0).all()
    # identical frames give identical results wherever they sit in the batch (no cross-frame leakage)
    for t in (r.mean_curvature, r.max_curvature, r.n_points, r.n_edge):
        assert torch.equal(t.view(B // 4, 4), t[:4].expand(B // 4, 4))
    assert torch.equal(r.bin_counts.sum(dim=1), r.n_points)
    valid = ((masks > 0) & (depth.to(torch.int32) > 0)).flatten(1).sum(dim=1).to(torch.int32)
    assert torch.equal(valid, r.n_points)
    for i in range(4):
        _check_frame(r, i, frames[i])


def test_more_knot_frames_than_sms_use_the_small_cta_and_agree(manifest):
    """With more frames than SMs k_fit_knots runs 128-thread CTAs (two frames per SM): same knots, same curvature."""
    from rdp_b200 import geometry
    names = ["geom_f32_480x640_f552_knots136", "geom_f32_480x640_f555_knots167", "geom_f32_480x640_f552_knots12"]
    metas = [manifest["cases"][n] for n in names]
    frames = [synth.make_frame(v["seed"], v["H"], v["W"], v["fill"], v["depth"]) for v in metas]
    n_sm = torch.cuda.get_device_properties(0).multi_processor_count
    B = n_sm + 12
    masks = torch.from_numpy(np.stack([frames[i % 3]["mask_gt"] for i in range(B)])).cuda()
    depth = torch.from_numpy(np.stack([frames[i % 3]["depth"] for i in range(B)])).cuda()
    r = geometry.curvature_batch(masks, depth, frames[0]["K"], frames[0]["depth_scale"])
    one = [geometry.curvature_batch(masks[i:i + 1], depth[i:i + 1], frames[0]["K"], frames[0]["depth_scale"]) for i in range(3)]
    torch.cuda.synchronize()
    assert int((r.status == 0).sum()) == B
    for i in range(B):
        v, o = metas[i % 3], one[i % 3]
        assert int(r.n_knots[i]) == v["n_knots"] == int(o.n_knots[0])
        for got, single, ref in ((float(r.mean_curvature[i]), float(o.mean_curvature[0]), v["mean_curvature"]),
                                 (float(r.max_curvature[i]), float(o.max_curvature[0]), v["max_curvature"])):
            assert abs(got - single) <= 1e-9 * abs(single) and abs(got - ref) <= 1e-8 * abs(ref)


@pytest.mark.parametrize("H,W,amp", [(64, 96, 300.0), (64, 96, 3000.0), (96, 128, 3000.0)])
def test_interpolation_knots_regime(H, W, amp):
    """Depth noise so large that FITPACK runs out of knots to add (n = nmax = m + 4: one knot per data point, the
    least-squares spline interpolates, fp = 0 < s) and then smooths back up to fp = s."""
    f = synth.make_frame(40, H, W, 0.5, "f32")
    rng = np.random.default_rng(int(amp) + H)
    f["depth"] = (f["depth"] * (1.0 + amp * rng.random((H, W)))).astype(np.float32)
    o = gp.curvature_profile(f["mask_gt"], f["depth"], f["K"], f["depth_scale"])
    assert o.status == gp.OK and o.n_knots == o.n_edge + 4            # the oracle is in the regime this test is about
    r = _run([f])
    assert int(r.status[0]) == 0 and int(r.n_edge[0]) == o.n_edge
    assert int(r.n_knots[0]) == o.n_knots and int(r.fit_ier[0]) == o.ier
    assert abs(float(r.fit_residual[0]) - o.fp) <= 1e-6 * o.fp
    assert abs(float(r.mean_curvature[0]) - o.mean_curvature) <= 1e-6 * o.mean_curvature
    assert abs(float(r.max_curvature[0]) - o.max_curvature) <= 1e-6 * o.max_curvature
