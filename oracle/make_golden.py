#!/usr/bin/env python
"""Generate `tests/golden/*.npz` by running the UNMODIFIED reference from /root/reference.

Run in the authoring container only (the GPU box has no /root/reference):

    python oracle/make_golden.py

The reference ships no tests or golden vectors, so these files are the pin: outputs of the
reference's own `UNet` (`pkg/segmentation_model.py`), `_get_pcd_from_mask`, `_find_point_cloud_edge`,
`compute_curvature_profile` (`pkg/geometry_utils.py`) and the server's mask rule (`server.py:124`) on inputs
that `rdp_b200.synth` regenerates from seeds.  Library versions are recorded in `manifest.json`.
"""
from __future__ import annotations

import hashlib
import json
import os
import sys
import warnings

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "robotic-discovery-platform_b200"))
REF = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden")

UNET_CASES = [
    # name, weight seed, planted, H, W, frame seeds, fill
    ("unet_random_48x64", 0, False, 48, 64, (0, 1), 0.30),
    ("unet_planted_48x64", 1, True, 48, 64, (2, 3), 0.30),
    ("unet_planted_odd_37x53", 1, True, 37, 53, (4,), 0.30),
    ("unet_planted_96x128", 2, True, 96, 128, (5,), 0.20),
]
GEOM_CASES = [
    # name, seed, H, W, fill, depth flavour
    ("geom_f32_240x320_f05", 10, 240, 320, 0.05, "f32"),
    ("geom_f32_240x320_f15", 11, 240, 320, 0.15, "f32"),
    ("geom_f32_240x320_f30", 12, 240, 320, 0.30, "f32"),
    ("geom_f32_480x640_f15", 13, 480, 640, 0.15, "f32"),
    ("geom_f32_480x640_f30", 14, 480, 640, 0.30, "f32"),
    ("geom_f32_480x640_f55_branchB", 15, 480, 640, 0.55, "f32"),
    ("geom_f32_480x640_f552_knots12", 15, 480, 640, 0.552, "f32"),
    ("geom_f32_480x640_f552_knots136", 21, 480, 640, 0.552, "f32"),
    ("geom_f32_480x640_f555_knots167", 15, 480, 640, 0.555, "f32"),
    ("geom_f32_480x640_f60_knots2039", 15, 480, 640, 0.60, "f32"),   # chaotic regime: reported, not gated
    ("geom_u16_480x640_f15_ties", 16, 480, 640, 0.15, "u16"),
    ("geom_u16_240x320_f30_ties", 17, 240, 320, 0.30, "u16"),
]
# Round 2: the knot-insertion regime at HD resolution, and on what the service really sees (uint16 depth; a mask that
# was up-sized from a low-resolution one, so its edge is a staircase).  Fields as above plus
#   stair  0 = ground-truth mask, n = mask_gt[::n, ::n] up-sized with OpenCV's INTER_NEAREST rule (server.py:125)
#   ties   "none"   = tie-free input, the unmodified reference is a deterministic oracle
#          "stable" = the input has ties (equal x among edge points, equal y at a bin's k-th boundary) and NumPy's
#                     argpartition / default argsort leave their order unspecified: the reference's own
#                     code is run with those two calls made deterministic (larger y first, then the smaller pixel
#                     index; equal x by pixel index - the rule DESIGN.md states), and the unmodified reference's
#                     result is recorded next to it as `*_unmodified` so the tie-induced gap is on file.
GEOM_CASES_R2 = [
    ("geom_f32_720x1280_f48_knots3942", 35, 720, 1280, 0.48, "f32", 0, "none"),
    ("geom_f32_720x1280_f50_knots5409", 33, 720, 1280, 0.50, "f32", 0, "none"),
    ("geom_f32_480x640_f15_stair3", 22, 480, 640, 0.15, "f32", 3, "none"),
    ("geom_u16_480x640_f56_knots_ties", 16, 480, 640, 0.56, "u16", 0, "stable"),
    ("geom_u16_480x640_f15_stair3_ties", 16, 480, 640, 0.15, "u16", 3, "stable"),
    ("geom_u16_240x320_f30_stair2_ties", 17, 240, 320, 0.30, "u16", 2, "stable"),
]


def _import_reference():
    if not os.path.isdir(REF):
        raise SystemExit("make_golden.py needs /root/reference (authoring container only)")
    # the reference's top-level package is called `pkg`, as is our drop-in mirror: load it under a private name
    import importlib.util
    mods = {}
    for name in ("segmentation_model", "geometry_utils"):
        spec = importlib.util.spec_from_file_location(f"_rdp_reference_{name}", os.path.join(REF, "pkg", f"{name}.py"))
        m = importlib.util.module_from_spec(spec)
        sys.modules[spec.name] = m
        spec.loader.exec_module(m)
        mods[name] = m
    return mods["segmentation_model"], mods["geometry_utils"]


class _StableNumpy:
    """`numpy` as the reference module sees it, with the two tie-dependent calls made deterministic.

    `argpartition(y, -k)` (`geometry_utils.py:139`): a full order by (y ascending, index descending) so that the last k
    entries are the k largest y with ties won by the smaller index, those k then listed by index (= pixel order, the
    cloud is row-major).  `argsort(x)` (`:74`): stable, so equal x keep that order.  Everything else is NumPy's own."""

    def __init__(self, real):
        self._np = real

    def __getattr__(self, name):
        return getattr(self._np, name)

    def argpartition(self, a, kth, *args, **kw):
        np_ = self._np
        a = np_.asarray(a)
        order = np_.lexsort((-np_.arange(a.shape[0]), a))
        k = -kth if kth < 0 else a.shape[0] - kth
        order[a.shape[0] - k:] = np_.sort(order[a.shape[0] - k:])
        return order

    def argsort(self, a, *args, **kw):
        return self._np.argsort(a, kind="stable")


def staircase_mask(mask, n, H, W):
    """mask[::n, ::n] brought back to (H, W) by cv2.resize(..., INTER_NEAREST), as the server does (`server.py:125`)."""
    import cv2
    return np.ascontiguousarray(cv2.resize(np.ascontiguousarray(mask[::n, ::n]), (W, H), interpolation=cv2.INTER_NEAREST))


def geometry_record(geo, mask, depth, K, scale):
    """Everything the tests compare for one frame, from the reference module `geo` (possibly with stable ties)."""
    from scipy.interpolate import splprep
    pcd = geo._get_pcd_from_mask(mask, depth, K, scale)
    edge = geo._find_point_cloud_edge(pcd)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        res = geo.compute_curvature_profile(mask=mask, depth_image=depth, intrinsics=K, depth_scale=scale)
        srt = edge[np.argsort(edge[:, 0], kind="stable")]
        (tck, _u), fp, ier, _ = splprep([srt[:, 0], srt[:, 1], srt[:, 2]], s=0.1, k=3, full_output=1)
    xs = np.sort(edge[:, 0])
    return pcd, edge, res, tck, float(fp), int(ier), int((np.diff(xs) == 0).sum())


def sample_positions(shape, n=48, seed=7):
    rng = np.random.default_rng(seed)
    return tuple(rng.integers(0, s, size=n) for s in shape)


def round2_cases(geo, manifest):
    from rdp_b200 import synth
    for name, seed, H, W, fill, flavour, stair, ties in GEOM_CASES_R2:
        f = synth.make_frame(seed, H, W, fill, flavour)
        mask = staircase_mask(f["mask_gt"], stair, H, W) if stair else f["mask_gt"]
        depth, K, scale = f["depth"], f["K"], f["depth_scale"]
        real_np = geo.np
        unmod = geometry_record(geo, mask, depth, K, scale)
        if ties == "stable":
            geo.np = _StableNumpy(real_np)
            try:
                pcd, edge, res, tck, fp, ier, x_ties = geometry_record(geo, mask, depth, K, scale)
            finally:
                geo.np = real_np
        else:
            pcd, edge, res, tck, fp, ier, x_ties = unmod
            assert x_ties == 0, (name, "is listed as tie-free but has equal x among its edge points", x_ties)
        rec = dict(
            n_points=np.int64(pcd.shape[0]),
            pcd_sha256=np.frombuffer(hashlib.sha256(np.ascontiguousarray(pcd, dtype='<f8').tobytes()).digest(), dtype=np.uint8),
            edge_sorted=edge[np.lexsort((edge[:, 1], edge[:, 0]))],
            mean_curvature=np.float64(res.mean_curvature), max_curvature=np.float64(res.max_curvature),
            spline_points=np.array([[p.x, p.y, p.z] for p in res.spline_points], dtype=np.float64),
            t=np.asarray(tck[0]), fp=np.float64(fp), ier=np.int64(ier), x_ties=np.int64(x_ties),
            mean_curvature_unmodified=np.float64(unmod[2].mean_curvature),
            max_curvature_unmodified=np.float64(unmod[2].max_curvature),
        )
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **rec)
        manifest["cases"][name] = dict(kind="geometry", seed=seed, H=H, W=W, fill=fill, depth=flavour, stair=stair, ties=ties,
                                       n_points=int(pcd.shape[0]), n_edge=int(edge.shape[0]), n_knots=int(len(tck[0])),
                                       ier=ier, fp=fp, x_ties=x_ties, mean_curvature=float(res.mean_curvature),
                                       max_curvature=float(res.max_curvature),
                                       mean_curvature_unmodified=float(unmod[2].mean_curvature),
                                       max_curvature_unmodified=float(unmod[2].max_curvature))
        print(name, {k: manifest["cases"][name][k] for k in ("n_points", "n_edge", "n_knots", "ier", "x_ties", "mean_curvature",
                                                              "max_curvature", "mean_curvature_unmodified")})


def main():
    import scipy
    import torch

    from rdp_b200 import synth

    seg, geo = _import_reference()
    os.makedirs(OUT, exist_ok=True)
    torch.manual_seed(0)
    if "--round2" in sys.argv:      # only the cases added in round 2, merged into the existing manifest
        with open(os.path.join(OUT, "manifest.json")) as fh:
            manifest = json.load(fh)
        assert (manifest["numpy"], manifest["scipy"]) == (np.__version__, scipy.__version__), "library versions changed"
        round2_cases(geo, manifest)
        with open(os.path.join(OUT, "manifest.json"), "w") as fh:
            json.dump(manifest, fh, indent=1, sort_keys=True)
        return
    manifest = {"torch": torch.__version__, "numpy": np.__version__, "scipy": scipy.__version__, "cases": {}}

    # ---------------- U-Net ----------------
    for name, wseed, planted, H, W, fseeds, fill in UNET_CASES:
        sd = synth.make_state_dict(wseed, planted=planted)
        model = seg.UNet(3, 1)
        model.load_state_dict(sd)
        model.eval()
        batch = synth.make_batch(fseeds, H, W, fill, "u16")
        x = torch.from_numpy(batch["rgb"]).permute(0, 3, 1, 2).float() / 255.0
        taps = {}
        hooks = []
        for mod_name, mod in model.named_modules():
            if isinstance(mod, torch.nn.ReLU):
                hooks.append(mod.register_forward_hook(
                    lambda m, i, o, key=mod_name: taps.__setitem__(key, o.detach().clone())))
        with torch.no_grad():
            logits = model(x)
            mask = (torch.sigmoid(logits) > 0.5).to(torch.uint8)       # server.py:124
        for h in hooks:
            h.remove()
        rec = {"logits": logits.numpy(), "mask": mask.numpy()}
        for key, val in taps.items():
            # ReLU modules sit at double_conv.2 / .5 -> name the tap after the conv (.0 / .3) that feeds it
            head, idx = key.rsplit(".", 1)
            conv_key = f"{head}.{int(idx) - 2}"
            v = val.numpy()
            pos = sample_positions(v.shape)
            rec[f"tap_mean/{conv_key}"] = np.float64(v.mean(dtype=np.float64))
            rec[f"tap_absmax/{conv_key}"] = np.float64(np.abs(v).max())
            rec[f"tap_samples/{conv_key}"] = v[pos]
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **rec)
        manifest["cases"][name] = dict(kind="unet", weight_seed=wseed, planted=planted, H=H, W=W,
                                       frame_seeds=list(fseeds), fill=fill,
                                       coverage=float(mask.float().mean()))
        print(name, "coverage", manifest["cases"][name]["coverage"],
              "logit range", float(logits.min()), float(logits.max()))

    # ---------------- geometry ----------------
    for name, seed, H, W, fill, flavour in GEOM_CASES:
        f = synth.make_frame(seed, H, W, fill, flavour)
        mask, depth, K, scale = f["mask_gt"], f["depth"], f["K"], f["depth_scale"]
        pcd = geo._get_pcd_from_mask(mask, depth, K, scale)
        edge = geo._find_point_cloud_edge(pcd)
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            res = geo.compute_curvature_profile(mask=mask, depth_image=depth, intrinsics=K, depth_scale=scale)
            # the fit itself, for n_knots / fp / ier (same call as geometry_utils.py:74-78)
            from scipy.interpolate import splprep
            srt = edge[np.argsort(edge[:, 0], kind="stable")]
            (tck, _u), fp, ier, _ = splprep([srt[:, 0], srt[:, 1], srt[:, 2]], s=0.1, k=3, full_output=1)
        # tie census on this frame: equal neighbouring x in the sorted edge, equal y at the k-th boundary
        xs = np.sort(edge[:, 0])
        x_ties = int((np.diff(xs) == 0).sum())
        rec = dict(
            n_points=np.int64(pcd.shape[0]),
            pcd_sha256=np.frombuffer(hashlib.sha256(np.ascontiguousarray(pcd, dtype='<f8').tobytes()).digest(), dtype=np.uint8),
            pcd_head=pcd[:64].copy(),
            pcd_tail=pcd[-64:].copy(),
            edge_sorted=edge[np.lexsort((edge[:, 1], edge[:, 0]))],
            mean_curvature=np.float64(res.mean_curvature),
            max_curvature=np.float64(res.max_curvature),
            spline_points=np.array([[p.x, p.y, p.z] for p in res.spline_points], dtype=np.float64),
            t=np.asarray(tck[0]), c=np.asarray(tck[1]), fp=np.float64(fp), ier=np.int64(ier),
            x_ties=np.int64(x_ties),
        )
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **rec)
        manifest["cases"][name] = dict(kind="geometry", seed=seed, H=H, W=W, fill=fill, depth=flavour,
                                       n_points=int(pcd.shape[0]), n_edge=int(edge.shape[0]),
                                       n_knots=int(len(tck[0])), ier=int(ier), fp=float(fp), x_ties=x_ties,
                                       mean_curvature=float(res.mean_curvature),
                                       max_curvature=float(res.max_curvature))
        print(name, {k: manifest["cases"][name][k] for k in ("n_points", "n_edge", "n_knots", "ier", "x_ties",
                                                              "mean_curvature", "max_curvature")})

    # ---------------- degenerate inputs: what the reference returns ----------------
    degenerate = {}
    H, W = 120, 160
    K = synth.camera_intrinsics(H, W)
    base = synth.make_frame(20, H, W, 0.3, "f32")
    cases = {
        "empty_mask": (np.zeros((H, W), np.uint8), base["depth"]),
        "zero_depth": (base["mask_gt"], np.zeros((H, W), np.float32)),
        "few_points_99": (None, base["depth"]),
        "single_column": (None, base["depth"]),
        "few_edge": (None, base["depth"]),
    }
    m = np.zeros((H, W), np.uint8); m.reshape(-1)[np.flatnonzero(base["depth"].reshape(-1) > 0)[:99]] = 1
    cases["few_points_99"] = (m, base["depth"])
    m = np.zeros((H, W), np.uint8); m[:, 80] = 1
    d = base["depth"].copy(); d[:, 80] = 0.5
    cases["single_column"] = (m, d)      # x identical only if z identical: constant depth column -> bin_width 0
    m = np.zeros((H, W), np.uint8); m[60, 10:150] = 1
    cases["few_edge"] = (m, np.abs(base["depth"]) + 0.1)   # 140 points, 50 bins, k=1 each -> E~50 (>=20): a valid tiny fit
    for key, (mm, dd) in cases.items():
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            r = geo.compute_curvature_profile(mask=mm, depth_image=dd, intrinsics=K, depth_scale=1.0)
        degenerate[key] = dict(mean=float(r.mean_curvature), max=float(r.max_curvature), n_spline=len(r.spline_points))
        print("degenerate", key, degenerate[key])
    manifest["degenerate"] = degenerate
    round2_cases(geo, manifest)

    with open(os.path.join(OUT, "manifest.json"), "w") as fh:
        json.dump(manifest, fh, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
