#!/usr/bin/env python
"""Secondary measurements for BASELINE.json configs C3, C4 and C5 (the headline config C2 is `bench.py`).

    python tools/bench_configs.py [--c3] [--c4] [--c5 N_FRAMES] > profiles/<name>.json

C3  U-Net + curvature at 1280x720, batch 32
C4  geometry only, 640x480 and 1280x720, mask fill 5..60 %, batch 256, each line labelled with the spline regime
C5  streaming: N frames from a ring of 1024 pinned host frames through rdp_b200.pipeline.FramePipeline
S   the server-shaped workload: resize to 256x256 -> U-Net -> nearest up-size -> curvature at camera
    resolution (rdp_b200.service.analyze_frames), device-resident, and the same through MicroBatcher from 10 host threads
Timing: CUDA events on the launching stream, 3 warm-up passes, inputs > L2 or rotated.
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "robotic-discovery-platform_b200"))

from pkg.segmentation_model import UNet  # noqa: E402
from rdp_b200 import geometry, synth  # noqa: E402
from rdp_b200.pipeline import FramePipeline  # noqa: E402


def timed(fn, iters, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def model_on_gpu():
    m = UNet(3, 1)
    m.load_state_dict(synth.make_state_dict(2, planted=True))
    return m.cuda().eval()


def c3():
    B, H, W = 32, 720, 1280
    model = model_on_gpu()
    frames = [synth.make_frame(7000 + i, H, W, 0.15, "u16") for i in range(8)]
    rgb = torch.from_numpy(np.stack([frames[i % 8]["rgb"] for i in range(B)])).cuda()
    depth = torch.from_numpy(np.stack([frames[i % 8]["depth"] for i in range(B)])).cuda()
    ws = model._engine_for(rgb.device).new_workspace(B, H, W)
    gws = torch.empty(geometry.workspace_bytes(B, H, W), dtype=torch.uint8, device="cuda")
    K = frames[0]["K"]

    def step():
        mask, _, _ = model.forward_mask(rgb, workspace=ws)
        return geometry.curvature_batch(mask[:, 0], depth, K, 0.001, workspace=gws)

    ms = timed(step, 5)
    r = step()
    torch.cuda.synchronize()
    return {"config": "C3: UNet+curvature 1280x720 batch 32 bf16", "ms_per_step": ms, "frames_per_s": B / ms * 1e3,
            "tflops_unet_equiv": B * 1124.44e9 / (ms * 1e-3) / 1e12, "frames_ok": int((r.status == 0).sum()),
            "workspace_gb": ws.numel() / 1e9}


def c4():
    out = []
    B = 256
    for (H, W) in ((480, 640), (720, 1280)):
        K = synth.camera_intrinsics(H, W)
        gws = torch.empty(geometry.workspace_bytes(B, H, W), dtype=torch.uint8, device="cuda")
        for fill in (0.05, 0.10, 0.15, 0.20, 0.30, 0.40, 0.50, 0.60):
            frames = [synth.make_frame(8000 + i, H, W, fill, "u16") for i in range(8)]
            masks = torch.from_numpy(np.stack([frames[i % 8]["mask_gt"] for i in range(B)])).cuda()
            depth = torch.from_numpy(np.stack([frames[i % 8]["depth"] for i in range(B)])).cuda()
            ms = timed(lambda: geometry.curvature_batch(masks, depth, K, 0.001, workspace=gws), 5)
            lm = []
            r = geometry.curvature_batch(masks, depth, K, 0.001, workspace=gws, launch_ms=lm)
            torch.cuda.synchronize()
            n = float(r.n_points.float().mean())
            P = H * W
            dp = lm[geometry.LAUNCH_NAMES.index("deproject")]
            out.append({
                "config": f"C4: geometry only {W}x{H} fill {fill:.2f} batch {B}", "ms_per_batch": ms,
                "frames_per_s": B / ms * 1e3, "points_per_frame": n, "edge_per_frame": float(r.n_edge.float().mean()),
                "knots_min": int(r.n_knots.min()), "knots_max": int(r.n_knots.max()),
                "regime": "8-knot cubic (ier=-2)" if int(r.n_knots.max()) == 8 else "knot insertion + smoothing",
                "frames_ok": int((r.status == 0).sum()),
                "launch_ms": {k: round(v, 4) for k, v in zip(geometry.LAUNCH_NAMES, lm)},
                "deproject_gbs_algorithmic": B * (3.0 * P + 4.0 * n) / (dp * 1e-3) / 1e9,
                "survey_bytes_per_frame": 3.0 * P + 24.0 * n,
            })
            print(json.dumps(out[-1]), file=sys.stderr)
    return out


def c5(n_frames):
    B, H, W = 64, 480, 640
    model = model_on_gpu()
    base = [synth.make_frame(9000 + i, H, W, 0.15, "u16") for i in range(32)]
    ring = 1024 // B
    host_rgb = [torch.from_numpy(np.stack([base[(i * 7 + j) % 32]["rgb"] for j in range(B)])).pin_memory() for i in range(ring)]
    host_depth = [torch.from_numpy(np.stack([base[(i * 7 + j) % 32]["depth"] for j in range(B)])).pin_memory() for i in range(ring)]
    pipe = FramePipeline(model, base[0]["K"], 0.001, B, H, W, slots=2)
    pipe.submit(host_rgb[0], host_depth[0]); pipe.collect()
    torch.cuda.synchronize()
    steps = n_frames // B
    s_in, _, s_out = pipe.events()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(s_in)
    ok = 0
    for i in range(steps):
        if pipe.in_flight == 2:
            ok += int((pipe.collect(copy=False)["status"] == 0).sum())
        pipe.submit(host_rgb[i % ring], host_depth[i % ring])
    while pipe.in_flight:
        ok += int((pipe.collect(copy=False)["status"] == 0).sum())
    e1.record(s_out)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    return {"config": f"C5: streaming {steps * B} frames 640x480 from a ring of 1024 pinned host frames, 1 GPU",
            "seconds": ms * 1e-3, "frames_per_s": steps * B / ms * 1e3, "frames_ok": ok}


def server_shaped():
    import threading
    import time

    from rdp_b200 import service
    from rdp_b200.batching import MicroBatcher
    out = []
    model = model_on_gpu()
    for (H, W) in ((480, 640), (720, 1280)):
        frames = [synth.make_frame(9500 + i, H, W, 0.30, "u16") for i in range(8)]
        K, scale = frames[0]["K"], frames[0]["depth_scale"]
        for B in (64, 256):
            rgb = torch.from_numpy(np.stack([frames[i % 8]["rgb"] for i in range(B)])).cuda()
            depth = torch.from_numpy(np.stack([frames[i % 8]["depth"] for i in range(B)])).cuda()
            ms = timed(lambda: service.analyze_frames(model, rgb, depth, K, scale), 5)
            r = service.analyze_frames(model, rgb, depth, K, scale)
            lm = []
            geometry.curvature_batch(r.final_mask, depth, K, scale, launch_ms=lm)
            out.append({"config": f"S1: server loop body {W}x{H} -> 256x256 U-Net -> curvature at {W}x{H}, batch {B}, device-resident",
                        "ms_per_batch": ms, "frames_per_s": B / ms * 1e3, "frames_ok": int((r.curvature.status == 0).sum()),
                        "knots_min": int(r.curvature.n_knots.min()), "knots_max": int(r.curvature.n_knots.max()),
                        "coverage_percent": float(r.mask_coverage_percent.mean()),
                        "geometry_launch_ms": {k: round(v, 4) for k, v in zip(geometry.LAUNCH_NAMES, lm)}})
            print(json.dumps(out[-1]), file=sys.stderr)
        # N2: ten request threads (server.py:172) with host frames, results back on the host (mask included)
        n_threads, per_thread = int(os.environ.get("RDP_S2_THREADS", 10)), 64   # 10 = the servicer's max_workers
        # max_batch = the servicer's worker count: a batch holding every outstanding request goes at once, without max_wait
        lanes = int(os.environ.get("RDP_S2_LANES", 1))
        with MicroBatcher(model, K, scale, max_batch=n_threads, max_wait_ms=2.0, bgr=False, want_mask=False, want_png=True,
                          lanes=lanes) as mb:
            def worker(t):
                for j in range(per_thread):
                    f = frames[(t + j) % 8]
                    mb.analyze(f["rgb"], f["depth"])
            for rep in range(2):          # the first pass warms the staging buffers and the workspace allocator
                threads = [threading.Thread(target=worker, args=(t,)) for t in range(n_threads)]
                t0 = time.perf_counter()
                for t in threads:
                    t.start()
                for t in threads:
                    t.join()
                dt = time.perf_counter() - t0
            out.append({"config": f"S2: MicroBatcher, {n_threads} request threads x {per_thread} frames {W}x{H} from host memory, "
                                  "PNG mask (encoded on the GPU) + curvature back on the host (wall clock)",
                        "frames_per_s": n_threads * per_thread / dt, "lanes": lanes, "batches": mb.batches_run, "frames": mb.frames_run,
                        "mean_batch": mb.frames_run / max(1, mb.batches_run),
                        "dispatcher_seconds": {k: round(v, 4) for k, v in mb._run.seconds.items()}, "wall_seconds_last_pass": dt})
            print(json.dumps(out[-1]), file=sys.stderr)
        # N3: the same closed loop from the request's own payloads (JPEG colour + 16-bit PNG depth, client.py:65-67):
        # (a) host decode in the request thread (cv2.imdecode, server.py:117-118) then analyze(); (b) analyze_encoded(): decoded on the GPU
        import cv2
        enc = [(cv2.imencode(".jpg", np.ascontiguousarray(f["rgb"][:, :, ::-1]))[1].tobytes(), cv2.imencode(".png", f["depth"])[1].tobytes())
               for f in frames]
        for mode in ("host_decode", "gpu_decode"):
            for n_thr in (10, 64):
                with MicroBatcher(model, K, scale, max_batch=n_thr, max_wait_ms=2.0, bgr=True, want_mask=False, want_png=True,
                                  lanes=2 if n_thr > 10 else 1) as mb:
                    def worker(t):
                        for j in range(32):
                            jb, pb = enc[(t + j) % 8]
                            if mode == "host_decode":
                                color = cv2.imdecode(np.frombuffer(jb, np.uint8), cv2.IMREAD_COLOR)
                                depth = cv2.imdecode(np.frombuffer(pb, np.uint8), cv2.IMREAD_UNCHANGED)
                                mb.analyze(color, depth)
                            else:
                                mb.analyze_encoded(jb, pb)
                    for rep in range(2):
                        threads = [threading.Thread(target=worker, args=(t,)) for t in range(n_thr)]
                        t0 = time.perf_counter()
                        for t in threads:
                            t.start()
                        for t in threads:
                            t.join()
                        dt = time.perf_counter() - t0
                    out.append({"config": f"S3: MicroBatcher from encoded payloads ({mode}), {n_thr} request threads x 32 frames {W}x{H}, wall clock",
                                "frames_per_s": n_thr * 32 / dt, "mean_batch": mb.frames_run / max(1, mb.batches_run),
                                "jpeg_kb": len(enc[0][0]) / 1e3, "png_kb": len(enc[0][1]) / 1e3})
                    print(json.dumps(out[-1]), file=sys.stderr)
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--c3", action="store_true")
    ap.add_argument("--c4", action="store_true")
    ap.add_argument("--c5", type=int, default=0)
    ap.add_argument("--server", action="store_true")
    a = ap.parse_args()
    res = {}
    if a.c3:
        res["C3"] = c3()
    if a.c4:
        res["C4"] = c4()
    if a.c5:
        res["C5"] = c5(a.c5)
    if a.server:
        res["S"] = server_shaped()
    print(json.dumps(res, indent=1))
