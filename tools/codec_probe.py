"""Throughput of the GPU wire-format decoders against cv2.imdecode on one host core (dev aid; profiles/r02_codec_probe.json)."""
import json, os, sys, time
import numpy as np, torch, cv2
sys.path[:0] = ['/root/repo', '/root/repo/robotic-discovery-platform_b200']
from rdp_b200 import service, synth

out = {}
for (H, W) in ((480, 640), (720, 1280)):
    frames = [synth.make_frame(900 + i, H, W, 0.15, "u16") for i in range(16)]
    jp = [cv2.imencode(".jpg", np.ascontiguousarray(f["rgb"][:, :, ::-1]))[1].tobytes() for f in frames]
    pn = [cv2.imencode(".png", f["depth"])[1].tobytes() for f in frames]
    t0 = time.perf_counter()
    for x in jp: cv2.imdecode(np.frombuffer(x, np.uint8), cv2.IMREAD_COLOR)
    t_j = (time.perf_counter() - t0) / len(jp)
    t0 = time.perf_counter()
    for x in pn: cv2.imdecode(np.frombuffer(x, np.uint8), cv2.IMREAD_UNCHANGED)
    t_p = (time.perf_counter() - t0) / len(pn)
    row = {"jpeg_kb": float(np.mean([len(x) for x in jp]) / 1e3), "png_kb": float(np.mean([len(x) for x in pn]) / 1e3),
           "raw_kb": H * W * 5 / 1e3, "cv2_jpeg_ms_per_frame_one_core": 1e3 * t_j, "cv2_png_ms_per_frame_one_core": 1e3 * t_p}
    for B in (64, 256, 592):
        jb, pb = [jp[i % 16] for i in range(B)], [pn[i % 16] for i in range(B)]
        for name, fn, payload in (("jpeg", service.decode_color_jpeg, jb), ("png", service.decode_depth_png, pb)):
            fn(payload, check=False); torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0 = time.perf_counter(); e0.record()
            for _ in range(3): fn(payload, check=False)
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 3
            row[f"gpu_{name}_B{B}"] = {"ms_per_batch": ms, "frames_per_s": B / ms * 1e3, "wall_ms_per_batch": 1e3 * (time.perf_counter() - t0) / 3}
    out[f"{W}x{H}"] = row
print(json.dumps(out, indent=1))
