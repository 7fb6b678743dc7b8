"""Times the N1 pre/post-processing kernels and the whole server loop body (dev aid; prints JSON lines)."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "robotic-discovery-platform_b200")]
from rdp_b200 import service, synth  # noqa: E402


def timed(fn, reps=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    from pkg.segmentation_model import UNet
    B = int(os.environ.get("B", 64))
    for H, W in [(480, 640), (720, 1280)]:
        base = synth.make_batch(range(8), H, W, 0.30, "u16")
        batch = {k: (np.concatenate([v] * (B // 8)) if isinstance(v, np.ndarray) and v.ndim >= 3 else v) for k, v in base.items()}
        rgb = torch.from_numpy(batch["rgb"]).cuda()
        depth = torch.from_numpy(batch["depth"]).cuda()
        small = (torch.rand((B, 256, 256), device="cuda") > 0.7).to(torch.uint8)
        t_pre = timed(lambda: service.preprocess(rgb))
        t_post = timed(lambda: service.upsize_mask(small, H, W))
        model = UNet(3, 1).eval()
        model.load_state_dict(synth.make_state_dict(5, planted=True))
        model = model.cuda()
        t_all = timed(lambda: service.analyze_frames(model, rgb, depth, batch["K"], batch["depth_scale"]), reps=10)
        from rdp_b200 import geometry
        r = service.analyze_frames(model, rgb, depth, batch["K"], batch["depth_scale"])
        x = r.model_input
        eng = model._engine_for(x.device)
        t_unet = timed(lambda: eng.forward(x, want_logits=False, want_mask=True), reps=10)
        t_geom = timed(lambda: geometry.curvature_batch(r.final_mask, depth, batch["K"], batch["depth_scale"]), reps=10)
        ms = []
        g = geometry.curvature_batch(r.final_mask, depth, batch["K"], batch["depth_scale"], launch_ms=ms)
        print(json.dumps(dict(unet256_ms=round(t_unet, 3), geom_ms=round(t_geom, 3),
                              geom_launch_ms=dict(zip(geometry.LAUNCH_NAMES, [round(v, 3) for v in ms])),
                              status=np.bincount(g.status.cpu().numpy(), minlength=6).tolist(),
                              knots=sorted(set(g.n_knots.cpu().numpy().tolist())),
                              coverage=round(float(r.mask_coverage_percent.mean()), 2))))
        t_png = timed(lambda: service.encode_mask_png(r.final_mask, to_host=False))
        import time
        import cv2
        fm = r.final_mask.cpu().numpy()
        t0 = time.perf_counter()
        for i in range(8):
            cv2.imencode(".png", fm[i] * 255)
        t_cv = (time.perf_counter() - t0) / 8 * 1e3
        pngs = service.encode_mask_png(r.final_mask)
        print(json.dumps(dict(png_gpu_ms_per_batch=round(t_png, 4), png_gpu_us_per_frame=round(t_png / B * 1e3, 2),
                              png_cv2_ms_per_frame_one_core=round(t_cv, 3), png_bytes_mean=int(np.mean([len(p) for p in pngs])),
                              cv2_bytes_first=len(cv2.imencode(".png", fm[0] * 255)[1]))))
        pre_bytes = B * (H * W * 3 + 3 * 256 * 256 * 4)
        post_bytes = B * (256 * 256 + H * W)
        print(json.dumps(dict(H=H, W=W, B=B, pre_ms=round(t_pre, 4), pre_GBs=round(pre_bytes / t_pre / 1e6, 1),
                              post_ms=round(t_post, 4), post_GBs=round(post_bytes / t_post / 1e6, 1),
                              loop_body_ms=round(t_all, 3), loop_body_fps=round(B / t_all * 1e3, 1))))


if __name__ == "__main__":
    main()
