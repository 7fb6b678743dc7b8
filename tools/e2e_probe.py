"""Where the end-to-end step goes: FramePipeline with 2 / 3 slots, compute only, copies only (dev aid)."""
import json, os, sys, time
import numpy as np, torch
sys.path[:0] = ['/root/repo', '/root/repo/robotic-discovery-platform_b200']
from pkg.segmentation_model import UNet
from rdp_b200 import geometry, synth
from rdp_b200.pipeline import FramePipeline

B, H, W = 64, 480, 640
dev = torch.device("cuda", 0)
model = UNet(3, 1); model.load_state_dict(synth.make_state_dict(2, planted=True)); model = model.to(dev).eval()
base = [synth.make_frame(5000 + i, H, W, 0.15, "u16") for i in range(16)]
K, scale = base[0]["K"], base[0]["depth_scale"]
RING = 3
host_rgb = [torch.from_numpy(np.stack([base[(i + 5 * r) % 16]["rgb"] for i in range(B)])).pin_memory() for r in range(RING)]
host_depth = [torch.from_numpy(np.stack([base[(i + 5 * r) % 16]["depth"] for i in range(B)])).pin_memory() for r in range(RING)]
out = {}
steps = 20
for slots in (2, 3, 4):
    pipe = FramePipeline(model, K, scale, B, H, W, slots=slots, device=dev)
    for i in range(4):
        pipe.submit(host_rgb[i % RING], host_depth[i % RING]); pipe.collect()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    host_t = 0.0
    for i in range(steps):
        if pipe.in_flight == slots:
            pipe.collect(copy=False)
        h0 = time.perf_counter()
        pipe.submit(host_rgb[i % RING], host_depth[i % RING])
        host_t += time.perf_counter() - h0
    while pipe.in_flight:
        pipe.collect(copy=False)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    out[f"slots{slots}"] = {"ms_per_step": 1e3 * dt / steps, "submit_host_ms": 1e3 * host_t / steps}
    del pipe
    torch.cuda.empty_cache()
# compute only, same loop without copies
rgb = host_rgb[0].to(dev); depth = host_depth[0].to(dev)
uws = model._engine_for(dev).new_workspace(B, H, W)
gws = torch.empty(geometry.workspace_bytes(B, H, W), dtype=torch.uint8, device=dev)
def step():
    mask, _, _ = model.forward_mask(rgb, workspace=uws)
    return geometry.curvature_batch(mask[:, 0], depth, K, scale, workspace=gws)
for _ in range(4): step()
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(steps): step()
torch.cuda.synchronize(); out["compute_only_ms"] = 1e3 * (time.perf_counter() - t0) / steps
# compute with a concurrent H2D stream hammering
s2 = torch.cuda.Stream()
d_rgb = torch.empty_like(rgb); d_depth = torch.empty_like(depth)
torch.cuda.synchronize(); t0 = time.perf_counter()
for i in range(steps):
    with torch.cuda.stream(s2):
        d_rgb.copy_(host_rgb[i % RING], non_blocking=True); d_depth.copy_(host_depth[i % RING], non_blocking=True)
    step()
torch.cuda.synchronize(); out["compute_plus_h2d_stream_ms"] = 1e3 * (time.perf_counter() - t0) / steps
torch.cuda.synchronize(); t0 = time.perf_counter()
for i in range(steps):
    d_rgb.copy_(host_rgb[i % RING], non_blocking=True); d_depth.copy_(host_depth[i % RING], non_blocking=True)
torch.cuda.synchronize(); out["h2d_only_ms"] = 1e3 * (time.perf_counter() - t0) / steps
print(json.dumps(out, indent=1))
