#!/usr/bin/env python
"""Benchmark of the vision hot path: U-Net forward + curvature profile on 640x480 RGB-D frames (BASELINE.json).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

A step is one pass of the hot path over one batch of synthetic frames: `UNet.forward_mask` (18 tensor-core convs
+ fused head) followed by `curvature_batch` (deprojection ... curvature).  Workload = BASELINE.json `configs[1]`:
640x480, batch 64 per GPU, bf16 on 1xB200; with N > 1 every rank runs its own batches (frames shard, no collective).

One JSON line on stdout (rank 0).  Besides the contract keys it carries
  roofline      tensor-core roofline of the dominant kernel (`k_conv3x3`, the 18 conv launches of a step together)
  roofline_hbm  HBM roofline of the geometry streaming kernel (`k_deproject`)
  cpu_baseline  the oracle port of the reference (PyTorch-CPU fp32 U-Net + NumPy/SciPy geometry) on this host
  e2e           the same metric through the public API from pinned HOST buffers (H2D + compute + D2H per step)
`--impl reference` times only that CPU port, on a bounded sample per step.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "robotic-discovery-platform_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

METRIC = "end-to-end RGB-D frames/sec (U-Net+curvature) at 640x480"
UNIT = "frames/s"
FILL = 0.15
WEIGHT_SEED = 2


# --------------------------------------------------------------------------------------------------------------
# work model (DESIGN.md "Measurement"): algorithmic FLOPs / bytes per frame
# --------------------------------------------------------------------------------------------------------------
def conv_flops_per_frame(H: int, W: int):
    """2*MAC of each of the 18 3x3 convs (+ the 1x1 head folded into the last) for one H x W frame."""
    from rdp_b200.synth import unet_layer_table
    level = [0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 3, 3, 2, 2, 1, 1, 0, 0]
    out = []
    for i, (_, _, cin, cout) in enumerate(unet_layer_table(3)):
        h, w = H >> level[i], W >> level[i]
        f = 2.0 * h * w * 9 * cin * cout
        if i == 17:
            f += 2.0 * h * w * 64
        out.append(f)
    return out


def geometry_bytes_per_frame(P: int, n_points: float) -> float:
    """mask u8 + depth u16 read once (3P), fp32 xyz written once (12N) and read once (12N)."""
    return 3.0 * P + 24.0 * n_points


def load_traffic():
    """DRAM bytes from the committed ncu --set full capture (profiles/), or None."""
    path = os.path.join(ROOT, "profiles", "r01_traffic.json")
    if os.path.exists(path):
        with open(path) as fh:
            return json.load(fh)
    return None


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as fh:
            p = json.load(fh)
        return {"tflops": p.get("bf16_tflops_sustained", p.get("bf16_tflops")), "hbm_gbs": p["hbm_gbs"],
                "source": "MEASURED_PEAKS.json (bf16 sustained, hbm copy)"}
    return {"tflops": 1400.0, "hbm_gbs": 6650.0, "source": "fallback (B200_PROFILING.md)"}


# --------------------------------------------------------------------------------------------------------------
# clocks
# --------------------------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU while the timed region runs (pynvml, else nvidia-smi)."""

    def __init__(self, index: int, period: float = 0.02):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop_evt = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _reasons_nvml(self):
        nv = self.nv
        try:
            bits = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
        except Exception:
            bits = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
        names = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
                 0x80: "hw_power_brake_slowdown", 0x2: "applications_clocks_setting"}
        return {n for b, n in names.items() if bits & b}

    def run(self):
        while not self._stop_evt.is_set():
            try:
                if self.nv is not None:
                    self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                    self.reasons |= self._reasons_nvml()
                else:
                    out = subprocess.run(
                        ["nvidia-smi", f"--id={self.index}", "--format=csv,noheader,nounits",
                         "--query-gpu=clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
                         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
                         "clocks_event_reasons.sw_power_cap"], capture_output=True, text=True, timeout=5).stdout
                    f = [x.strip() for x in out.strip().split(",")]
                    self.samples.append(int(f[0]))
                    self.max_mhz = int(f[1])
                    for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[2:]):
                        if val.lower().startswith("active"):
                            self.reasons.add(name)
            except Exception:
                pass
            self._stop_evt.wait(self.period)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=5)
        return {"sm_mhz": statistics.median(self.samples) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


# --------------------------------------------------------------------------------------------------------------
# CPU port of the reference (oracle) - only used by the cpu_baseline leg and --impl reference
# --------------------------------------------------------------------------------------------------------------
def use_all_host_threads():
    """torchrun exports OMP_NUM_THREADS=1; the reference arm is meant to use every host core, as the reference would."""
    import torch
    n = os.cpu_count() or 1
    try:
        n = len(os.sched_getaffinity(0))
    except (AttributeError, OSError):
        pass
    if torch.get_num_threads() < n:
        torch.set_num_threads(n)
    return torch.get_num_threads()


def cpu_reference_frames(sd, frames, K, scale):
    """Reference path as shipped: fp32 U-Net on all host threads, server mask rule, then per-frame geometry."""
    import torch
    from oracle import geometry_port, unet_port
    n = 0
    for f in frames:
        logits = unet_port.unet_forward(sd, unet_port.rgb_u8_to_input(f["rgb"][None]))
        mask = unet_port.mask_from_logits(logits)[0, 0].numpy()
        geometry_port.curvature_profile(mask, f["depth"], K, scale, fit="scipy")
        n += 1
    return n


def time_cpu_reference(H, W, budget_s=15.0, max_frames=64):
    import torch
    from rdp_b200 import synth
    use_all_host_threads()
    sd = synth.make_state_dict(WEIGHT_SEED, planted=True)
    frames = [synth.make_frame(1000 + i, H, W, FILL, "u16") for i in range(2)]
    K, scale = frames[0]["K"], frames[0]["depth_scale"]
    cpu_reference_frames(sd, frames[:1], K, scale)                 # warm-up (oneDNN primitive creation)
    t0 = time.perf_counter()
    cpu_reference_frames(sd, frames[1:2], K, scale)
    per_frame = time.perf_counter() - t0
    n = int(max(2, min(max_frames, budget_s / max(per_frame, 1e-3))))
    frames = [synth.make_frame(2000 + i, H, W, FILL, "u16") for i in range(n)]
    t0 = time.perf_counter()
    cpu_reference_frames(sd, frames, K, scale)
    dt = time.perf_counter() - t0
    return {"value": n / dt, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
            "host_cpus": os.cpu_count(),
            "sample": f"{n} frames {W}x{H} (fill {FILL}, u16 depth), oracle port: torch-CPU fp32 U-Net on "
                      f"{torch.get_num_threads()} threads + NumPy/SciPy geometry, {dt:.1f} s"}


def run_reference(args):
    """`--impl reference`: the reference's CPU implementation (oracle port) on this box's host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import torch
    from rdp_b200 import synth
    use_all_host_threads()
    H, W = args.height, args.width
    sd = synth.make_state_dict(WEIGHT_SEED, planted=True)
    per_step = args.ref_frames
    frames = [synth.make_frame(3000 + i, H, W, FILL, "u16") for i in range(per_step)]
    K, scale = frames[0]["K"], frames[0]["depth_scale"]
    for _ in range(args.warmup):
        cpu_reference_frames(sd, frames[:1], K, scale)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_reference_frames(sd, frames, K, scale)
    dt = time.perf_counter() - t0
    value = per_step * args.steps / dt
    sample = (f"{per_step} frames per step x {args.steps} steps, {W}x{H}, oracle port of the reference on "
              f"{torch.get_num_threads()} threads")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
        "config": {"workload": f"configs[1] sample: UNet+curvature {W}x{H}, {per_step} frames/step on host CPU",
                   "height": H, "width": W, "fill": FILL},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
                         "host_cpus": os.cpu_count(), "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


# --------------------------------------------------------------------------------------------------------------
# the B200 arm
# --------------------------------------------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist

    from pkg.segmentation_model import UNet
    from rdp_b200 import _lib, geometry, synth

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the B200 arm has no CPU fallback; use --impl reference for the CPU port)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    _lib.load()

    B, H, W = args.batch, args.height, args.width
    P = H * W
    sd = synth.make_state_dict(WEIGHT_SEED, planted=True)
    model = UNet(3, 1)
    model.load_state_dict(sd)
    model = model.to(dev).eval()

    # input ring: RING distinct batches, built from 16 distinct seeded frames each rotated differently per slot
    RING = 3
    base = [synth.make_frame(5000 + 97 * rank + i, H, W, FILL, "u16") for i in range(16)]
    K, scale = base[0]["K"], base[0]["depth_scale"]
    host_rgb, host_depth = [], []
    for r in range(RING):
        order = [(i + 5 * r) % 16 for i in range(B)]
        host_rgb.append(torch.from_numpy(np.stack([base[i]["rgb"] for i in order])).pin_memory())
        host_depth.append(torch.from_numpy(np.stack([base[i]["depth"] for i in order])).pin_memory())
    dev_rgb = [t.to(dev) for t in host_rgb]
    dev_depth = [t.to(dev) for t in host_depth]
    unet_ws = model._engine_for(dev).new_workspace(B, H, W)
    geom_ws = torch.empty(geometry.workspace_bytes(B, H, W), dtype=torch.uint8, device=dev)

    def step(i, rgb=None, depth=None):
        s = i % RING
        mask, _, _ = model.forward_mask(dev_rgb[s] if rgb is None else rgb, workspace=unet_ws)
        return geometry.curvature_batch(mask[:, 0], dev_depth[s] if depth is None else depth, K, scale, workspace=geom_ws)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    for i in range(args.warmup):
        res = step(i)
    torch.cuda.synchronize(dev)
    status = res.status.cpu().numpy()
    n_points = float(res.n_points.float().mean())

    # ---- timed region 1: inputs resident in HBM ----
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        step(i)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop()
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    value = world * B * args.steps / (ms * 1e-3)

    # ---- timed region 2: end to end from pinned host buffers through the public API ----
    # FramePipeline = the per-GPU stream pipeline a user runs: H2D of micro-batch i+1 and D2H of i-1 overlap compute of i
    from rdp_b200.pipeline import FramePipeline
    pipe = FramePipeline(model, K, scale, B, H, W, depth_dtype=dev_depth[0].dtype, slots=2, device=dev)
    pipe.submit(host_rgb[0], host_depth[0])
    pipe.collect()
    barrier()
    s_in, _, s_out = pipe.events()
    e0.record(s_in)
    for i in range(args.steps):
        if pipe.in_flight == 2:
            pipe.collect(copy=False)
        pipe.submit(host_rgb[i % RING], host_depth[i % RING])
    while pipe.in_flight:
        last = pipe.collect(copy=False)
    e1.record(s_out)
    barrier()
    ms_e2e = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms_e2e], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_e2e = float(t.item())
    e2e_value = world * B * args.steps / (ms_e2e * 1e-3)
    h2d, d2h = int(pipe.h2d_bytes), int(pipe.d2h_bytes)
    e2e_ok = int((last["status"] == 0).sum())

    # ---- per-launch timing (CUDA events between launches on the launching stream), same inputs ----
    eng = model._engine_for(dev)
    unet_ms = np.zeros(eng.launches_per_forward)
    geom_ms = np.zeros(len(geometry.LAUNCH_NAMES))
    prof_steps = max(1, min(args.steps, 5))
    for i in range(prof_steps):
        lm, gm = [], []
        _, mask, _ = eng.forward(dev_rgb[i % RING], want_logits=False, want_mask=True, workspace=unet_ws, launch_ms=lm)
        geometry.curvature_batch(mask[:, 0], dev_depth[i % RING], K, scale, workspace=geom_ws, launch_ms=gm)
        unet_ms += np.array(lm)
        geom_ms += np.array(gm)
    unet_ms /= prof_steps
    geom_ms /= prof_steps
    names = eng.LAUNCH_NAMES
    conv_idx = [i for i, n in enumerate(names) if n.startswith("conv")]
    flops = conv_flops_per_frame(H, W)
    peaks = load_peaks()
    conv_ms = float(unet_ms[conv_idx].sum())
    conv_tflops = B * sum(flops) / (conv_ms * 1e-3) / 1e12
    per_layer = {names[i]: round(B * flops[j] / (unet_ms[i] * 1e-3) / 1e12, 1) for j, i in enumerate(conv_idx)}
    dp_ms = float(geom_ms[geometry.LAUNCH_NAMES.index("deproject")])
    # k_deproject's own algorithmic bytes: 3P read, 4N pixel ids written (fp32 xyz only when the caller asks for it)
    dp_bytes = B * (3.0 * P + 4.0 * n_points)
    step_ms_prof = float(unet_ms.sum() + geom_ms.sum())
    ncu = load_traffic() if (B, H, W) == (64, 480, 640) else None   # the capture was taken on this workload

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic",
        "config": {"workload": f"configs[1]: UNet+curvature {W}x{H}, batch {B} per GPU, bf16 storage / fp32 accumulate",
                   "batch_per_gpu": B, "height": H, "width": W, "fill": FILL, "depth": "u16",
                   "weights": f"seeded random init (seed {WEIGHT_SEED}) with a planted colour path so that masks follow the object",
                   "parallelism": f"frames sharded over {world} GPU(s), no collective",
                   "l2": f"ring of {RING} input batches ({RING * B * P * 5 / 1e6:.0f} MB) and {unet_ws.numel() / 1e9:.1f} GB of "
                         "activations per step, both larger than the 126 MB L2",
                   "frames_ok": int((status == 0).sum()), "mean_points_per_frame": n_points},
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": ms_e2e / args.steps,
                "frames_ok_last_batch": e2e_ok,
                "how": "rdp_b200.pipeline.FramePipeline: pinned host rgb u8 + depth u16 -> H2D -> UNet.forward_mask -> "
                       "curvature_batch -> D2H of mean/max/status/points, three streams, two slots"},
        "gpu_launches": int((eng.launches_per_forward + len(geometry.LAUNCH_NAMES) - 1) * args.steps),
        "roofline": {"bound": "tensor", "kernel": "k_conv3x3 (18 launches per step)", "achieved": conv_tflops,
                     "peak": peaks["tflops"], "unit": "TFLOP/s", "frac": conv_tflops / peaks["tflops"],
                     "traffic": (ncu["conv_dram_bytes_per_step"] / ncu["conv_launches"]) if ncu else None,
                     "traffic_note": (f"mean DRAM bytes per conv launch, {ncu['source']}" if ncu else None),
                     "peak_source": peaks["source"] + ", of measured",
                     "flops_per_frame": sum(flops), "conv_ms_per_step": conv_ms, "share_of_step": conv_ms / step_ms_prof,
                     "per_layer_tflops": per_layer,
                     "how": f"CUDA events between launches on the launching stream, mean of {prof_steps} steps"},
        "roofline_hbm": {"bound": "hbm", "kernel": "k_deproject", "achieved": dp_bytes / (dp_ms * 1e-3) / 1e9,
                         "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": dp_bytes / (dp_ms * 1e-3) / 1e9 / peaks["hbm_gbs"],
                         "traffic": ncu["deproject_dram_bytes_per_launch"] if ncu else None, "algorithmic_bytes": dp_bytes,
                         "ms": dp_ms, "geometry_ms_per_step": float(geom_ms.sum()),
                         "geometry_launch_ms": {n: round(float(v), 4) for n, v in zip(geometry.LAUNCH_NAMES, geom_ms)},
                         "survey_bytes_per_frame": geometry_bytes_per_frame(P, n_points)},
        "launch_ms": {n: round(float(v), 4) for n, v in zip(names, unet_ms)},
    }
    if rank == 0 and world == 1:
        line["server_shaped"] = time_server_shaped(model, dev_rgb[0], dev_depth[0], K, scale, B, H, W)
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = time_cpu_reference(H, W, budget_s=args.cpu_budget)
    elif rank == 0:
        line["cpu_baseline"] = None
    if rank == 0:
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


_REAL_STDOUT = None


def time_server_shaped(model, rgb, depth, K, scale, B, H, W, reps=5):
    """Secondary figure (not the headline metric): the reference server's own loop body on the same frames - resize to
    256x256, U-Net, nearest up-size, curvature at camera resolution, PNG of the mask (`server.py:120-143`)."""
    import torch

    from rdp_b200 import service

    def step():
        r = service.analyze_frames(model, rgb, depth, K, scale)
        service.encode_mask_png(r.final_mask, to_host=False)
        return r

    for _ in range(3):
        r = step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        r = step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    # the same steps on two alternating streams: one batch's spline fits (one CTA per frame) overlap the next batch's U-Net
    ring = service.StreamRing(2)
    for _ in range(2):
        with ring.next():
            step()
    ring.synchronize()
    e0.record()
    for _ in range(2 * reps):
        with ring.next():
            step()
    ring.synchronize()
    e1.record()
    torch.cuda.synchronize()
    ms2 = e0.elapsed_time(e1) / (2 * reps)
    return {"value": B / ms * 1e3, "unit": "frames/s", "ms_per_step": ms, "two_streams": {"value": B / ms2 * 1e3, "ms_per_step": ms2},
            "frames_ok": int((r.curvature.status == 0).sum()),
            "knots_min": int(r.curvature.n_knots.min()), "knots_max": int(r.curvature.n_knots.max()),
            "workload": f"{B} frames {W}x{H}: rdp_pre_resize -> UNet 256x256 -> rdp_post_mask_nearest -> rdp_geom_curvature at "
                        f"{W}x{H} -> rdp_post_mask_png, device-resident inputs"}


def emit(line: dict) -> None:
    """The JSON line goes to the real stdout; everything else (NCCL banners, warnings) was rerouted to stderr."""
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)      # keep stdout for the one JSON line; libraries that print to fd 1 now land on stderr
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["b200", "reference"], default="b200")
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--height", type=int, default=480)
    ap.add_argument("--width", type=int, default=640)
    ap.add_argument("--ref-frames", type=int, default=4, help="frames per step of the CPU reference arm")
    ap.add_argument("--cpu-budget", type=float, default=15.0, help="seconds of CPU work for the cpu_baseline leg")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else max(args.warmup, 1)
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
