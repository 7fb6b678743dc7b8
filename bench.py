#!/usr/bin/env python
"""Benchmark of the vision hot path: U-Net forward + curvature profile on 640x480 RGB-D frames (BASELINE.json).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--config c2|c3|c4|c5]

A step is one pass of the hot path over one batch of synthetic frames: `UNet.forward_mask` (18 tensor-core convs
+ fused head) followed by `curvature_batch` (deprojection ... curvature).  Workload = BASELINE.json `configs[1]`:
640x480, batch 64 per GPU, bf16 on 1xB200; with N > 1 every rank runs its own batches (frames shard, no collective).

One JSON line on stdout (rank 0).  Besides the contract keys it carries
  roofline      tensor-core roofline of the dominant kernel (`k_conv3x3`, the 18 conv launches of a step together)
  roofline_hbm  HBM roofline of the geometry streaming kernel (`k_deproject`)
  cpu_baseline  the oracle port of the reference (PyTorch-CPU fp32 U-Net + NumPy/SciPy geometry) on this host
  e2e           the same metric through the public API from pinned HOST buffers (H2D + compute + D2H per step)
  cudnn_reference  (c2, measurement only) the reference U-Net's own PyTorch ops on the same GPU through cuDNN: fp32 as
                shipped and bf16 autocast + channels_last - the kernels the hand-written convolutions have to beat
  latency_b1    (c2) ms per frame of the server's unchanged call pattern at batch 1 (`server.py:121-130`)
  training_step (c2, with the cuDNN leg) one optimisation step of the reference's training loop at 256x256 through the
                drop-in module, and through the reference module + cuDNN autograd on the same GPU (measurement only)
`--impl reference` times the reference's own modules (oracle/_ref, else the oracle port) on the host cores, on a
bounded sample per step.

`--config` selects the BASELINE.json configuration (default c2 = configs[1], the one the metric is quoted on):
  c3  configs[2]: U-Net + curvature at 1280x720, batch 32
  c4  configs[3]: geometry only, 640x480 and 1280x720, mask fill 5-60 %, batch 256; `roofline` is the stage-level HBM
      fraction against SURVEY 8(d)'s bytes P*(3+24*fill)
  c5  configs[4]: 65,536 frames streamed from pinned host memory through FramePipeline, sharded over the ranks
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
    for name in ("r02_traffic.json", "r01_traffic.json"):
        path = os.path.join(ROOT, "profiles", name)
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


_REF_MODULES = None


def reference_modules():
    """The UNMODIFIED reference modules (`pkg/segmentation_model.py`, `pkg/geometry_utils.py`) from the git-ignored
    oracle/_ref/ that oracle/make_ref.sh fills in the authoring container, loaded under private names (our drop-in
    mirror owns the name `pkg`).  None when that directory did not travel: the oracle port stands in."""
    global _REF_MODULES
    if _REF_MODULES is None:
        import importlib.util
        base = os.path.join(ROOT, "oracle", "_ref", "pkg")
        mods = []
        for name in ("segmentation_model", "geometry_utils"):
            path = os.path.join(base, name + ".py")
            if not os.path.exists(path):
                mods = None
                break
            spec = importlib.util.spec_from_file_location(f"_rdp_ref_{name}", path)
            m = importlib.util.module_from_spec(spec)
            sys.modules[spec.name] = m
            spec.loader.exec_module(m)
            mods.append(m)
        _REF_MODULES = tuple(mods) if mods else ()
    return _REF_MODULES or None


def reference_kind():
    return "reference" if reference_modules() else "port"


def reference_unet(sd, device="cpu"):
    """The reference's `UNet(3, 1)` with the benchmark's weights, eval mode (`server.py:82-83`), or None."""
    mods = reference_modules()
    if not mods:
        return None
    model = mods[0].UNet(3, 1)
    model.load_state_dict(sd)
    return model.to(device).eval()


def cpu_reference_frames(sd, frames, K, scale, model=None):
    """Reference path as shipped (`server.py:121-130`): fp32 U-Net on all host threads, the server's mask rule, then
    `compute_curvature_profile` per frame - the reference's own modules when oracle/_ref is there, else the port."""
    import warnings

    import torch
    mods = reference_modules()
    n = 0
    if mods:
        model = model if model is not None else reference_unet(sd)
        for f in frames:
            x = torch.from_numpy(f["rgb"]).permute(2, 0, 1).float().div(255.0).unsqueeze(0)       # ToTensor
            with torch.no_grad():
                out = model(x)
                mask = (torch.sigmoid(out) > 0.5).squeeze(0).cpu().numpy().transpose(1, 2, 0).astype(np.uint8)
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                mods[1].compute_curvature_profile(depth_image=f["depth"], mask=mask[:, :, 0], intrinsics=K, depth_scale=scale)
            n += 1
        return n
    from oracle import geometry_port, unet_port
    for f in frames:
        logits = unet_port.unet_forward(sd, unet_port.rgb_u8_to_input(f["rgb"][None]))
        mask = unet_port.mask_from_logits(logits)[0, 0].numpy()
        geometry_port.curvature_profile(mask, f["depth"], K, scale, fit="scipy")
        n += 1
    return n


def _all_cores_worker(args):
    seed, H, W = args
    import torch
    torch.set_num_threads(1)
    from rdp_b200 import synth
    sd = synth.make_state_dict(WEIGHT_SEED, planted=True)
    model = reference_unet(sd)
    f = synth.make_frame(seed, H, W, FILL, "u16")
    cpu_reference_frames(sd, [f], f["K"], f["depth_scale"], model)          # warm-up
    t0 = time.perf_counter()
    cpu_reference_frames(sd, [f], f["K"], f["depth_scale"], model)
    return time.perf_counter() - t0


def time_cpu_all_cores(H, W):
    """BASELINE.md 4.4's throughput-oriented variant: one single-threaded worker per host core, one frame each."""
    import concurrent.futures
    import multiprocessing
    n = os.cpu_count() or 1
    try:
        n = len(os.sched_getaffinity(0))
    except (AttributeError, OSError):
        pass
    t0 = time.perf_counter()
    with concurrent.futures.ProcessPoolExecutor(max_workers=n, mp_context=multiprocessing.get_context("spawn")) as ex:
        per = list(ex.map(_all_cores_worker, [(4000 + i, H, W) for i in range(n)]))
    # steady state: every worker turns out one frame per `t` seconds
    return {"value": sum(1.0 / t for t in per), "unit": UNIT, "workers": n, "threads_per_worker": 1,
            "seconds_per_frame_per_worker": statistics.median(per), "wall_s": time.perf_counter() - t0}


def time_cpu_reference(H, W, budget_s=15.0, max_frames=64, all_cores=False):
    import torch
    from rdp_b200 import synth
    use_all_host_threads()
    sd = synth.make_state_dict(WEIGHT_SEED, planted=True)
    frames = [synth.make_frame(1000 + i, H, W, FILL, "u16") for i in range(2)]
    K, scale = frames[0]["K"], frames[0]["depth_scale"]
    model = reference_unet(sd)
    cpu_reference_frames(sd, frames[:1], K, scale, model)          # warm-up (oneDNN primitive creation)
    t0 = time.perf_counter()
    cpu_reference_frames(sd, frames[1:2], K, scale, model)
    per_frame = time.perf_counter() - t0
    n = int(max(2, min(max_frames, budget_s / max(per_frame, 1e-3))))
    frames = [synth.make_frame(2000 + i, H, W, FILL, "u16") for i in range(n)]
    t0 = time.perf_counter()
    cpu_reference_frames(sd, frames, K, scale, model)
    dt = time.perf_counter() - t0
    what = ("the reference's own modules (oracle/_ref)" if reference_modules() else "oracle port")
    out = {"value": n / dt, "unit": UNIT, "cores": torch.get_num_threads(), "kind": reference_kind(),
           "host_cpus": os.cpu_count(),
           "sample": f"{n} frames {W}x{H} (fill {FILL}, u16 depth), {what}: torch-CPU fp32 U-Net on "
                     f"{torch.get_num_threads()} threads + NumPy/SciPy geometry, {dt:.1f} s"}
    if all_cores:
        try:
            out["all_cores"] = time_cpu_all_cores(H, W)
        except Exception as e:  # a measurement extra must not cost the line
            out["all_cores"] = {"error": repr(e)}
    return out


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
    model = reference_unet(sd)
    for _ in range(args.warmup):
        cpu_reference_frames(sd, frames[:1], K, scale, model)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_reference_frames(sd, frames, K, scale, model)
    dt = time.perf_counter() - t0
    value = per_step * args.steps / dt
    what = "the reference's own modules (oracle/_ref)" if reference_modules() else "oracle port of the reference"
    sample = f"{per_step} frames per step x {args.steps} steps, {W}x{H}, {what} on {torch.get_num_threads()} threads"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
        "config": {"workload": f"{args.config_name} sample: UNet+curvature {W}x{H}, {per_step} frames/step on host CPU",
                   "height": H, "width": W, "fill": FILL},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": torch.get_num_threads(), "kind": reference_kind(),
                         "host_cpus": os.cpu_count(), "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


# --------------------------------------------------------------------------------------------------------------
# measurement-only legs of the c2 line
# --------------------------------------------------------------------------------------------------------------
def time_cudnn_reference(sd, rgb_u8, flops_per_frame, reps=3):
    """"the reference's Blackwell kernel" is whatever PyTorch/cuDNN runs for the unmodified `UNet` on
    this GPU.  Measurement only - nothing of it is on the product path: the reference module itself (oracle/_ref, else the
    oracle's functional restatement of the same torch ops) on CUDA, same batch, same event timing: fp32 as shipped
    (`server.py:57,121-123`; PyTorch's defaults allow TF32 inside cuDNN convolutions), strict fp32, and bf16 autocast +
    channels_last (the strongest stock variant)."""
    import torch
    from oracle import unet_port
    dev = rgb_u8.device
    B = rgb_u8.shape[0]
    x = rgb_u8.permute(0, 3, 1, 2).float().div(255.0).contiguous()
    model = reference_unet(sd, dev)
    sd_dev = None if model is not None else {k: v.to(dev) for k, v in sd.items()}
    torch.backends.cudnn.benchmark = True
    out = {"what": ("the reference's own nn.Module (oracle/_ref)" if model is not None else "oracle/unet_port.py (the same torch ops)")
                   + " on CUDA through PyTorch/cuDNN, measurement only", "batch": B,
           "torch": torch.__version__, "cudnn": torch.backends.cudnn.version()}

    def fwd(inp):
        return model(inp) if model is not None else unet_port.unet_forward(sd_dev, inp)

    def timed(fn):
        for _ in range(2):
            fn()
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize(dev)
        return e0.elapsed_time(e1) / reps

    variants = {}
    try:
        torch.backends.cudnn.allow_tf32 = False
        torch.backends.cuda.matmul.allow_tf32 = False
        with torch.no_grad():
            variants["fp32_strict"] = timed(lambda: fwd(x))              # no TF32 anywhere
            torch.backends.cudnn.allow_tf32 = True
            variants["fp32_as_shipped"] = timed(lambda: fwd(x))          # PyTorch's defaults, which the reference keeps: TF32 convolutions
            torch.backends.cudnn.allow_tf32 = False
            if model is not None:
                model_cl = model.to(memory_format=torch.channels_last)
                xcl = x.contiguous(memory_format=torch.channels_last)
                with torch.autocast("cuda", dtype=torch.bfloat16):
                    variants["bf16_autocast_channels_last"] = timed(lambda: model_cl(xcl))
            else:
                with torch.autocast("cuda", dtype=torch.bfloat16):
                    variants["bf16_autocast"] = timed(lambda: fwd(x))
    except Exception as e:  # out of memory on a smaller GPU etc.: report, do not lose the line
        out["error"] = repr(e)
    for k, ms in variants.items():
        out[k] = {"ms_per_batch": ms, "frames_per_s": B / ms * 1e3, "tflops": B * flops_per_frame / (ms * 1e-3) / 1e12}
    del model, x
    torch.cuda.empty_cache()
    return out


def time_training_step(device, batches=(4, 32), steps=6):
    """one optimisation step of the reference's loop (`scripts/train_segmenter.py:151-165`: forward in
    train mode, BCEWithLogitsLoss, backward, Adam) at 256x256 through the drop-in `UNet` - the reference's batch of 4 and
    a batch of 32 - and, measurement only, the reference module itself through PyTorch/cuDNN autograd on the same GPU."""
    import torch
    from pkg.segmentation_model import UNet
    from rdp_b200 import synth
    out = {"what": "full step (zero_grad, forward, loss, backward, Adam) at 256x256, CUDA events, mean of %d steps after 3" % steps,
           "flops_per_image": "3 x 79.96 GFLOP (forward, input gradient, weight gradient)"}
    sd = synth.make_state_dict(7, planted=False)
    g = torch.Generator(device="cpu").manual_seed(5)

    def timed(model, x, t):
        opt = torch.optim.Adam(model.parameters(), lr=1e-4)
        crit = torch.nn.BCEWithLogitsLoss()

        def step():
            opt.zero_grad()
            loss = crit(model(x), t)
            loss.backward()
            opt.step()
            return loss
        for _ in range(3):
            step()
        torch.cuda.synchronize(device)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            loss = step()
        e1.record()
        torch.cuda.synchronize(device)
        return e0.elapsed_time(e1) / steps, float(loss.detach())

    try:
        for B in batches:
            x = torch.rand((B, 3, 256, 256), generator=g).to(device)
            t = (torch.rand((B, 1, 256, 256), generator=g) > 0.7).float().to(device)
            ours = UNet(3, 1)
            ours.load_state_dict(sd)
            ms, loss = timed(ours.to(device).train(), x, t)
            row = {"ms_per_step": ms, "images_per_s": B / ms * 1e3, "tflops_equiv": 3 * 79.960e9 * B / (ms * 1e-3) / 1e12, "loss": loss}
            del ours
            ref = reference_unet(sd, device)
            if ref is not None:
                torch.backends.cudnn.benchmark = True
                torch.backends.cudnn.allow_tf32 = True     # PyTorch's default, which train_segmenter.py keeps
                rms, rloss = timed(ref.train(), x, t)
                row["reference_cudnn_as_shipped"] = {"ms_per_step": rms, "images_per_s": B / rms * 1e3, "loss": rloss,
                                                     "precision": "fp32 tensors, TF32 allowed inside cuDNN convolutions (PyTorch default)"}
                row["ours_over_reference"] = rms / ms
                del ref
            out["batch%d" % B] = row
            torch.cuda.empty_cache()
    except Exception as e:  # keep the line
        out["error"] = repr(e)
    return out


def time_latency_b1(model, frame, K, scale, reps=30):
    """The production call pattern, unchanged (`server.py:121-130`): one 256x256 frame through `model(x)`, the server's
    `sigmoid > 0.5` + `.cpu()`, `cv2.resize(INTER_NEAREST)` back to camera size, `compute_curvature_profile` on NumPy
    arrays.  Milliseconds per frame, wall clock around the whole body (it ends on the host)."""
    import cv2
    import torch
    from pkg.geometry_utils import compute_curvature_profile
    dev = next(model.parameters()).device
    H, W = frame["depth"].shape
    small = cv2.resize(frame["rgb"], (256, 256), interpolation=cv2.INTER_AREA)
    x_host = torch.from_numpy(small).permute(2, 0, 1).float().div(255.0).unsqueeze(0)

    def body():
        t = [time.perf_counter()]
        input_tensor = x_host.to(dev)
        with torch.no_grad():
            output = model(input_tensor)
            mask_resized = (torch.sigmoid(output) > 0.5).squeeze(0).cpu().numpy().transpose(1, 2, 0).astype(np.uint8)
        t.append(time.perf_counter())
        final_mask = cv2.resize(mask_resized, (W, H), interpolation=cv2.INTER_NEAREST)
        t.append(time.perf_counter())
        res = compute_curvature_profile(depth_image=frame["depth"], mask=final_mask, intrinsics=K, depth_scale=scale)
        t.append(time.perf_counter())
        return res, t

    for _ in range(5):
        res, _ = body()
    parts = []
    for _ in range(reps):
        res, t = body()
        parts.append([1e3 * (t[i + 1] - t[i]) for i in range(3)] + [1e3 * (t[3] - t[0])])
    med = np.median(np.array(parts), axis=0)
    return {"ms_per_frame": float(med[3]), "unet_and_mask_to_host_ms": float(med[0]), "cv2_resize_ms": float(med[1]),
            "compute_curvature_profile_ms": float(med[2]), "frames_per_s_one_thread": 1e3 / float(med[3]),
            "result_points": len(res.spline_points), "reps": reps,
            "how": "wall clock, median; model((1,3,256,256)) -> sigmoid>0.5 -> .cpu() -> cv2.resize NEAREST -> "
                   f"compute_curvature_profile at {W}x{H} on NumPy arrays (server.py:121-130 unchanged)"}


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
    pipe = FramePipeline(model, K, scale, B, H, W, depth_dtype=dev_depth[0].dtype, slots=args.slots, device=dev)
    for i in range(max(args.warmup, args.slots + 1)):       # every slot once: allocator blocks, tensor maps, pinned buffers
        pipe.submit(host_rgb[i % RING], host_depth[i % RING])
        pipe.collect()
    barrier()
    s_in, _, s_out = pipe.events()
    e0.record(s_in)
    for i in range(args.steps):
        if pipe.in_flight == args.slots:
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
    # k_deproject's own algorithmic bytes: 3P read; 4N pixel ids + 16N fp64 x/y keys written (fp32 xyz only on request)
    dp_bytes = B * (3.0 * P + 20.0 * n_points)
    step_ms_prof = float(unet_ms.sum() + geom_ms.sum())
    ncu = load_traffic() if (B, H, W) == (64, 480, 640) else None   # the capture was taken on this workload
    l2_note_gb = unet_ws.numel() / 1e9

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic",
        "config": {"workload": f"{args.config_name}: UNet+curvature {W}x{H}, batch {B} per GPU, bf16 storage / fp32 accumulate",
                   "batch_per_gpu": B, "height": H, "width": W, "fill": FILL, "depth": "u16",
                   "weights": f"seeded random init (seed {WEIGHT_SEED}) with a planted colour path so that masks follow the object",
                   "parallelism": f"frames sharded over {world} GPU(s), no collective",
                   "l2": f"ring of {RING} input batches ({RING * B * P * 5 / 1e6:.0f} MB) and {l2_note_gb:.1f} GB of "
                         "activations per step, both larger than the 126 MB L2",
                   "frames_ok": int((status == 0).sum()), "mean_points_per_frame": n_points},
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": ms_e2e / args.steps,
                "frames_ok_last_batch": e2e_ok,
                "how": "rdp_b200.pipeline.FramePipeline: pinned host rgb u8 + depth u16 -> H2D -> UNet.forward_mask -> "
                       f"curvature_batch -> D2H of mean/max/status/points, three streams, {args.slots} slots"},
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
                         "survey_bytes_per_frame": geometry_bytes_per_frame(P, n_points),
                         "stage": {"what": "all geometry launches of a step against SURVEY 8(d)'s bytes P*(3+24*fill)",
                                   "achieved": B * geometry_bytes_per_frame(P, n_points) / (float(geom_ms.sum()) * 1e-3) / 1e9,
                                   "frac": B * geometry_bytes_per_frame(P, n_points) / (float(geom_ms.sum()) * 1e-3) / 1e9 / peaks["hbm_gbs"]}},
        "launch_ms": {n: round(float(v), 4) for n, v in zip(names, unet_ms)},
    }
    if rank == 0 and world == 1 and args.config == "c2":
        line["server_shaped"] = time_server_shaped(model, dev_rgb[0], dev_depth[0], K, scale, B, H, W)
        line["latency_b1"] = time_latency_b1(model, base[0], K, scale)
        if not args.no_cudnn:
            del unet_ws, geom_ws, pipe          # room for the fp32 reference's activations
            torch.cuda.empty_cache()
            line["cudnn_reference"] = time_cudnn_reference(sd, dev_rgb[0], sum(flops))
            best = max((v["frames_per_s"] for v in line["cudnn_reference"].values() if isinstance(v, dict)), default=None)
            if best:
                line["cudnn_reference"]["our_unet_frames_per_s"] = B / (float(unet_ms.sum()) * 1e-3)
                line["cudnn_reference"]["ours_over_best_cudnn"] = B / (float(unet_ms.sum()) * 1e-3) / best
            line["training_step"] = time_training_step(dev)
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = time_cpu_reference(H, W, budget_s=args.cpu_budget, all_cores=(args.config == "c2"))
    elif rank == 0:
        line["cpu_baseline"] = None
    if rank == 0:
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def _dist_setup():
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the B200 arm has no CPU fallback; use --impl reference for the CPU port)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    return world, rank, local_rank, dev


def _max_ms(ms, world, dev):
    import torch
    import torch.distributed as dist
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    return ms


def run_c4(args):
    """configs[3]: geometry only - deprojection, binned edge, spline curvature - at 640x480 and 1280x720, mask fill
    5-60 %, batch 256.  A step is one pass of `curvature_batch` over each of the 16 (resolution, fill) batches."""
    import torch
    import torch.distributed as dist

    from rdp_b200 import _lib, geometry, synth
    world, rank, local_rank, dev = _dist_setup()
    _lib.load()
    B = args.batch
    peaks = load_peaks()
    fills = (0.05, 0.10, 0.15, 0.20, 0.30, 0.40, 0.50, 0.60)
    points = []
    for (H, W) in ((480, 640), (720, 1280)):
        K = synth.camera_intrinsics(H, W)
        gws = torch.empty(geometry.workspace_bytes(B, H, W), dtype=torch.uint8, device=dev)
        for fill in fills:
            frames = [synth.make_frame(8000 + 31 * rank + i, H, W, fill, "u16") for i in range(8)]
            masks = torch.from_numpy(np.stack([frames[i % 8]["mask_gt"] for i in range(B)])).to(dev)
            depth = torch.from_numpy(np.stack([frames[i % 8]["depth"] for i in range(B)])).to(dev)
            points.append(dict(H=H, W=W, fill=fill, K=K, masks=masks, depth=depth, gws=gws))
    for pt in points:
        for _ in range(args.warmup):
            r = geometry.curvature_batch(pt["masks"], pt["depth"], pt["K"], 0.001, workspace=pt["gws"])
        pt["ok"] = int((r.status == 0).sum())
        pt["n_points"] = float(r.n_points.float().mean())
        pt["knots"] = (int(r.n_knots.min()), int(r.n_knots.max()))
    sampler = ClockSampler(local_rank)
    sampler.start()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(len(points) + 1)]
    acc = np.zeros(len(points))
    t_all0, t_all1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_all0.record()
    for _ in range(args.steps):
        ev[0].record()
        for j, pt in enumerate(points):
            geometry.curvature_batch(pt["masks"], pt["depth"], pt["K"], 0.001, workspace=pt["gws"])
            ev[j + 1].record()
        ev[-1].synchronize()
        acc += np.array([ev[j].elapsed_time(ev[j + 1]) for j in range(len(points))])
    t_all1.record()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    ms = _max_ms(t_all0.elapsed_time(t_all1), world, dev)
    clocks = sampler.stop()
    acc /= args.steps
    rows, bytes_a, ms_a = [], 0.0, 0.0
    for j, pt in enumerate(points):
        P = pt["H"] * pt["W"]
        lm = []
        geometry.curvature_batch(pt["masks"], pt["depth"], pt["K"], 0.001, workspace=pt["gws"], launch_ms=lm)
        stage_bytes = B * geometry_bytes_per_frame(P, pt["n_points"])
        cubic = pt["knots"][1] == 8
        if cubic:
            bytes_a += stage_bytes
            ms_a += acc[j]
        rows.append({"resolution": f"{pt['W']}x{pt['H']}", "fill": pt["fill"], "ms_per_batch": round(float(acc[j]), 4),
                     "frames_per_s": B / acc[j] * 1e3, "points_per_frame": pt["n_points"], "frames_ok": pt["ok"],
                     "knots_min": pt["knots"][0], "knots_max": pt["knots"][1],
                     "regime": "8-knot cubic (ier=-2)" if cubic else "knot insertion + smoothing (latency-bound fits)",
                     "stage_gbs": stage_bytes / (acc[j] * 1e-3) / 1e9, "stage_frac_of_hbm": stage_bytes / (acc[j] * 1e-3) / 1e9 / peaks["hbm_gbs"],
                     "launch_ms": {k: round(v, 4) for k, v in zip(geometry.LAUNCH_NAMES, lm)}})
    frames_per_step = B * len(points)
    value = world * frames_per_step * args.steps / (ms * 1e-3)
    line = {
        "metric": "geometry-only frames/sec (deprojection + binned edge + spline curvature), 640x480 and 1280x720, fill 5-60 %",
        "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"{args.config_name}: geometry only, 2 resolutions x 8 fills, batch {B} per point, masks supplied",
                   "batch_per_point": B, "depth": "u16", "l2": "16 input batches (7.6 GB) rotate through every step, each larger than the 126 MB L2",
                   "parallelism": f"every rank runs the sweep on its own frames ({world} GPU(s)), no collective"},
        "clocks": clocks, "gpu_launches": int(len(geometry.LAUNCH_NAMES) - 1) * len(points) * args.steps,
        "e2e": None,
        "roofline": {"bound": "hbm", "kernel": "geometry stage (all launches), points in the single-cubic regime",
                     "achieved": bytes_a / (ms_a * 1e-3) / 1e9 if ms_a else None, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                     "frac": bytes_a / (ms_a * 1e-3) / 1e9 / peaks["hbm_gbs"] if ms_a else None, "traffic": None,
                     "algorithmic_bytes": "P*(3+24*fill) per frame (SURVEY 8d): mask+depth read once, fp32 xyz written and read once",
                     "peak_source": peaks["source"]},
        "points": rows,
    }
    if rank == 0:
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def run_c5(args):
    """configs[4]: 65,536 frames of 640x480 streamed from pinned host memory, sharded over the ranks (contiguous blocks,
    `pipeline.frame_shard`), every rank its own `FramePipeline`; no collective on the data path."""
    import torch
    import torch.distributed as dist

    from pkg.segmentation_model import UNet
    from rdp_b200 import _lib, synth
    from rdp_b200.pipeline import FramePipeline, frame_shard
    world, rank, local_rank, dev = _dist_setup()
    _lib.load()
    B, H, W = args.batch, args.height, args.width
    total = args.total_frames if args.steps is None else args.steps * B * world
    mine = frame_shard(total, world, rank)
    steps = len(mine) // B
    model = UNet(3, 1)
    model.load_state_dict(synth.make_state_dict(WEIGHT_SEED, planted=True))
    model = model.to(dev).eval()
    base = [synth.make_frame(9000 + 131 * rank + i, H, W, FILL, "u16") for i in range(32)]
    ring = max(1, 1024 // B)                 # a ring of 1,024 distinct frame slots in pinned memory
    host_rgb = [torch.from_numpy(np.stack([base[(i * 7 + j) % 32]["rgb"] for j in range(B)])).pin_memory() for i in range(ring)]
    host_depth = [torch.from_numpy(np.stack([base[(i * 7 + j) % 32]["depth"] for j in range(B)])).pin_memory() for i in range(ring)]
    slots = args.slots
    pipe = FramePipeline(model, base[0]["K"], 0.001, B, H, W, slots=slots, device=dev)
    for i in range(max(3, args.warmup)):
        pipe.submit(host_rgb[i % ring], host_depth[i % ring])
        pipe.collect()
    sampler = ClockSampler(local_rank)
    sampler.start()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    s_in, _, s_out = pipe.events()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(s_in)
    ok = 0
    for i in range(steps):
        if pipe.in_flight == slots:
            ok += int((pipe.collect(copy=False)["status"] == 0).sum())
        pipe.submit(host_rgb[i % ring], host_depth[i % ring])
    while pipe.in_flight:
        ok += int((pipe.collect(copy=False)["status"] == 0).sum())
    e1.record(s_out)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    ms = _max_ms(e0.elapsed_time(e1), world, dev)
    clocks = sampler.stop()
    frames = steps * B * world
    if world > 1:
        t = torch.tensor([ok], device=dev)
        dist.all_reduce(t)
        ok = int(t.item())
    value = frames / (ms * 1e-3)
    eng = model._engine_for(dev)
    from rdp_b200 import geometry
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": max(3, args.warmup),
        "ms_per_step": ms / max(1, steps), "higher_is_better": True, "scaling": "strong" if args.steps is None else "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": f"{args.config_name}: {frames} frames {W}x{H} streamed from pinned host memory, micro-batches of {B}, "
                               f"{world} GPU(s), {steps} steps per rank", "total_frames": frames, "batch_per_gpu": B, "slots": slots,
                   "ring": f"{ring * B} distinct frame slots per rank in pinned host memory ({ring * B * H * W * 5 / 1e9:.2f} GB), larger than L2",
                   "parallelism": f"contiguous frame blocks per rank ({world} ranks), no collective", "frames_ok": ok},
        "clocks": clocks, "seconds": ms * 1e-3,
        "gpu_launches": int((eng.launches_per_forward + len(geometry.LAUNCH_NAMES) - 1) * steps),
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": int(pipe.h2d_bytes), "d2h_bytes_per_step": int(pipe.d2h_bytes),
                "how": "this configuration is end to end by construction: every micro-batch is copied from pinned host memory "
                       "and its results are copied back inside the timed region"},
        "roofline": None, "cpu_baseline": None,
    }
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
    ap.add_argument("--steps", type=int, default=None, help="timed steps (default: 40 for c2, 20 for c3, 3 for c4, the whole stream for c5)")
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["b200", "reference"], default="b200")
    ap.add_argument("--config", choices=["c2", "c3", "c4", "c5"], default="c2", help="BASELINE.json configs[1..4]")
    ap.add_argument("--batch", type=int, default=None)
    ap.add_argument("--height", type=int, default=None)
    ap.add_argument("--width", type=int, default=None)
    ap.add_argument("--total-frames", type=int, default=65536, help="c5: frames in the stream (all ranks together)")
    ap.add_argument("--slots", type=int, default=3, help="c5: micro-batches in flight per GPU")
    ap.add_argument("--no-cudnn", action="store_true", help="c2: skip the cuDNN measurement leg")
    ap.add_argument("--ref-frames", type=int, default=4, help="frames per step of the CPU reference arm")
    ap.add_argument("--cpu-budget", type=float, default=15.0, help="seconds of CPU work for the cpu_baseline leg")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    # default step counts: long enough that filling and draining the three-slot pipeline (one exposed host-to-device copy,
    # ~3.5 ms) is a small part of the end-to-end region: 2-4 % of it at 10 steps, 0.7 % at 40
    defaults = {"c2": (64, 480, 640, 40, "configs[1]"), "c3": (32, 720, 1280, 20, "configs[2]"),
                "c4": (256, 0, 0, 3, "configs[3]"), "c5": (64, 480, 640, None, "configs[4]")}[args.config]
    args.batch = defaults[0] if args.batch is None else args.batch
    args.height = defaults[1] if args.height is None else args.height
    args.width = defaults[2] if args.width is None else args.width
    args.config_name = defaults[4]
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else max(args.warmup, 1)
    if args.impl == "reference":
        if args.config in ("c4", "c5"):          # the CPU arm is defined on the U-Net + curvature step (640x480)
            args.height, args.width = 480, 640
        args.steps = 10 if args.steps is None else args.steps
        return run_reference(args)
    if args.config == "c5":
        return run_c5(args)
    args.steps = defaults[3] if args.steps is None else args.steps
    if args.config == "c4":
        return run_c4(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
