"""Per-GPU streaming pipeline: pinned host frames -> H2D -> U-Net mask -> curvature -> D2H of the small results.

Frames are independent end to end (eval-mode BatchNorm, no cross-frame statistic; reference `server.py:83,116-152`),
so a box with N GPUs runs N of these pipelines, one per process, and gathers results on the host: there is no
data-path collective.  Inside one GPU three streams overlap the stages of consecutive micro-batches:

    copy-in stream   : rgb u8 + depth (pinned host) -> device slot
    compute stream   : UNet.forward_mask -> curvature_batch   (one shared workspace: compute is serial on this stream)
    copy-out stream  : mean / max / status / spline points -> pinned host slot

`frame_shard` is the partitioning rule used across ranks (contiguous blocks, sizes differ by at most one).
"""
from __future__ import annotations

import numpy as np
import torch

from . import geometry


def frame_shard(total: int, world: int, rank: int) -> range:
    """Contiguous block of frame indices owned by `rank`: frame i belongs to rank floor(i*world/total)-style blocks."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    base, extra = divmod(total, world)
    start = rank * base + min(rank, extra)
    return range(start, start + base + (1 if rank < extra else 0))


def max_over_ranks(value: float, group=None, device="cpu") -> float:
    """Timed regions are reported as the slowest rank's time (works with gloo on CPU and nccl on GPU)."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())


class _Slot:
    def __init__(self, B, H, W, depth_dtype, device):
        self.rgb = torch.empty((B, H, W, 3), dtype=torch.uint8, device=device)
        self.depth = torch.empty((B, H, W), dtype=depth_dtype, device=device)
        self.out = {
            "mean": torch.empty(B, dtype=torch.float64).pin_memory(),
            "max": torch.empty(B, dtype=torch.float64).pin_memory(),
            "status": torch.empty(B, dtype=torch.int32).pin_memory(),
            "points": torch.empty((B, geometry.NUM_SAMPLES, 3), dtype=torch.float64).pin_memory(),
        }
        self.loaded = torch.cuda.Event()
        self.computed = torch.cuda.Event()
        self.done = torch.cuda.Event()
        self.free = torch.cuda.Event()       # compute has consumed the device inputs
        self.result = None
        self.busy = False


class FramePipeline:
    """Micro-batches of B frames flow through `submit`; `collect` returns results in submission order."""

    def __init__(self, model, intrinsics, depth_scale, B: int, H: int, W: int, depth_dtype=torch.uint16, slots: int = 2,
                 device=None):
        self.model = model
        self.K = np.asarray(intrinsics, dtype=np.float64)
        self.scale = depth_scale
        self.B, self.H, self.W = B, H, W
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        with torch.cuda.device(self.device):
            self.s_in, self.s_compute, self.s_out = (torch.cuda.Stream(self.device) for _ in range(3))
            self.slots = [_Slot(B, H, W, depth_dtype, self.device) for _ in range(slots)]
            self.unet_ws = model._engine_for(self.device).new_workspace(B, H, W)
            self.geom_ws = torch.empty(geometry.workspace_bytes(B, H, W), dtype=torch.uint8, device=self.device)
        self.depth_fmt = geometry.depth_format(depth_dtype, depth_scale)
        self._next = 0
        self._pending = []
        self.h2d_bytes = B * H * W * (3 + torch.empty(0, dtype=depth_dtype).element_size())
        self.d2h_bytes = sum(t.numel() * t.element_size() for t in self.slots[0].out.values())

    def submit(self, host_rgb: torch.Tensor, host_depth: torch.Tensor) -> None:
        """host tensors should be pinned; returns immediately (work is queued on the three streams)."""
        slot = self.slots[self._next % len(self.slots)]
        self._next += 1
        if slot.busy:                                   # reuse: its previous results must have been collected
            raise RuntimeError("pipeline slot still holds uncollected results: call collect() more often")
        slot.busy = True
        with torch.cuda.stream(self.s_in):
            self.s_in.wait_event(slot.free)             # previous occupant's compute has read the device inputs
            slot.rgb.copy_(host_rgb, non_blocking=True)
            slot.depth.copy_(host_depth, non_blocking=True)
            slot.loaded.record(self.s_in)
        with torch.cuda.stream(self.s_compute):
            self.s_compute.wait_event(slot.loaded)
            mask, _, _ = self.model.forward_mask(slot.rgb, workspace=self.unet_ws)
            res = geometry.curvature_batch(mask[:, 0], slot.depth, self.K, self.scale, workspace=self.geom_ws,
                                           depth_fmt=self.depth_fmt)
            slot.free.record(self.s_compute)
            slot.computed.record(self.s_compute)
        with torch.cuda.stream(self.s_out):
            self.s_out.wait_event(slot.computed)
            slot.out["mean"].copy_(res.mean_curvature, non_blocking=True)
            slot.out["max"].copy_(res.max_curvature, non_blocking=True)
            slot.out["status"].copy_(res.status, non_blocking=True)
            slot.out["points"].copy_(res.spline_points, non_blocking=True)
            slot.done.record(self.s_out)
        slot.result = res                               # keeps the device tensors alive until collected
        for t in (res.mean_curvature, res.max_curvature, res.status, res.spline_points, mask):
            t.record_stream(self.s_out)
        self._pending.append(slot)

    def collect(self, copy: bool = True):
        """Blocks until the oldest submitted micro-batch has finished; returns its host-side results."""
        slot = self._pending.pop(0)
        slot.done.synchronize()
        out = {k: (v.numpy().copy() if copy else v.numpy()) for k, v in slot.out.items()}
        slot.result = None
        slot.busy = False
        return out

    @property
    def in_flight(self) -> int:
        return len(self._pending)

    def events(self):
        return self.s_in, self.s_compute, self.s_out
