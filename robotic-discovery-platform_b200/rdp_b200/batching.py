"""Request batching shim for the unchanged gRPC servicer.

The reference server handles one frame per loop iteration on each of its 10 worker threads
(`services/vision_analysis/server.py:113-152`, `ThreadPoolExecutor(max_workers=10)` at `:172`).  `MicroBatcher`
sits behind those threads: every thread hands over its decoded frame and blocks; one dispatcher thread coalesces
what has arrived (up to `max_batch` frames, waiting at most `max_wait_ms` after the first) into one GPU micro-batch,
runs the loop body once (`rdp_b200.service.analyze_frames`) and hands each caller its own frame's results.  The wire
format (`protos/vision.proto:21-35`) and the servicer's response-building code stay as they are.
"""
from __future__ import annotations

import queue
import threading
import time
from concurrent.futures import Future
from dataclasses import dataclass
from typing import Callable, List, Optional

import numpy as np


@dataclass
class FrameResult:
    """What one iteration of the server loop needs to fill its response (`server.py:127-143`)."""
    curvature_results: object            # pkg.geometry_utils.CurvatureResult, as compute_curvature_profile returns it
    final_mask: Optional[np.ndarray]     # (H,W) uint8 {0,1}; None when the batcher was built with want_mask=False
    mask_coverage: float                 # percent, server.py:133
    batch_size: int = 1                  # how many frames shared the GPU pass (diagnostics)
    mask_png: Optional[bytes] = None     # the response's `mask` field (server.py:142-143), encoded on the GPU (want_png=True)


@dataclass
class _Request:
    color: object                        # the frame, or a _Staged slot when the runner stages on the request thread
    depth: object
    future: Future
    key: tuple


class _Staged:
    """A frame already copied into a pinned slot by its request thread."""
    __slots__ = ("pool", "slot", "shape")

    def __init__(self, pool, slot, shape):
        self.pool, self.slot, self.shape = pool, slot, shape


class _PinnedPool:
    def __init__(self, n: int, cshape, dshape, depth_dtype):
        import torch
        self.color = torch.empty((n,) + tuple(cshape), dtype=torch.uint8).pin_memory()
        self.depth = torch.empty((n,) + tuple(dshape), dtype=depth_dtype).pin_memory()
        self.color_np, self.depth_np = self.color.numpy(), self.depth.numpy()
        self.free: "queue.Queue[int]" = queue.Queue()
        for i in range(n):
            self.free.put(i)


def _gpu_runner(model, intrinsics, depth_scale, bgr: bool, size: int, device, want_mask: bool, want_png: bool,
                max_batch: int) -> Callable:
    """The default `run_batch`: staged host frames -> list of FrameResult through the CUDA loop body.

    `run.stage(color, depth)` is called by the request thread itself: it copies the frame into a pinned slot (the
    copy releases the GIL, so ten request threads stage in parallel and the dispatcher only issues the H2D copies)."""
    import torch

    from pkg.geometry_utils import _to_result

    from . import service

    dev = torch.device(device)
    pools, pools_lock = {}, threading.Lock()
    lane = threading.local()      # every dispatcher lane (thread) issues its batches on a stream of its own

    def stage(color: np.ndarray, depth: np.ndarray) -> _Staged:
        key = (color.shape, depth.shape, depth.dtype.str)
        with pools_lock:
            pool = pools.get(key)
            if pool is None:  # bounded: one pool per camera format; a full pool makes request threads wait their turn
                pool = pools[key] = _PinnedPool(2 * max_batch + 16, color.shape, depth.shape, torch.from_numpy(depth[:0]).dtype)
        slot = pool.free.get()
        pool.color_np[slot][...] = color
        pool.depth_np[slot][...] = depth
        return _Staged(pool, slot, color.shape)

    def run_encoded(colors: List[bytes], depths: List[bytes]) -> List[FrameResult]:
        """Batches that arrived as the request's own payloads (JPEG colour, 16-bit PNG depth): decoded on the GPU."""
        B = len(colors)
        if getattr(lane, "stream", None) is None:
            with torch.cuda.device(dev):
                lane.stream = torch.cuda.Stream()
        try:
            with torch.cuda.device(dev), torch.cuda.stream(lane.stream):
                r = service.analyze_encoded(model, colors, depths, intrinsics, depth_scale, size=size)
                status = r.curvature.status.cpu().numpy()
                mean, mx = r.curvature.mean_curvature.cpu().numpy(), r.curvature.max_curvature.cpu().numpy()
                pts = r.curvature.spline_points.cpu().numpy()
                cov = r.mask_coverage_percent.cpu().numpy()
                pngs = service.encode_mask_png(r.final_mask) if want_png else [None] * B
                masks = r.final_mask.cpu().numpy() if want_mask else [None] * B
        finally:
            lane.stream.synchronize()
        return [FrameResult(_to_result(int(status[i]), mean[i], mx[i], pts[i]), masks[i], float(cov[i]), B, pngs[i])
                for i in range(B)]

    def run(colors, depths) -> List[FrameResult]:
        if isinstance(colors[0], (bytes, bytearray, memoryview)):
            return run_encoded(colors, depths)
        t0 = time.perf_counter()
        B, pool = len(colors), colors[0].pool
        if getattr(lane, "stream", None) is None:
            with torch.cuda.device(dev):
                lane.stream = torch.cuda.Stream()
        try:
            with torch.cuda.device(dev), torch.cuda.stream(lane.stream):
                dc = torch.empty((B,) + tuple(pool.color.shape[1:]), dtype=torch.uint8, device=dev)
                dd = torch.empty((B,) + tuple(pool.depth.shape[1:]), dtype=pool.depth.dtype, device=dev)
                for i, c in enumerate(colors):
                    dc[i].copy_(pool.color[c.slot], non_blocking=True)
                    dd[i].copy_(pool.depth[c.slot], non_blocking=True)
                t1 = time.perf_counter()
                r = service.analyze_frames(model, dc, dd, intrinsics, depth_scale, size=size, bgr=bgr)
                status = r.curvature.status.cpu().numpy()       # (synchronises: the H2D copies are done from here on)
                mean, mx = r.curvature.mean_curvature.cpu().numpy(), r.curvature.max_curvature.cpu().numpy()
                pts = r.curvature.spline_points.cpu().numpy()
                cov = r.mask_coverage_percent.cpu().numpy()
                pngs = service.encode_mask_png(r.final_mask) if want_png else [None] * B   # a few KB per frame, not H*W
                masks = r.final_mask.cpu().numpy() if want_mask else [None] * B
        finally:
            # The slots must go back whatever happens: after a CUDA error the synchronize raises again, and a pool that
            # never refills would leave every later request thread blocked in stage().
            try:
                lane.stream.synchronize()
            finally:
                for c in colors:
                    pool.free.put(c.slot)
        t2 = time.perf_counter()
        out = [FrameResult(_to_result(int(status[i]), mean[i], mx[i], pts[i]), masks[i], float(cov[i]), B, pngs[i])
               for i in range(B)]
        t3 = time.perf_counter()
        with pools_lock:
            run.seconds["issue_h2d"] += t1 - t0
            run.seconds["gpu_and_copies"] += t2 - t1
            run.seconds["unpack"] += t3 - t2
        return out

    run.stage = stage
    run.seconds = {"issue_h2d": 0.0, "gpu_and_copies": 0.0, "unpack": 0.0}   # where the dispatcher's time went (diagnostics)
    return run


class MicroBatcher:
    """Thread-safe: any number of request threads call `analyze` (blocking) or `submit` (returns a Future)."""

    def __init__(self, model=None, intrinsics=None, depth_scale=None, *, max_batch: int = 64, max_wait_ms: float = 2.0,
                 bgr: bool = True, size: int = 256, device="cuda", want_mask: bool = True, want_png: bool = False,
                 lanes: int = 1, run_batch: Optional[Callable] = None):
        """`lanes` dispatcher threads form and run batches independently (each on its own CUDA stream): with many
        concurrent requests a second lane overlaps one batch's latency-bound spline fits with the next batch's U-Net."""
        if max_batch < 1 or lanes < 1:
            raise ValueError("max_batch and lanes must be >= 1")
        self.max_batch, self.max_wait = int(max_batch), float(max_wait_ms) / 1e3
        self._run = run_batch if run_batch is not None else _gpu_runner(model, intrinsics, depth_scale, bgr, size, device,
                                                                             want_mask, want_png, self.max_batch)
        self._q: "queue.Queue[Optional[_Request]]" = queue.Queue()
        self._closed = False
        self._gate = threading.Lock()      # orders submit()'s enqueue against close()'s stop tokens
        self.batches_run, self.frames_run = 0, 0
        self._stats_lock = threading.Lock()
        self._threads = [threading.Thread(target=self._loop, name=f"rdp-microbatcher-{i}", daemon=True) for i in range(lanes)]
        for t in self._threads:
            t.start()

    # ---- request side ------------------------------------------------------------------------------------
    def submit(self, color_image: np.ndarray, depth_image: np.ndarray) -> Future:
        if self._closed:
            raise RuntimeError("MicroBatcher is closed")
        color, depth = np.asarray(color_image), np.asarray(depth_image)
        if color.ndim != 3 or color.dtype != np.uint8 or depth.shape != color.shape[:2]:
            raise ValueError(f"expected (H,W,C) uint8 colour and (H,W) depth, got {color.shape} {color.dtype} / {depth.shape}")
        fut: Future = Future()
        key = (color.shape, depth.dtype.str)          # frames of different camera formats cannot share a batch
        stage = getattr(self._run, "stage", None)
        if stage is not None:                         # copy into pinned memory here, on the request thread
            color, depth = stage(color, depth), None
        with self._gate:                              # a request is either queued before the stop tokens or refused
            if self._closed:
                if stage is not None:
                    color.pool.free.put(color.slot)
                raise RuntimeError("MicroBatcher is closed")
            self._q.put(_Request(color, depth, fut, key))
        return fut

    def analyze(self, color_image: np.ndarray, depth_image: np.ndarray, timeout: Optional[float] = None) -> FrameResult:
        return self.submit(color_image, depth_image).result(timeout)

    def submit_encoded(self, color_bytes: bytes, depth_bytes: bytes) -> Future:
        """The request's payloads as they came off the wire (`request.color_image.data`, `request.depth_image.data`,
        `server.py:117-118`): JPEG colour and 16-bit PNG depth.  They are decoded on the GPU with the rest of the loop
        body, so the request thread does no image work at all and a fifth of the bytes cross the host-device link."""
        from .service import jpeg_size, png_size
        hw = jpeg_size(color_bytes)
        if png_size(depth_bytes) != hw:
            raise ValueError(f"colour image is {hw}, depth image {png_size(depth_bytes)}")
        fut: Future = Future()
        with self._gate:
            if self._closed:
                raise RuntimeError("MicroBatcher is closed")
            self._q.put(_Request(bytes(color_bytes), bytes(depth_bytes), fut, ("encoded",) + tuple(hw)))
        return fut

    def analyze_encoded(self, color_bytes: bytes, depth_bytes: bytes, timeout: Optional[float] = None) -> FrameResult:
        return self.submit_encoded(color_bytes, depth_bytes).result(timeout)

    def close(self) -> None:
        with self._gate:
            if self._closed:
                return
            self._closed = True
            for _ in self._threads:        # one stop token per lane; what was accepted before it is still served
                self._q.put(None)
        for t in self._threads:
            t.join()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # ---- dispatcher --------------------------------------------------------------------------------------
    def _loop(self) -> None:
        held: List[_Request] = []          # arrived, but for a different camera format than the batch being built
        stop = False
        while not (stop and not held):
            first = held.pop(0) if held else self._q.get()
            if first is None:
                stop = True
                continue
            batch, deadline = [first], time.monotonic() + self.max_wait
            rest, held = held, []
            for r in rest:                  # older leftovers first, in arrival order
                (batch if r.key == first.key and len(batch) < self.max_batch else held).append(r)
            while len(batch) < self.max_batch and not stop:
                left = deadline - time.monotonic()
                try:
                    r = self._q.get(timeout=left) if left > 0 else self._q.get_nowait()
                except queue.Empty:
                    break
                if r is None:
                    stop = True
                elif r.key == first.key:
                    batch.append(r)
                else:
                    held.append(r)
            self._dispatch(batch)

    def _dispatch(self, batch: List[_Request]) -> None:
        try:
            results = self._run([r.color for r in batch], [r.depth for r in batch])
            if len(results) != len(batch):
                raise RuntimeError(f"run_batch returned {len(results)} results for {len(batch)} frames")
        except BaseException as e:  # every waiting request thread gets the error (server.py:154-158 reports it)
            for r in batch:
                r.future.set_exception(e)
            return
        with self._stats_lock:
            self.batches_run += 1
            self.frames_run += len(batch)
        for r, res in zip(batch, results):
            r.future.set_result(res)
