"""Host logic of the request batching shim: coalescing, order, formats, errors - no GPU needed."""
import threading
import time

import numpy as np
import pytest

from rdp_b200.batching import FrameResult, MicroBatcher


def _fake_runner(log, delay=0.0):
    def run(colors, depths):
        log.append(len(colors))
        assert len({c.shape for c in colors}) == 1 and len({d.dtype for d in depths}) == 1
        if delay:
            time.sleep(delay)
        # the "result" identifies the frame: its first colour byte and depth value
        return [FrameResult(None, np.full(c.shape[:2], c[0, 0, 0], np.uint8), float(d[0, 0]), len(colors))
                for c, d in zip(colors, depths)]
    return run


def _frame(tag, H=4, W=6, dtype=np.uint16):
    return np.full((H, W, 3), tag, np.uint8), np.full((H, W), tag, dtype)


def test_every_caller_gets_its_own_frame_and_frames_are_coalesced():
    log = []
    with MicroBatcher(run_batch=_fake_runner(log, delay=0.02), max_batch=8, max_wait_ms=20) as mb:
        out = {}

        def worker(tag):
            r = mb.analyze(*_frame(tag))
            out[tag] = (int(r.final_mask[0, 0]), r.mask_coverage, r.batch_size)

        threads = [threading.Thread(target=worker, args=(t,)) for t in range(1, 31)]   # 30 request threads
        for t in threads:
            t.start()
        for t in threads:
            t.join(10)
        assert sorted(out) == list(range(1, 31))
        assert all(out[t][0] == t and out[t][1] == float(t) for t in out)
        assert sum(log) == 30 and max(log) <= 8 and len(log) < 30          # batches formed, never above max_batch
        assert mb.frames_run == 30 and mb.batches_run == len(log)


def test_single_request_is_not_held_longer_than_max_wait():
    log = []
    with MicroBatcher(run_batch=_fake_runner(log), max_batch=64, max_wait_ms=5) as mb:
        t0 = time.monotonic()
        r = mb.analyze(*_frame(7))
        assert time.monotonic() - t0 < 1.0 and r.batch_size == 1 and log == [1]


def test_full_batch_goes_without_waiting_and_order_is_arrival_order():
    log = []
    with MicroBatcher(run_batch=_fake_runner(log), max_batch=4, max_wait_ms=10_000) as mb:
        futs = [mb.submit(*_frame(t)) for t in range(1, 9)]
        res = [f.result(5) for f in futs]                       # would time out if it waited for max_wait
        assert [int(r.final_mask[0, 0]) for r in res] == list(range(1, 9))
        assert log == [4, 4]


def test_camera_formats_do_not_share_a_batch():
    log = []
    with MicroBatcher(run_batch=_fake_runner(log), max_batch=16, max_wait_ms=30) as mb:
        futs = [mb.submit(*_frame(1)), mb.submit(*_frame(2, H=8, W=8)), mb.submit(*_frame(3)),
                mb.submit(*_frame(4, dtype=np.float32)), mb.submit(*_frame(5, H=8, W=8))]
        res = [f.result(5) for f in futs]
        assert [int(r.final_mask[0, 0]) for r in res] == [1, 2, 3, 4, 5]
        assert [r.final_mask.shape for r in res] == [(4, 6), (8, 8), (4, 6), (4, 6), (8, 8)]
        assert sorted(log) == [1, 2, 2]


def test_errors_reach_every_waiting_caller_and_the_batcher_survives():
    calls = []

    def run(colors, depths):
        calls.append(len(colors))
        if colors[0][0, 0, 0] == 99:
            raise RuntimeError("engine failure")
        return [FrameResult(None, c[:, :, 0], 0.0, len(colors)) for c in colors]

    with MicroBatcher(run_batch=run, max_batch=4, max_wait_ms=20) as mb:
        bad = [mb.submit(*_frame(99)) for _ in range(3)]
        for f in bad:
            with pytest.raises(RuntimeError, match="engine failure"):
                f.result(5)
        assert mb.analyze(*_frame(1)).batch_size == 1                      # still serving
    with pytest.raises(ValueError):
        MicroBatcher(run_batch=run, max_batch=0)


def test_bad_frames_are_refused_at_submit_and_close_is_final():
    mb = MicroBatcher(run_batch=_fake_runner([]), max_batch=2, max_wait_ms=1)
    with pytest.raises(ValueError):
        mb.submit(np.zeros((4, 6, 3), np.float32), np.zeros((4, 6), np.uint16))
    with pytest.raises(ValueError):
        mb.submit(np.zeros((4, 6, 3), np.uint8), np.zeros((5, 6), np.uint16))
    pending = [mb.submit(*_frame(t)) for t in range(1, 6)]
    mb.close()                                                             # drains what was accepted
    assert [int(f.result(1).final_mask[0, 0]) for f in pending] == [1, 2, 3, 4, 5]
    with pytest.raises(RuntimeError):
        mb.submit(*_frame(1))
    mb.close()


def test_two_lanes_serve_everything_and_overlap_batches():
    log, active, peak = [], [0], [0]
    lock = threading.Lock()

    def run(colors, depths):
        with lock:
            active[0] += 1
            peak[0] = max(peak[0], active[0])
            log.append(len(colors))
        time.sleep(0.03)
        with lock:
            active[0] -= 1
        return [FrameResult(None, np.full(c.shape[:2], c[0, 0, 0], np.uint8), float(d[0, 0]), len(colors)) for c, d in zip(colors, depths)]

    with MicroBatcher(run_batch=run, max_batch=4, max_wait_ms=5, lanes=2) as mb:
        out = {}

        def worker(tag):
            out[tag] = int(mb.analyze(*_frame(tag)).final_mask[0, 0])

        threads = [threading.Thread(target=worker, args=(t,)) for t in range(1, 41)]
        for t in threads:
            t.start()
        for t in threads:
            t.join(10)
        pending = [mb.submit(*_frame(t)) for t in range(50, 56)]
    assert out == {t: t for t in range(1, 41)}
    assert [int(f.result(1).final_mask[0, 0]) for f in pending] == list(range(50, 56))      # close() drained both lanes
    assert sum(log) == 46 and max(log) <= 4 and peak[0] == 2                                  # two batches in flight
    with pytest.raises(ValueError):
        MicroBatcher(run_batch=run, lanes=0)


def test_encoded_requests_form_their_own_batches_and_keep_their_payloads():
    """`submit_encoded` takes the request's JPEG / PNG bytes: batches are keyed by image size, never mixed with decoded
    frames, and every caller gets the result of its own payload."""
    cv2 = pytest.importorskip("cv2")
    seen = []

    def run(colors, depths):
        kinds = {isinstance(c, bytes) for c in colors}
        assert len(kinds) == 1
        seen.append((kinds.pop(), len(colors)))
        if isinstance(colors[0], bytes):
            return [FrameResult(None, None, float(len(c)), len(colors)) for c in colors]
        return [FrameResult(None, None, -1.0, len(colors)) for _ in colors]

    def enc(H, W, v):
        return (cv2.imencode(".jpg", np.full((H, W, 3), v, np.uint8))[1].tobytes(),
                cv2.imencode(".png", np.full((H, W), v, np.uint16))[1].tobytes())

    with MicroBatcher(run_batch=run, max_batch=8, max_wait_ms=30) as mb:
        small, big = [enc(8, 8, v) for v in range(3)], [enc(16, 24, v) for v in range(2)]
        futs = [mb.submit_encoded(*p) for p in small] + [mb.submit(*_frame(5))] + [mb.submit_encoded(*p) for p in big]
        res = [f.result(5) for f in futs]
        assert [r.mask_coverage for r in res[:3]] == [float(len(p[0])) for p in small]
        assert res[3].mask_coverage == -1.0
        assert [r.mask_coverage for r in res[4:]] == [float(len(p[0])) for p in big]
        assert sorted(seen) == sorted([(True, 3), (False, 1), (True, 2)])
        with pytest.raises(ValueError):
            mb.submit_encoded(small[0][0], big[0][1])            # colour and depth of different sizes
