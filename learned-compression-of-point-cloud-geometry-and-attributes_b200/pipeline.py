"""Frame pipeline: several frames in flight on one GPU.

Two schedulers over the same per-frame generators (results are identical, frame by frame):
``threads`` (default) gives every in-flight frame its own launch thread and CUDA stream — the blocking
points of a frame (point counts read back by the coordinate kernels, the coder) release the GIL, so
they stall that frame only; ``cooperative`` resumes all frames round-robin from ONE thread and is
kept for comparison (a count read-back of one frame holds up the launches of the others: measured
55 ms of blocked launch thread per 56 ms of kernels per frame in the tensor-core mode).


One frame's encode+decode is a dependency chain with two long CPU-only links — the reference's
sequential rANS coder (model/entropy_models.py:372,408; ~2.3 M symbols per direction at 800 k
points) — during which the GPU has nothing to do for that frame.  Frames are independent
(evaluate.py:95-110 codes them one at a time), so a sequence (BASELINE.json configs[3]) is coded with
`depth` frames in flight: each frame is a generator (ColorModel.compress_gen / .decompress_gen) on
its own CUDA stream that yields where it waits for a device->host copy or for the coder (which runs
in compressai_compat's coder threads, GIL released).  The scheduler resumes whichever frame is
ready, oldest first, so while one frame sits in the coder the other frames' kernels keep the SMs
busy.  A single Python thread issues all launches: no GIL hand-offs between launch threads.
No frame's result depends on the interleaving: every kernel of a frame is enqueued on that frame's
stream in program order.
"""
import concurrent.futures as cf
import threading
import time
from collections import deque

import torch

from .compressai_compat import Wait, _pinned, _pinned_release, drive


class _Frame:
    __slots__ = ("index", "gen", "stream", "wait", "started")

    def __init__(self, index, gen, stream):
        self.index, self.gen, self.stream, self.wait, self.started = index, gen, stream, None, False


class FramePipeline:
    def __init__(self, model, depth=3, device=None, scheduler="threads"):
        self.model = model
        self.depth = max(1, int(depth))
        self.device = torch.device(device) if device is not None else model.device
        self._streams = [torch.cuda.Stream(device=self.device) for _ in range(self.depth)]
        self.stats = {}  # host-side accounting of the last roundtrip_many call
        if scheduler not in ("threads", "cooperative"):
            raise ValueError(scheduler)
        self.scheduler = scheduler
        self._pool = cf.ThreadPoolExecutor(max_workers=self.depth, thread_name_prefix="pcc-frame") \
            if scheduler == "threads" and self.depth > 1 else None

    def _roundtrip_gen(self, x, Q, host_io):
        m = self.model
        strings, shape, k, coords = yield from m.compress_gen(x, Q)
        if host_io:
            coords = coords.cpu()  # the latent coordinates leave the device with the strings
        out = yield from m.decompress_gen(None, coords, strings, shape, k)
        if host_io:
            h = _pinned(out.shape, out.dtype)
            h.copy_(out, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record()
            yield Wait(event=ev)
            out = h.clone()
            _pinned_release(h)
        return out, strings, coords

    def _worker(self, stream, ready, work, frames, results, host_io, busy):
        """One launch thread: codes frames from the shared queue on its own stream until it is empty."""
        torch.cuda.set_device(self.device)  # the current device and the grad mode are per thread
        done = None
        with torch.no_grad(), torch.cuda.stream(stream):
            stream.wait_event(ready)
            while True:
                try:
                    i = work.popleft()  # deque.popleft is atomic
                except IndexError:
                    break
                t0 = time.perf_counter()
                x, Q = frames[i]
                results[i] = drive(self._roundtrip_gen(x, Q, host_io))
                busy.append((time.perf_counter() - t0) * 1e3)
            done = torch.cuda.Event()
            done.record(stream)
        return done

    def _roundtrip_threads(self, frames, host_io):
        cur = torch.cuda.current_stream(self.device)
        ready = torch.cuda.Event()
        ready.record(cur)
        work, results, busy = deque(range(len(frames))), [None] * len(frames), []
        n = min(self.depth, len(frames))
        futs = [self._pool.submit(self._worker, self._streams[j], ready, work, frames, results, host_io, busy)
                for j in range(n)]
        for f in futs:
            done = f.result()  # re-raises a worker's exception
            if done is not None:
                cur.wait_event(done)  # the caller's stream sees every frame's results
        self.stats = {"scheduler": "threads", "launch_threads": n, "frames": len(frames),
                      "frame_wall_ms_mean": sum(busy) / max(len(busy), 1), "frame_wall_ms_max": max(busy, default=0.0)}
        return results

    @torch.no_grad()
    def roundtrip_many(self, frames, host_io=False):
        """frames: list of (x, Q) -> list of (decoded cloud, strings, latent coordinates), in order.
        Inputs may live on the host (H2D happens on the frame's stream) or on the device; with
        host_io the latent coordinates and the decoded cloud are returned in host memory."""
        if self._pool is not None and len(frames) > 1:
            return self._roundtrip_threads(frames, host_io)
        cur = torch.cuda.current_stream(self.device)
        ready = torch.cuda.Event()
        ready.record(cur)
        pending = deque(enumerate(frames))
        free = list(self._streams)
        active, results = [], [None] * len(frames)
        st = self.stats = {"resumes": 0, "resume_ms": 0.0, "max_resume_ms": 0.0, "idle_ms": 0.0}
        while pending or active:
            while pending and free:
                i, (x, Q) = pending.popleft()
                s = free.pop()
                s.wait_event(ready)
                active.append(_Frame(i, self._roundtrip_gen(x, Q, host_io), s))
            progressed = False
            for f in list(active):
                if f.started and not f.wait.ready():
                    continue
                progressed = True
                t0 = time.perf_counter()
                with torch.cuda.stream(f.stream):
                    try:
                        f.wait = f.gen.send(None)
                        f.started = True
                    except StopIteration as e:
                        results[f.index] = e.value
                        done = torch.cuda.Event()
                        done.record(f.stream)
                        cur.wait_event(done)  # the caller's stream sees this frame's results
                        active.remove(f)
                        free.append(f.stream)
                dt = (time.perf_counter() - t0) * 1e3
                st["resumes"] += 1
                st["resume_ms"] += dt
                st["max_resume_ms"] = max(st["max_resume_ms"], dt)
                break  # rescan from the oldest frame after every step
            if not progressed:
                t0 = time.perf_counter()
                time.sleep(5e-5)
                st["idle_ms"] += (time.perf_counter() - t0) * 1e3
        return results

    def close(self):
        if self._pool is not None:
            self._pool.shutdown(wait=True)
            self._pool = None
