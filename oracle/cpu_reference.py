"""cpu_reference.py -- TEST / BASELINE INFRASTRUCTURE ONLY: the reference's CPU path for one clip, timed by
bench.py's `cpu_baseline` leg and `--impl reference` arm and used by tests as the end-to-end checker.

Path per frame pair, as the reference runs it on the host (optical_flow_seg.py:413-461):
  stage 1  frame-mode synapse simulator (simulation/simulationcode_v4_transistor_vehicle.m, numpy restatement in
           stage1_oracle.py -- there is no MATLAB on these boxes) -> current -> uint8 map -> transition_pic ->
           4-connected components -> merged rectangle (FLAG 2)
  stage 2  cv2.calcOpticalFlowFarneback on the crop (the reference's own call; `kind` = "reference"); when cv2
           cannot be imported, the C restatement oracle/farneback_oracle.c (`kind` = "port").
The product path never imports this module.
"""
import os
import sys
import time

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(_HERE))
from oracle import stage1_oracle as s1  # noqa: E402

FARNEBACK_PARAMS = dict(pyr_scale=0.5, levels=3, winsize=15, iterations=3, poly_n=5, poly_sigma=1.2, flags=0)
GRASP = dict(MEMSIZE=80, THRES=250, ext=(20, 20, 20, 20))

try:
    import cv2
    KIND = "reference"
except Exception:  # pragma: no cover
    cv2 = None
    KIND = "port"

_oracle = None


def farneback(prev, nxt, **p):
    global _oracle
    if cv2 is not None:
        return cv2.calcOpticalFlowFarneback(prev, nxt, None, **p)
    if _oracle is None:
        sys.path.insert(0, os.path.join(os.path.dirname(_HERE), "tests"))
        import helpers
        _oracle = helpers.load_oracle()
    return _oracle.farneback(np.ascontiguousarray(prev), np.ascontiguousarray(nxt), **p)


def clip_rects(frames, cfg=GRASP, sim=None, roi_state="scripts"):
    """Stage 1 for one clip (F, H, W) uint8 -> (memimg (F-1, Hc, Wc), rects list of (x0,y0,x1,y1) or None).
    roi_state "scripts": pair t is gated by the array state BEFORE it (slice OFFSET + t of constructed3DMatrix; all
    four scripts overwrite memimg2 with memimg1, optical_flow_seg.py:435); "notebook": by the state after it
    (slice OFFSET + t + 1, demo.ipynb cell 11).  SURVEY.md 3.5."""
    assert roi_state in ("scripts", "notebook")
    sim = sim or {}
    M = cfg["MEMSIZE"]
    H, W = frames.shape[1:]
    cells = np.stack([s1.compress_image(f, M, M) for f in frames])
    _, cur = s1.simulate_memristor_array_vec(cells, **sim)
    memimg = s1.current_to_u8(cur[:-1] if roi_state == "scripts" else cur[1:])  # cur[0] = the initial state
    rects = []
    for t in range(len(frames) - 1):
        n, _, _, r = s1.roi_from_memimg(memimg[t], (H, W), M, M, cfg["THRES"], cfg["ext"], 2)
        rects.append(r[0] if r else None)
    return memimg, rects


def process_clip(frames, cfg=GRASP, params=FARNEBACK_PARAMS, full_frame=False, keep_flow=False, sim=None,
                 roi_state="scripts"):
    """The reference path over one clip.  Returns (pairs, seconds, rects, flows or None)."""
    t0 = time.perf_counter()
    F, H, W = frames.shape
    flows = []
    if full_frame:
        rects = [(0, 0, W, H)] * (F - 1)
    else:
        _, rects = clip_rects(frames, cfg, sim, roi_state)
    for t in range(F - 1):
        flow = np.zeros((H, W, 2))  # optical_flow_seg.py:215
        r = rects[t]
        if r is not None:
            x0, y0, x1, y1 = r
            a, b = frames[t, y0:y1, x0:x1], frames[t + 1, y0:y1, x0:x1]
            if a.size > 0:
                flow[y0:y1, x0:x1] = farneback(a, b, **params)
        if keep_flow:
            flows.append(flow)
    return F - 1, time.perf_counter() - t0, rects, (flows if keep_flow else None)


_POOL_CLIPS = None   # per worker process: the clips, memory-mapped from the pool's scratch directory


def _pool_init(paths):
    global _POOL_CLIPS
    if cv2 is not None:
        cv2.setNumThreads(1)
    _POOL_CLIPS = [np.load(p, mmap_mode="r") for p in paths]


def _worker(args):
    idx, full_frame, roi_state = args
    n, dt, _, _ = process_clip(np.asarray(_POOL_CLIPS[idx]), full_frame=full_frame, roi_state=roi_state)
    return n


class ClipPool:
    """ONE persistent pool of `workers` processes over a fixed list of clips ((F, H, W) uint8 arrays), one clip per
    task -- OpenCV's Farneback core is single-threaded, so this is the CPU path's best case.  Created BEFORE any timed
    region; `run()` is what gets timed.  The workers are SPAWNED (fresh interpreters importing numpy / cv2 only), never
    forked: bench.py's GPU arm calls this from a process that holds a CUDA context and thread pools, and forked
    children of such a process deadlocked in their first cv2 / BLAS call.  The clips reach the workers through .npy files
    in a scratch directory (memory-mapped once per worker, at pool start)."""

    def __init__(self, clips, workers, roi_state="scripts"):
        import multiprocessing as mp
        import tempfile
        self.n, self.workers, self.roi_state = len(clips), max(1, int(workers)), roi_state
        base = "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else None
        self.dir = tempfile.mkdtemp(prefix="nsof_clips_", dir=base)
        self.paths = []
        seen = {}
        for i, c in enumerate(clips):
            if id(c) not in seen:       # the same array listed twice is stored once
                seen[id(c)] = os.path.join(self.dir, f"clip{i}.npy")
                np.save(seen[id(c)], np.ascontiguousarray(c))
            self.paths.append(seen[id(c)])
        self.pool = mp.get_context("spawn").Pool(self.workers, initializer=_pool_init, initargs=(self.paths,))

    def run(self, full_frame=False):
        """-> (pairs processed, wall seconds) for one pass over all clips."""
        t0 = time.perf_counter()
        tasks = [(i, full_frame, self.roi_state) for i in range(self.n)]
        n = sum(self.pool.map(_worker, tasks, chunksize=1))
        return n, time.perf_counter() - t0

    def close(self):
        import shutil
        if self.pool is not None:
            self.pool.close()
            self.pool.join()
            self.pool = None
        shutil.rmtree(self.dir, ignore_errors=True)


def process_clips_parallel(clips, workers, full_frame=False, roi_state="scripts"):
    """One-shot convenience over ClipPool (tests); bench.py keeps one ClipPool across its timed steps."""
    pool = ClipPool(clips, workers, roi_state)
    try:
        return pool.run(full_frame)
    finally:
        pool.close()
