#!/usr/bin/env python
"""bench.py -- ROI-gated Farneback frame-pairs/s at 1080p (BASELINE.json metric), one process per GPU.

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
  python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path on the host cores

Workload (config.workload): BASELINE.json configs[3] -- synthetic 1920x1080 clips, sequence-parallel, no
collective.  Every rank holds `--clips` clips of `--frames` frames resident in HBM (weak scaling: 8 clips per GPU
-> 64 clips at 8 GPUs).  One step = one pass of the hot path over all the rank's clips: stage 1 (compress_image ->
synapse array update -> ROI rectangles) and stage 2 (Farneback on every ROI crop, pasted into zeroed flow frames).
`value` = pairs of all ranks / max-over-ranks device time.  See DESIGN.md "Measurement".
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
_T0 = time.time()


def note(msg):
    """Progress on stderr (stdout carries the one JSON line); NSOF_BENCH_QUIET=1 silences it."""
    if not os.environ.get("NSOF_BENCH_QUIET") and int(os.environ.get("RANK", "0")) == 0:
        print(f"[bench +{time.time() - _T0:6.1f}s] {msg}", file=sys.stderr, flush=True)

METRIC = "roi_gated_farneback_frame_pairs_per_sec_1080p"
UNIT = "frame-pairs/s"
H, W = 1080, 1920


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="nsof", choices=["nsof", "reference"])
    ap.add_argument("--clips", type=int, default=8, help="clips per GPU")
    ap.add_argument("--frames", type=int, default=32, help="frames per clip")
    ap.add_argument("--roi-state", default="scripts", choices=["scripts", "notebook"],
                    help="which array state gates a pair (SURVEY.md 3.5): before it, as the scripts do, or after it")
    ap.add_argument("--min-seconds", type=float, default=1.0,
                    help="repeat the timed block of exactly --steps steps until the device-resident leg has run this long")
    ap.add_argument("--chunks", default="", help="pair boundaries of the step's Farneback batches, e.g. 0,12,31 (default: one batch)")
    ap.add_argument("--s2-streams", type=int, default=1, help="streams the batches of a step alternate over")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the full-frame and per-kernel legs")
    return ap.parse_args()


def workload_config(args, n_gpus):
    """Identical in both arms (`--impl nsof` and `--impl reference`): it names the workload, not how an arm runs it."""
    return {"workload": "configs[3]: synthetic 1920x1080 clips of 32 frames, sequence-parallel (no collective); 64 clips "
                        "over 8 GPUs = 8 clips per GPU, weak-scaled (per-GPU work fixed: 8 clips at every N)",
            "frame": [H, W], "clips_per_gpu": args.clips, "frames_per_clip": args.frames,
            "pairs_per_step_per_gpu": args.clips * (args.frames - 1), "clips_total": args.clips * n_gpus,
            "roi": "synapse-array gate (MEMSIZE 80, THRES 250, EXT 20, FLAG 2)", "roi_state": args.roi_state,
            "farneback": [0.5, 3, 15, 3, 5, 1.2, 0],
            "stage1": "frame mode, nSubSteps 1000 (parity unpinned: no MATLAB; held to independent 50-digit fixtures)",
            "l2": "inputs exceed L2 (frames + flow frames per step >> 126 MB)",
            "parallelism": f"sequence-parallel x{n_gpus}"}


GPU_SCHEDULE = ("DeviceStream (pipeline.py): one Farneback batch of all pairs per step on the caller's stream; stage 1 of "
                "step k+1 on a high-priority side stream beside it; every timed block = exactly K steps submitted and "
                "drained between two events")


# ------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "20"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1]))
                mx.append(float(c[2]))
            except ValueError:
                continue
            for nm, v in zip(names, c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        try:
            os.unlink(self.f.name)
        except OSError:
            pass
        if sm:
            sm.sort()
            out.update(sm_mhz=sm[len(sm) // 2], sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm))
        return out


# ------------------------------------------------------------------------------------------- CPU legs
def host_clips(n_clips, frames, seed0=1234):
    import importlib
    syn = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.synthetic")
    return [syn.make_clip(seed0 + c, H, W, frames, "cpu").numpy() for c in range(n_clips)]


def _host_workers():
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    return max(1, cores)


def _ref_what(ref):
    return f"numpy stage 1 + {'cv2 ' + ref.cv2.__version__ if ref.cv2 is not None else 'C oracle'} Farneback on the ROI crop"


def cpu_baseline_leg(args, clip_np):
    """Bounded sample on rank 0: the reference CPU path on all host cores (one persistent process per core, one clip of
    the workload's own length per task -- OpenCV's Farneback core is single-threaded), plus the same path in ONE
    process for the per-core figure."""
    from oracle import cpu_reference as ref
    workers = _host_workers()
    clips = host_clips(min(workers, 16), args.frames)
    clips = [clips[i % len(clips)] for i in range(workers)]
    pool = ref.ClipPool(clips, workers, args.roi_state)  # created (and warmed) before the timed pass
    pool.run()
    pairs, dt_all = pool.run()
    pool.close()
    short = clip_np[:min(len(clip_np), 6)]
    ref.process_clip(short[:3], roi_state=args.roi_state)
    n1, dt1, _, _ = ref.process_clip(short, roi_state=args.roi_state)
    nf, dtf, _, _ = ref.process_clip(short[:3], full_frame=True)
    return {"value": pairs / dt_all, "unit": UNIT, "cores": workers, "kind": ref.KIND,
            "sample": f"one pass over {workers} clips x {args.frames} frames = {pairs} pairs of the same synthetic 1080p "
                      f"workload (roi_state {args.roi_state}), one persistent process per host core, {_ref_what(ref)}",
            "single_process_value": n1 / dt1, "single_process_sample": f"{n1} pairs (first {len(short)} frames of clip 0)",
            "single_process_full_frame_value": nf / dtf}


def reference_arm(args):
    """--impl reference: the reference CPU path on all host cores, on the GPU arm's config.  One step = one pass over
    one clip (of the workload's own `--frames`) per host core; the process pool is persistent and created before the
    timed loop."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from oracle import cpu_reference as ref
    workers = _host_workers()
    clips = host_clips(min(workers, 16), args.frames)
    clips = [clips[i % len(clips)] for i in range(workers)]
    pool = ref.ClipPool(clips, workers, args.roi_state)
    for _ in range(max(1, min(args.warmup, 1))):
        pool.run()
    t0 = time.perf_counter()
    pairs = 0
    for _ in range(args.steps):
        n, _ = pool.run()
        pairs += n
    dt = time.perf_counter() - t0
    pool.close()
    val = pairs / dt
    sample = (f"each step = {workers} clips x {args.frames} frames ({workers * (args.frames - 1)} pairs) of the same "
              f"synthetic 1080p workload (roi_state {args.roi_state}), one persistent process per host core, {_ref_what(ref)}")
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, args.gpus),
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": workers, "kind": ref.KIND, "sample": sample},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


def parity_check(args, clip_u8, rects, n_rects, flow_of_pair, n_pairs=8):
    """Outside every timed region: the first `n_pairs` pairs of clip 0 of the timed workload through the reference CPU
    path (stage 1 is a recurrence from frame 0, so a prefix of the clip reproduces the prefix of the rectangles) against
    what the timed GPU step left in HBM."""
    import numpy as np
    from oracle import cpu_reference as ref
    F = min(len(clip_u8), n_pairs + 1)
    _, _, want_rects, flows = ref.process_clip(clip_u8[:F], keep_flow=True, roi_state=args.roi_state)
    ok, epe_sum, epe_max, epe_max_in, px, errs = True, 0.0, 0.0, 0.0, 0, []
    BAND = 24
    for t in range(F - 1):
        got = tuple(int(v) for v in rects[t]) if n_rects[t] else None
        ok = ok and (got == want_rects[t])
        if got is None or got != want_rects[t]:
            continue
        x0, y0, x1, y1 = got
        d = np.sqrt(((flow_of_pair(t)[y0:y1, x0:x1].astype(np.float64) - flows[t][y0:y1, x0:x1]) ** 2).sum(-1))
        epe_sum, epe_max, px = epe_sum + d.sum(), max(epe_max, float(d.max())), px + d.size
        if d.shape[0] > 2 * BAND and d.shape[1] > 2 * BAND:
            epe_max_in = max(epe_max_in, float(d[BAND:-BAND, BAND:-BAND].max()))
        errs.append(d.ravel()[::7])
    q999 = float(np.quantile(np.concatenate(errs), 0.999)) if errs else 0.0
    # epe_max sits on isolated near-singular pixels of the smooth synthetic background (cv2 vs the scalar C restatement
    # of the same algorithm differs there just as much; tests/test_gpu_parity.py states the per-pixel bounds)
    # ... and near the BORDER of a crop: on its last row / column OpenCV's UpdateMatrices switches between its bilinear branch
    # and its out-of-image fallback on the sign of a displacement of ~1e-7 px (y + dy >= h - 1) -- a discontinuity of the
    # algorithm itself, any two roundings of it land on different sides -- and the 15 x 15 window sums carry a flipped
    # pixel 7 px further per iteration (3 iterations: a band of ~24 px; the emulator build shows the same pixels)
    return {"pairs": F - 1, "rects_equal": bool(ok), "epe_mean": float(epe_sum / max(px, 1)), "epe_q999": q999,
            "epe_max": epe_max, "epe_max_excluding_a_24px_border_band": epe_max_in,
            "against": _ref_what(ref), "bar": "BASELINE.json: EPE <= 0.01 px, masks bit-exact"}


# ------------------------------------------------------------------------------------------- GPU arm
def gpu_arm(args):
    import importlib

    import torch
    import torch.distributed as dist

    par = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.parallel")
    rank, world, local = par.rank_info()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; libnsof has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    # stdout carries the one JSON line: NCCL's log (whatever level the environment asks for) goes to stderr
    os.environ.setdefault("NCCL_DEBUG", "WARN")
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    par.init("nccl", dev)  # plumbing only: barrier + max-over-ranks of the time; clips shard with no collective

    nsof = importlib.import_module("nsof_b200")
    syn = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.synthetic")
    pipe_mod = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.pipeline")
    eng = nsof.Engine(local)
    Cn, F = args.clips, args.frames
    cb = [int(v) for v in args.chunks.split(",")] if args.chunks else [0, F - 1]
    assert cb[0] == 0 and cb[-1] == F - 1 and cb == sorted(cb), "--chunks must run from 0 to frames - 1"
    pipe = pipe_mod.DeviceStream(eng, Cn, F, H, W, depth=2, roi_state=args.roi_state,
                                 chunks=list(zip(cb[:-1], cb[1:])), s2_streams=args.s2_streams)

    # synthetic clips, generated on the device, layout [F][C][H][W]
    frames = torch.empty((F, Cn, H, W), dtype=torch.uint8, device=dev)
    for i, c in enumerate(par.weak_scaling_clips(Cn, rank, world)):  # clip c -> GPU c mod G
        frames[:, i] = syn.make_clip(par.clip_seed(c), H, W, F, dev)
    torch.cuda.synchronize()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup, count=None):
        count = count or eng.launch_count
        for _ in range(warmup):
            fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = count()
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), count() - l0

    pairs_step = pipe.n_pairs
    note(f"clips on the device ({Cn} x {F} frames); value leg")
    # ---- value: device-resident step; the clock sampler runs from here to the end of the e2e leg
    clk = ClockSampler(local)
    if rank == 0:
        clk.start()
    def run_steps(k):  # k steps through the stream; the last one is finished (drained) before returning
        for _ in range(k):
            pipe.submit(frames)
        pipe.drain()

    run_steps(args.warmup)
    # the timed block is EXACTLY K steps between two events; it is repeated back to back (each block its own barrier +
    # synchronize on both sides) until the leg has run --min-seconds, so the clock record has samples under load.
    # value = pairs of all blocks / summed max-over-ranks block times; ms_per_step = that time / (blocks x K).
    block_ms, launches = [], 0
    while True:
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = pipe.launch_count()
        e0.record()                # stage 1 of the first step is ordered behind this event (side stream waits for it)
        run_steps(args.steps)      # exactly K steps, every Farneback batch on this stream
        e1.record()
        barrier()
        if not block_ms:
            launches = pipe.launch_count() - l0
        ms_t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms_t, op=dist.ReduceOp.MAX)
        block_ms.append(float(ms_t.item()))
        if sum(block_ms) >= args.min_seconds * 1e3 or len(block_ms) >= 200:   # every rank sees the same reduced times
            break
    ms = sum(block_ms) / len(block_ms)     # mean K-step block
    value = pairs_step * world * args.steps / (ms * 1e-3)
    roi_frac = pipe.slots[0].last_crop_pixels / (pairs_step * H * W)
    crop_px_step = pipe.slots[0].last_crop_pixels

    # ---- e2e: pinned host clips in, rectangles + packed ROI flow crops out, copies inside the timed region
    note(f"value leg done: {len(block_ms)} blocks of {args.steps} steps, {ms / args.steps:.3f} ms per step; parity check")
    # parity of what was just timed, outside the timed region (rank 0): first pairs of clip 0 vs the reference CPU path
    parity = None
    if rank == 0:
        sl = pipe.slots[(pipe.k - 1) % len(pipe.slots)]
        torch.cuda.synchronize()
        r_all, n_all = sl.last_rects
        parity = parity_check(args, frames[:, 0].cpu().numpy(), r_all.reshape(F - 1, Cn, 4)[:, 0],
                              n_all.reshape(F - 1, Cn)[:, 0], lambda t: sl.flow[t, 0].cpu().numpy())
    note(f"parity {parity}; e2e leg")
    hp = pipe_mod.HostStream(local, Cn, F, H, W, depth=3, roi_state=args.roi_state)
    clips_host = [torch.empty((F, H, W), dtype=torch.uint8).pin_memory() for _ in range(Cn)]
    for c in range(Cn):
        clips_host[c].copy_(frames[:, c])
    torch.cuda.synchronize()

    def timed_host(steps, warmup):
        for _ in range(warmup):
            hp.submit(clips_host)
        hp.drain()
        torch.cuda.synchronize()
        barrier()
        t0 = time.perf_counter()  # several streams: wall clock between device-wide synchronisations
        got = 0
        for _ in range(steps):    # every step: H2D of its clips, D2H of its rectangles + flow crops, results collected
            got += hp.submit(clips_host) is not None
        got += len(hp.drain())    # pipeline fill and drain are inside the timed region
        torch.cuda.synchronize()
        assert got == steps
        ms_ = torch.tensor([(time.perf_counter() - t0) * 1e3], dtype=torch.float64, device=dev)
        barrier()
        if world > 1:
            dist.all_reduce(ms_, op=dist.ReduceOp.MAX)
        return float(ms_.item())

    ms_e2e = timed_host(args.steps, 3)
    clocks = clk.stop() if rank == 0 else None  # sampled every 20 ms across both timed regions (value and e2e)
    s0 = hp.slots[0]

    def copy_floor(steps):
        """The same bytes over the same pinned buffers and copy streams with NO kernels: what the host<->device path of
        this box allows for the e2e leg (all ranks copy at once, as in the leg itself)."""
        lib, fsz = eng.lib, H * W
        px = s0.pipe.last_crop_pixels
        n_ch = len(s0.pipe.chunks)
        barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            for j, (p0, p1) in enumerate(s0.pipe.chunks):
                f0 = 0 if j == 0 else p0 + 1
                for c in range(Cn):
                    lib.check(lib.nsof_copy_2d_async(s0.frames.data_ptr() + (f0 * Cn + c) * fsz, Cn * fsz,
                                                     clips_host[c].data_ptr() + f0 * fsz, fsz, fsz, p1 + 1 - f0, 1,
                                                     s0._ptr(s0.s_in)), "copy floor H2D")
                a, b = 2 * px * j // n_ch, 2 * px * (j + 1) // n_ch
                with torch.cuda.stream(s0.s_out):
                    s0.flow_host[a:b].copy_(s0.pipe.packed[a:b], non_blocking=True)
        torch.cuda.synchronize()
        ms_ = torch.tensor([(time.perf_counter() - t0) * 1e3], dtype=torch.float64, device=dev)
        barrier()
        if world > 1:
            dist.all_reduce(ms_, op=dist.ReduceOp.MAX)
        return float(ms_.item()) / steps

    copy_floor(2)
    floor_ms = copy_floor(max(5, args.steps // 2))
    e2e = {"value": pairs_step * world * args.steps / (ms_e2e * 1e-3), "unit": UNIT,
           "h2d_bytes_per_step": int(hp.h2d_bytes), "d2h_bytes_per_step": int(hp.d2h_bytes),
           "ms_per_step": ms_e2e / args.steps, "copy_floor_ms_per_step": floor_ms,
           "copy_floor_GBps_per_gpu": (hp.h2d_bytes + hp.d2h_bytes) / (floor_ms * 1e-3) / 1e9,
           "frac_of_copy_floor": floor_ms / (ms_e2e / args.steps),
           "api": "HostStream.submit/drain (pinned uint8 clips -> rectangles + packed float32 ROI flow on the host), "
                  f"{len(hp.slots)} steps in flight, each on {len(s0.slots)} streams (H2D / stage 1 / stage 2 / D2H) "
                  f"in {len(s0.pipe.chunks)} time chunks; all {args.steps} steps submitted and drained inside the "
                  "timed region, host clock between device synchronisations"}
    del hp
    note(f"e2e leg done: {ms_e2e / args.steps:.3f} ms per step; per-kernel leg")
    return finish(args, locals())


def dropin_call_leg(eng):
    """extras["dropin_call"]: what a maintainer gets after the two-line edit of INTEGRATION.md -- the synchronous
    per-pair call `calcOpticalFlowFarneback(prev, next, None, **farneback_params)` on numpy arrays (pageable host
    memory in, a fresh numpy flow out), median of 15 calls, beside cv2's own time on the same arrays."""
    import numpy as np
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import helpers
    nsof = __import__("nsof_b200")
    try:
        import cv2
    except Exception:
        cv2 = None
    out = {}
    for name, (h, w) in (("crop_600x600", (600, 600)), ("frame_1080x1920", (1080, 1920))):
        a, b = helpers.textured_pair(5, h, w)
        ts = []
        for _ in range(18):
            t0 = time.perf_counter()
            f = eng.calcOpticalFlowFarneback(a, b, None, **nsof.FARNEBACK_PARAMS)
            ts.append((time.perf_counter() - t0) * 1e3)
        row = {"nsof_ms_median": float(np.median(ts[3:])), "nsof_ms_min": float(min(ts[3:]))}
        if cv2 is not None:
            tc = []
            for _ in range(3):
                t0 = time.perf_counter()
                g = cv2.calcOpticalFlowFarneback(a, b, None, **nsof.FARNEBACK_PARAMS)
                tc.append((time.perf_counter() - t0) * 1e3)
            d = np.sqrt(((f.astype(np.float64) - g) ** 2).sum(-1))
            row.update(cv2_ms_median=float(np.median(tc)), speedup=float(np.median(tc) / row["nsof_ms_median"]),
                       epe_mean=float(d.mean()), epe_max=float(d.max()))
        out[name] = row
    return out


def param_sets_leg(eng, frames, timed):
    """extras["param_sets"]: full-frame 1080p pairs/s and fraction of the HBM peak for the two OTHER parameter sets the
    reference ships (optical_flow_seg.py:61-112, data/*/Parameters.txt): they run on the generic kernels."""
    import importlib
    nsof = importlib.import_module("nsof_b200")
    pipe_mod = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.pipeline")
    peak, _ = measured_peak()
    out = {}
    F = 9
    fr = frames[:F, :1].contiguous()
    for name, p in (("autodriving_txt", dict(pyr_scale=0.6, levels=3, winsize=3, iterations=3, poly_n=10, poly_sigma=1.05, flags=0)),
                    ("tabletennis", dict(pyr_scale=0.6, levels=3, winsize=4, iterations=2, poly_n=1, poly_sigma=1.05, flags=0))):
        cfg = nsof.RoiConfig(farneback_params=p)
        pf = pipe_mod.SequencePipeline(eng, 1, F, H, W, cfg, full_frame=True)
        ms, _ = timed(lambda: pf.step(fr), 5, 2)
        pairs = 5 * pf.n_pairs
        bytes_pair = eng.algorithmic_bytes(H, W, **p)
        out[name] = {"params": list(p.values()), "value": pairs / (ms * 1e-3), "unit": UNIT,
                     "algorithmic_MB_per_pair": bytes_pair / 1e6,
                     "frac_of_peak": pairs * bytes_pair / (ms * 1e-3) / 1e9 / peak}
        del pf
    return out


def tiled_8k_leg(args, rank, world, local, dev, barrier, steps=20, pairs=8, halo=32):
    """extras["tiled_8k"]: BASELINE.json configs[4] -- synthetic 8K (7680 x 4320) frame pairs, each split by rows over
    ALL ranks (strong scaling), the level loop in one library call per batch with the halo rows stored straight into
    the row neighbours' workspaces over NVLink peer memory (tiled.TiledFarneback.run_linked).  Before timing, rank
    0 runs the same pairs untiled on its own GPU and every rank compares its band bit for bit."""
    import importlib

    import torch
    import torch.distributed as dist
    nsof = importlib.import_module("nsof_b200")
    syn = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.synthetic")
    tiled = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.tiled")
    H8, W8 = 4320, 7680
    P = nsof.FARNEBACK_PARAMS
    eng = nsof.Engine(local)
    frames = syn.make_clip(999, H8, W8, pairs + 1, dev, velocity=(6.0, -3.0))   # the same frames on every rank
    flow = torch.zeros((pairs, H8, W8, 2), dtype=torch.float32, device=dev)
    crops = lambda out: [(frames[t].data_ptr(), frames[t + 1].data_ptr(), H8, W8, W8, W8, out[t].data_ptr(), W8 * 8)
                         for t in range(min(pairs, out.shape[0]))]
    tf = tiled.TiledFarneback(eng, rank, world, halo_r1=halo)
    tf.run_linked(crops(flow), **P)
    torch.cuda.synchronize()
    # bit-identity of every rank's band against an untiled run (two pairs are enough to hold 8K workspaces twice)
    lo, hi = tiled.band_rows(H8, rank, world)
    ref_eng = nsof.Engine(local)
    ref = torch.empty((2, H8, W8, 2), dtype=torch.float32, device=dev)
    ref_eng.farneback_batch_device(crops(ref), torch.cuda.current_stream().cuda_stream, **P)
    torch.cuda.synchronize()
    ok = torch.tensor([int(torch.equal(flow[:2, lo:hi], ref[:, lo:hi]))], device=dev)
    if world > 1:
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    ref_eng.close()
    del ref
    torch.cuda.empty_cache()
    for _ in range(2):
        tf.run_linked(crops(flow), check_guard=False, **P)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = eng.launch_count()
    e0.record()
    for _ in range(steps):
        tf.run_linked(crops(flow), check_guard=False, **P)
    e1.record()
    barrier()
    launches = eng.launch_count() - l0
    tf._check_guard()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_step = float(ms.item()) / steps
    bytes_pair = eng.algorithmic_bytes(H8, W8)
    peak, _ = measured_peak()
    out = {"metric": "tiled_farneback_frame_pairs_per_sec_8k", "value": pairs / (ms_step * 1e-3), "unit": UNIT,
           "n_gpus": world, "scaling": "strong", "steps": steps, "ms_per_step": ms_step, "pairs_per_step": pairs,
           "frame": [H8, W8], "halo_r1": halo, "bands_bit_identical_to_untiled": bool(ok.item()),
           "launches_per_step": launches // steps,
           "algorithmic_GBps_per_gpu": pairs * bytes_pair / world / (ms_step * 1e-3) / 1e9,
           "frac_of_hbm_peak_per_gpu": pairs * bytes_pair / world / (ms_step * 1e-3) / 1e9 / peak,
           "exchange": "one store-to-peer kernel + one flag-wait kernel per stage (NVLink peer memory, CUDA IPC); no NCCL "
                       "on the data path, one host call per batch" if world > 1 else "none (one GPU)"}
    eng.close()
    del frames, flow
    torch.cuda.empty_cache()
    return out


# kernel class (nsof_profile rows) -> __global__ function
KERNEL_FUNCTION = {"fb_blur_um": "k_fb_blur_solve_t", "fb_blur_flow": "k_fb_blur_solve_t",
                   "fb_pyr_polyexp": "k_fb_pyr_polyexp_t", "fb_um_first": "k_fb_um_first", "fb_pyr": "k_fb_pyr_sep",
                   "fb_polyexp": "k_fb_polyexp", "s1_frame_sim": "k_s1_frame_run", "s1_lanczos": "k_s1_lanczos_rows+cols",
                   "s1_roi": "k_s1_roi", "fb_flow_clear": "k_fb_flow_clear", "table_upload": "k_table_upload",
                   "halo_copy": "k_copy_blocks"}


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic(kernel):
    """Per-launch DRAM traffic of `kernel` from the committed ncu --set full summary, if one exists."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            return json.load(f)[kernel]["dram_bytes_per_launch"]
    except Exception:
        return None


def finish(args, L):
    import importlib

    import torch
    eng, pipe, frames, rank, world, dev = L["eng"], L["pipe"], L["frames"], L["rank"], L["world"], L["dev"]
    timed, Cn, F = L["timed"], L["Cn"], L["F"]
    peak, peak_src = measured_peak()
    extras = {}
    # ---- per-kernel leg: CUDA event pair around every launch of a few extra steps (not part of `value`)
    def kernel_tables(step_fn, nsteps=2):
        eng.profile(True)
        for _ in range(nsteps):
            step_fn()
        eng.profile(False)
        st = eng.profile_read()
        tot = sum(s["ms"] for s in st.values()) or 1.0
        rows = {k: {"launches": s["launches"] // nsteps, "ms_per_step": round(s["ms"] / nsteps, 4),
                    "share": round(s["ms"] / tot, 4), "algorithmic_MB_per_step": round(s["algorithmic_bytes"] / nsteps / 1e6, 3),
                    "achieved_GBps": round(s["algorithmic_bytes"] / (s["ms"] * 1e-3) / 1e9, 1) if s["ms"] > 0 else None}
                for k, s in st.items()}
        cls = {}
        for k, s in st.items():
            c = cls.setdefault(k.split(".")[0], dict(launches=0, ms=0.0, algorithmic_bytes=0.0))
            for f in c:
                c[f] += s[f]
        return rows, cls

    # per-kernel leg on a serialised copy of the step (stage 1 on the same stream: no overlap between the event pairs)
    pipe_mod = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.pipeline")
    pipe_ser = pipe_mod.SequencePipeline(eng, Cn, F, H, W, overlap=False, chunks=[(0, F - 1)],
                                         roi_state=args.roi_state)  # one batch per step, as timed
    pipe_ser.step(frames)
    kernels, stats = kernel_tables(lambda: pipe_ser.step(frames))
    del pipe_ser
    # dominant kernel = the __global__ function with the most device time in the step (fb_blur_um and fb_blur_flow
    # are the two modes of k_fb_blur_solve_t); achieved = its algorithmic bytes / its time, summed over its launches
    fn_of = lambda cls: KERNEL_FUNCTION.get(cls, cls)
    per_fn = {}
    for cls, srow in stats.items():
        a = per_fn.setdefault(fn_of(cls), dict(launches=0, ms=0.0, algorithmic_bytes=0.0, classes=[]))
        a["launches"] += srow["launches"]; a["ms"] += srow["ms"]; a["algorithmic_bytes"] += srow["algorithmic_bytes"]
        a["classes"].append(cls)
    dom = max(per_fn, key=lambda k: per_fn[k]["ms"])
    d = per_fn[dom]
    ach = d["algorithmic_bytes"] / (d["ms"] * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": dom, "classes": sorted(d["classes"]), "achieved": ach, "peak": peak,
                "unit": "GB/s", "frac": ach / peak, "traffic": ncu_traffic(dom), "peak_source": peak_src,
                "algorithmic_bytes_per_launch": d["algorithmic_bytes"] / d["launches"],
                "avg_launch_ms": d["ms"] / d["launches"], "launches_per_step": d["launches"] // 2,
                "share_of_step_kernel_time": d["ms"] / sum(x["ms"] for x in per_fn.values())}
    fb_bytes = sum(s["algorithmic_bytes"] for k, s in stats.items() if k.startswith("fb_")) / 2
    fb_ms = sum(s["ms"] for k, s in stats.items() if k.startswith("fb_")) / 2
    extras["stage2_all_kernels"] = {"algorithmic_GB_per_step": fb_bytes / 1e9, "ms_per_step": fb_ms,
                                    "achieved_GBps": fb_bytes / (fb_ms * 1e-3) / 1e9 if fb_ms else None,
                                    "frac_of_peak": fb_bytes / (fb_ms * 1e-3) / 1e9 / peak if fb_ms else None}
    note("per-kernel leg done; full-frame leg")
    # ---- ROI = 100 % leg (full-frame Farneback, the configuration the >= 50 % roofline target is stated on)
    if not args.no_extras:
        for sl in pipe.slots:
            del sl.flow
        torch.cuda.empty_cache()
        ff_clips = min(Cn, 4)
        pf = pipe_mod.SequencePipeline(eng, ff_clips, F, H, W, full_frame=True)
        ff_frames = frames[:, :ff_clips].contiguous()
        ms_ff, _ = timed(lambda: pf.step(ff_frames), max(2, args.steps // 2), 2)
        n_ff = max(2, args.steps // 2) * pf.n_pairs
        bytes_pair = eng.algorithmic_bytes(H, W)
        extras["full_frame"] = {"value": n_ff * world / (ms_ff * 1e-3), "unit": UNIT,
                                "algorithmic_MB_per_pair": bytes_pair / 1e6,
                                "achieved_GBps_per_gpu": n_ff * bytes_pair / (ms_ff * 1e-3) / 1e9,
                                "frac_of_peak": n_ff * bytes_pair / (ms_ff * 1e-3) / 1e9 / peak,
                                "pairs_per_step_per_gpu": pf.n_pairs}
        extras["full_frame"]["kernels"] = kernel_tables(lambda: pf.step(ff_frames))[0]
    if not args.no_extras:
        if rank == 0 and world == 1:
            note("drop-in call and parameter-set legs")
            extras["dropin_call"] = dropin_call_leg(eng)
            extras["param_sets"] = param_sets_leg(eng, frames, timed)
        for sl in pipe.slots:   # release the device-resident pipelines before the 8K workspaces are built
            for e in [sl.eng, sl.s1_eng] + [e for e, _ in sl.s2_extra]:
                e.close()
        del pf
        torch.cuda.empty_cache()
        note("tiled 8K leg")
        extras["tiled_8k"] = tiled_8k_leg(args, rank, world, L["local"], dev, L["barrier"])
    if rank != 0:
        return 0
    line = {"metric": METRIC, "value": L["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": L["ms"] / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload_config(args, world),
            "clocks": L["clocks"], "e2e": L["e2e"], "gpu_launches": int(L["launches"]), "roofline": roofline,
            "timed_blocks": len(L["block_ms"]), "block_ms_min_max": [min(L["block_ms"]), max(L["block_ms"])],
            "schedule": GPU_SCHEDULE, "parity_check": L["parity"],
            "roi_fraction_of_frame": L["roi_frac"], "kernels": kernels, **extras}
    if world == 1 and not args.no_cpu_baseline:
        note("cpu_baseline leg")
        line["cpu_baseline"] = cpu_baseline_leg(args, frames[:, 0].cpu().numpy())
    note("done")
    print(json.dumps(line), flush=True)
    return 0


def main():
    args = parse_args()
    wd = float(os.environ.get("NSOF_BENCH_WATCHDOG", "0") or 0)
    if wd > 0:  # dump every thread's Python stack to stderr every `wd` seconds (diagnosing a stuck leg)
        import faulthandler
        faulthandler.dump_traceback_later(wd, repeat=True, file=sys.stderr)
    if args.impl == "reference":
        return reference_arm(args)
    rc = gpu_arm(args)
    try:
        import torch.distributed as dist
        if dist.is_initialized():
            dist.destroy_process_group()
    except Exception:
        pass
    return rc


if __name__ == "__main__":
    sys.exit(main())
