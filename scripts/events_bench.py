"""BASELINE.json configs[2] measurement: event stream -> synapse-array memory (scheme 1 / scheme 2 of
eventsim/event_mem_sim.py) on the device, next to the numpy loop of the reference (oracle/stage1_oracle.py restates it
bit for bit) on a bounded sample.  driving_data.hdf5 is a 134-byte LFS pointer and h5py is absent, so the stream is the
synthetic moving box (reference size 240x320, and the scaled 1080x1920 variant of SURVEY.md 8(d)).
usage: python scripts/events_bench.py"""
import ctypes as C
import importlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

nsof = importlib.import_module("nsof_b200")
syn = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.synthetic")
eng_mod = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.engine")
from oracle import stage1_oracle as s1  # noqa: E402  (CPU baseline leg only)

eng = nsof.Engine(0)
lib, h = eng.lib, eng.handle
sp = eng_mod._synapse_params(nsof.PARAMS)
st = lambda: C.c_void_p(torch.cuda.current_stream().cuda_stream)


def run(name, H, W, box, speed, slice_us=1000, cpu_slices=60):
    x, y, p, t = syn.synthetic_events(H, W, box, speed, 1.5)
    bounds = nsof.slice_bounds(t, slice_us).astype(np.int64)
    n_slices = len(bounds) - 1
    d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    dx, dy, dp, dt_, db = d(x.astype(np.int16)), d(y.astype(np.int16)), d(p.astype(np.int8)), d(t), d(bounds)
    out = {"frame": [H, W], "events": int(len(x)), "slices": int(n_slices), "slice_us": slice_us}

    def timed(fn, reps=5):
        fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    w = torch.empty((H, W), dtype=torch.float32, device="cuda")
    wb = torch.empty_like(w)
    na = torch.zeros((H, W), dtype=torch.int64, device="cuda")
    nb = torch.zeros_like(na)

    def scheme1():
        w.fill_(0.5)
        lib.check(lib.nsof_event_scheme1_f32(h, dx.data_ptr(), dy.data_ptr(), db.data_ptr(), n_slices, H, W, 1, -6.0, 0.3,
                                             C.byref(sp), 5e-4, w.data_ptr(), st()))

    def scheme2():
        w.fill_(0.5); wb.fill_(0.5); na.zero_(); nb.zero_()
        lib.check(lib.nsof_event_scheme2_f32(h, dx.data_ptr(), dy.data_ptr(), dp.data_ptr(), dt_.data_ptr(), db.data_ptr(),
                                             n_slices, H, W, 0, -6.0, 0.3, 800, C.byref(sp), 5e-4, w.data_ptr(), wb.data_ptr(),
                                             na.data_ptr(), nb.data_ptr(), st()))

    for nm, fn, arrays in (("scheme1", scheme1, 1), ("scheme2_split", scheme2, 2)):
        ms = timed(fn)
        # algorithmic bytes: 4 B / event (+1 polarity, scheme 2) + per 32-slice chunk a read+write of w (and of the
        # int64 refractory map) for every array
        chunks = (n_slices + 31) // 32
        by = len(x) * (4 if arrays == 1 else 5) + chunks * arrays * H * W * (8 if arrays == 1 else 24)
        # The bound of this path is NOT bytes: with a bias above `von` every pixel evaluates one powf per slice and array
        # (w stays in registers over the 32 slices of a chunk).  CUDA's powf is ~2 MUFU (lg2, ex2) + ~40 FP32
        # instructions, so the ceilings are 16 MUFU / clk / SM and 128 FP32 lanes / clk / SM at 148 SMs x 1.965 GHz.
        powf_s = n_slices * arrays * H * W / (ms * 1e-3)
        sm_clk = 148 * 1.965e9
        out[nm] = {"ms": ms, "slices_per_s": n_slices / (ms * 1e-3), "events_per_s": len(x) / (ms * 1e-3),
                   "algorithmic_GBps": by / (ms * 1e-3) / 1e9, "bound": "fp32 issue (powf), not HBM",
                   "powf_per_s": powf_s, "frac_of_mufu_peak": powf_s * 2 / (16 * sm_clk),
                   "frac_of_fp32_peak_at_50_instr_per_update": powf_s * 50 / (128 * sm_clk)}
    # CPU: the reference's numpy loop on the first cpu_slices slices (bias 0.3: every pixel updates every slice, as above)
    n_ev = int(bounds[cpu_slices])
    t0 = time.perf_counter()
    s1.scheme1(x[:n_ev], y[:n_ev], t[:n_ev], H, W, slice_us, -6.0, 0.3)
    c1 = time.perf_counter() - t0
    t0 = time.perf_counter()
    s1.scheme2(x[:n_ev], y[:n_ev], p[:n_ev], t[:n_ev], H, W, slice_us, -6.0, 0.3, "split")
    c2 = time.perf_counter() - t0
    ncpu = len(s1.slice_bounds(t[:n_ev], slice_us)) - 1
    out["cpu_numpy"] = {"sample_slices": ncpu, "cores": 1, "scheme1_slices_per_s": ncpu / c1, "scheme2_split_slices_per_s": ncpu / c2}
    print(json.dumps({name: out}), flush=True)


run("events_240x320", 240, 320, (50, 50), 300.0)
run("events_1080x1920", 1080, 1920, (270, 480), 1200.0, cpu_slices=20)
