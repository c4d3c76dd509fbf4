"""A/B of the blur+solve kernel's R1 window (shared-memory gathers, default) against the L2 gathers, and of the bulk L2
prefetches of its phase 1, on the bench workload: the device-resident step (8 clips x 32 frames of 1080p, scripts' gate) and the full-frame step (ROI = 100 %),
window on / off alternating in one process, plus the share of tiles that took the window.
usage: python scripts/r1_window_ab.py [steps] [default]"""
import ctypes as C
import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
nsof = importlib.import_module("nsof_b200")
syn = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.synthetic")
pm = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.pipeline")
K = int(sys.argv[1]) if len(sys.argv) > 1 else 20
Cn, F, H, W = 8, 32, 1080, 1920
dev = torch.device("cuda:0")
frames = torch.empty((F, Cn, H, W), dtype=torch.uint8, device=dev)
for c in range(Cn):
    frames[:, c] = syn.make_clip(1234 + c, H, W, F, dev)


def engines(ds):
    out = []
    for s in ds.slots:
        out += [s.eng] + [e for e, _ in s.s2_extra]
    return out


def timed(ds, fr, n):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        ds.submit(fr)
    ds.drain()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


for tag, full, clips in (("gated step (248 pairs)", False, Cn), ("full frame (62 pairs)", True, 2)):
    eng = nsof.Engine(0)
    fr = frames[:, :clips].contiguous()
    ds = pm.DeviceStream(eng, clips, F, H, W, depth=2, full_frame=full)
    for _ in range(3):
        ds.submit(fr)
    ds.drain()
    variants = [("windows in both kernels (default)", 1, -1), ("window in blur+solve only", 2, -1),
                ("blur+solve only, R0 + R1 prefetch", 2, 0), ("blur+solve only, R1 prefetch", 2, 1),
                ("blur+solve only, R0 prefetch", 2, 2), ("L2 gathers, prefetch (round 1)", 0, 0), ("L2 gathers, no prefetch", 0, 3)]
    if len(sys.argv) > 2 and sys.argv[2] == "default":   # only the default configuration (A/B of build variants)
        variants = variants[:1]
    res = {v[0]: [] for v in variants}
    for rep in range(3):
        for name, on, mask in variants:
            for e in engines(ds):
                e.lib.check(e.lib.nsof_debug_r1_window(e.handle, on))
                e.lib.check(e.lib.nsof_debug_bs_prefetch(e.handle, mask))
            timed(ds, fr, 2)
            res[name].append(timed(ds, fr, K))
    out = (C.c_ulonglong * 4)()
    for e in engines(ds):
        e.lib.check(e.lib.nsof_debug_r1_window(e.handle, 1))
        e.lib.check(e.lib.nsof_debug_bs_prefetch(e.handle, -1))
        e.lib.check(e.lib.nsof_debug_r1_window_stats(e.handle, out, None))
    timed(ds, fr, 2)
    tot = [0, 0, 0, 0]
    for e in engines(ds):
        e.lib.check(e.lib.nsof_debug_r1_window_stats(e.handle, out, None))
        for q in range(4):
            tot[q] += out[q]
    print(f"{tag}: blur+solve tiles via window {tot[0]} / L2 {tot[1]} = {tot[0] / max(1, tot[0] + tot[1]):.4f}; first "
          f"UpdateMatrices {tot[2]} / {tot[3]} = {tot[2] / max(1, tot[2] + tot[3]):.4f}")
    for name, _, _ in variants:
        print(f"    {name:34s} {min(res[name]):7.3f} ms per step (runs {['%.3f' % v for v in res[name]]})", flush=True)
    for s in ds.slots:
        for e in [s.eng, s.s1_eng] + [e for e, _ in s.s2_extra]:
            e.close()
