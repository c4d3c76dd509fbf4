"""Times the step with and without the motion-segmentation post-process on the bench
workload, the segmentation launches per class, and the reference's CPU path (cv2) on the same regions."""
import importlib
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
nsof = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200")
pm = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.pipeline")
syn = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.synthetic")
Cn, F, H, W = 8, 16, 1080, 1920
dev = torch.device("cuda:0")
frames = torch.stack([syn.make_clip(1234 + c, H, W, F, dev) for c in range(Cn)], dim=1).contiguous()


def t_step(pipe, n=5):
    for _ in range(3):
        pipe.step(frames)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        pipe.step(frames)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


res = {}
for seg in (False, True):
    eng = nsof.Engine(0)
    pipe = pm.SequencePipeline(eng, Cn, F, H, W, segment=seg)
    res[seg] = t_step(pipe)
    print(f"step segment={seg}: {res[seg]:.3f} ms  ({pipe.n_pairs / res[seg] * 1e3:.0f} pairs/s)", flush=True)
    if seg:
        for label, chunks in (("time chunks 1,2,4,4,4", None), ("one batch per step", [(0, F - 1)])):
            ser = pm.SequencePipeline(nsof.Engine(0), Cn, F, H, W, segment=True, overlap=False, chunks=chunks)
            ser.step(frames)
            torch.cuda.synchronize()
            ser.eng.profile(True)
            ser.step(frames)
            torch.cuda.synchronize()
            rows = [dict(name=k, **v) for k, v in ser.eng.profile_read().items() if k.startswith("seg_")]
            ser.eng.profile(False)
            print(f"   {label}: " + ", ".join(f"{r['name']} {r['launches']} x {r['ms'] / r['launches'] * 1e3:.1f} us" for r in rows)
                  + f" = {sum(r['ms'] for r in rows):.3f} ms per step", flush=True)
            del ser
        ser = pm.SequencePipeline(nsof.Engine(0), Cn, F, H, W, segment=True, overlap=False)
        ser.step(frames)
        torch.cuda.synchronize()
        ser.eng.profile(True)
        ser.step(frames)
        torch.cuda.synchronize()
        rows = [dict(name=k, **v) for k, v in ser.eng.profile_read().items() if k.startswith("seg_")]
        ser.eng.profile(False)
        for r in rows:
            print(f"   {r['name']:12s} {r['launches']:4d} launches  {r['ms']:.3f} ms/step", flush=True)
        print(f"   segmentation kernels: {sum(r['ms'] for r in rows):.3f} ms per step of {ser.n_items} regions, "
              f"{ser.last_crop_pixels / 1e6:.1f} Mpx", flush=True)
        # the reference's CPU path on the same regions (cv2, one thread of work per call as in the script)
        import cv2
        flow = pipe.flow[:2].cpu().numpy()
        rects, n_rects = pipe.last_rects
        k = cv2.getStructuringElement(cv2.MORPH_ELLIPSE, (10, 10))
        t0, px = time.perf_counter(), 0
        for i in range(2 * Cn):
            if not n_rects[i]:
                continue
            t, c = divmod(i, Cn)
            x0, y0, x1, y1 = (int(v) for v in rects[i])
            fr = -flow[t, c, y0:y1, x0:x1].astype(np.float64)
            mag, ang = cv2.cartToPolar(fr[..., 0], fr[..., 1])
            m = np.zeros(mag.shape, np.uint8)
            m[mag > 1] = 255
            for _ in range(5):
                m = cv2.erode(cv2.dilate(m, k), k)
            px += m.size
        dt = time.perf_counter() - t0
        print(f"   cv2 on the host (threshold + 5 x dilate/erode only): {dt / px * pipe.last_crop_pixels * 1e3:.1f} ms per step "
              f"equivalent ({px / dt / 1e6:.1f} Mpx/s)", flush=True)
print(f"segmentation adds {res[True] - res[False]:.3f} ms per step ({(res[True] / res[False] - 1) * 100:.1f} %)")
# one Farneback batch per step (DeviceStream, the bench's value leg): 12 segmentation launches per step
ds_ms = {}
for seg in (False, True):
    ds = pm.DeviceStream(nsof.Engine(0), Cn, F, H, W, segment=seg)
    for _ in range(3):
        ds.submit(frames)
    ds.drain()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(8):
        ds.submit(frames)
    ds.drain()
    e1.record()
    torch.cuda.synchronize()
    ds_ms[seg] = e0.elapsed_time(e1) / 8
    del ds
    torch.cuda.empty_cache()
print(f"DeviceStream (8 steps): {ds_ms[False]:.3f} ms per step, with segmentation {ds_ms[True]:.3f} ms (+{(ds_ms[True] / ds_ms[False] - 1) * 100:.1f} %)")
