"""Device timings of the frame-side neighbours of the path on the bench workload's
shapes, with cv2 on the host beside them: grey conversion of 128 BGR 1080p frames, prediction warp of 120 ROI regions."""
import importlib
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
nsof = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200")
eng = nsof.Engine(0)
H, W, NF, NR = 1080, 1920, 128, 120
rh, rw = 420, 560
g = torch.Generator(device="cuda").manual_seed(1)
bgr = torch.randint(0, 256, (NF, H, W, 3), dtype=torch.uint8, device="cuda", generator=g)
grey = torch.empty((NF, H, W), dtype=torch.uint8, device="cuda")
flow = torch.randn((NR, rh, rw, 2), device="cuda", generator=g) * 4
out = torch.empty((NR, rh, rw, 3), dtype=torch.uint8, device="cuda")
st = torch.cuda.current_stream().cuda_stream


def timed(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


ms = timed(lambda: eng.lib.check(eng.lib.nsof_rgb2gray_u8(eng.handle, bgr.data_ptr(), W * 3, H * W * 3, NF, H, W, grey.data_ptr(), W,
                                                          H * W, st)))
by = NF * H * W * 4
print(f"rgb2gray {NF} x 1080p: {ms:.3f} ms  ({by / ms / 1e6:.0f} GB/s algorithmic, 4 B per pixel)", flush=True)
regions = [(bgr[i % NF].data_ptr(), W * 3, H, W, 3, flow[i].data_ptr(), rw * 8, 300, 200, rh, rw, out[i].data_ptr(), rw * 3)
           for i in range(NR)]
ms2 = timed(lambda: eng.remap_flow_device(regions, st))
by2 = NR * rh * rw * (8 + 3 + 3)
print(f"remap_flow {NR} regions of {rh} x {rw} x 3: {ms2:.3f} ms  ({by2 / ms2 / 1e6:.0f} GB/s algorithmic, 14 B per pixel)", flush=True)
import cv2  # noqa: E402
f0 = bgr[0].cpu().numpy()
t0 = time.perf_counter()
for _ in range(8):
    cv2.cvtColor(f0, cv2.COLOR_RGB2GRAY)
tg = (time.perf_counter() - t0) / 8
fl = flow[0].cpu().numpy()
mx = (np.arange(300, 300 + rw)[None, :] + fl[..., 0].astype(np.float64)).astype(np.float32)
my = (np.arange(200, 200 + rh)[:, None] + fl[..., 1].astype(np.float64)).astype(np.float32)
t0 = time.perf_counter()
for _ in range(4):
    for c in range(3):
        cv2.remap(f0[:, :, c], mx, my, cv2.INTER_LINEAR, borderMode=cv2.BORDER_REPLICATE)
tr = (time.perf_counter() - t0) / 4
print(f"cv2 on the host: cvtColor {tg * 1e3 * NF:.1f} ms per {NF} frames; 3-channel remap {tr * 1e3 * NR:.1f} ms per {NR} regions")
