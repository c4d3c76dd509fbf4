"""H2D copy patterns of the host pipeline, alone and against a concurrent D2H stream: one flat copy per step, one
flat copy per time chunk, and the per-clip strided 2-D copies HostPipeline issues (clip-major host -> time-major
device)."""
import ctypes as C
import importlib
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
nsof = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200")
lib = nsof.Engine(0).lib
dev = torch.device("cuda:0")
Cn, F, H, W = 8, 16, 1080, 1920
fsz = H * W
host_clips = [torch.empty((F, H, W), dtype=torch.uint8).pin_memory() for _ in range(Cn)]
host_flat = torch.empty((F, Cn, H, W), dtype=torch.uint8).pin_memory()
frames = torch.empty((F, Cn, H, W), dtype=torch.uint8, device=dev)
obuf = torch.empty(226_511_200, dtype=torch.uint8).pin_memory()
dout = torch.empty_like(obuf, device=dev)
s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
chunks = [(0, 2), (2, 4), (4, 8), (8, 12), (12, 16)]  # frame ranges


def h2d(kind):
    if kind == "flat":
        frames.copy_(host_flat, non_blocking=True)
    elif kind == "chunk-flat":
        for f0, f1 in chunks:
            frames[f0:f1].copy_(host_flat[f0:f1], non_blocking=True)
    elif kind == "frame-flat":
        for f0, f1 in chunks:
            for c in range(Cn):
                for f in range(f0, f1):
                    frames[f, c].copy_(host_clips[c][f], non_blocking=True)
    elif kind == "clip-major":  # device layout (C, F, H, W): one flat copy per clip and chunk
        fr = frames.view(Cn, F, H, W)
        for f0, f1 in chunks:
            for c in range(Cn):
                fr[c, f0:f1].copy_(host_clips[c][f0:f1], non_blocking=True)
    else:
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        for f0, f1 in chunks:
            for c in range(Cn):
                lib.check(lib.nsof_copy_2d_async(frames.data_ptr() + (f0 * Cn + c) * fsz, Cn * fsz,
                                                 host_clips[c].data_ptr() + f0 * fsz, fsz, fsz, f1 - f0, 1, st), "copy")


def rate(kind, with_out, n=10):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(s1):
        e0.record()
        for _ in range(n):
            h2d(kind)
        e1.record()
    if with_out:
        with torch.cuda.stream(s2):
            for _ in range(n):
                if with_out == "flat":
                    obuf.copy_(dout, non_blocking=True)
                else:
                    q = obuf.numel() // 5
                    for j in range(5):
                        obuf[j * q:(j + 1) * q].copy_(dout[j * q:(j + 1) * q], non_blocking=True)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


rate("flat", "flat", 2)
for kind in ("flat", "chunk-flat", "clip-2d", "frame-flat", "clip-major"):
    print(f"H2D {kind:10s}: alone {rate(kind, None):6.3f} ms   with flat D2H {rate(kind, 'flat'):6.3f} ms   "
          f"with chunked D2H {rate(kind, 'chunk'):6.3f} ms", flush=True)
