"""Copy-engine rates while the SMs stream HBM: H2D / D2H alone, together, and each with a device-to-device copy
kernel (a stand-in for the HBM-bound Farneback kernels) running on another stream."""
import time

import torch

dev = torch.device("cuda:0")
hbuf = torch.empty(265_420_800, dtype=torch.uint8).pin_memory()
dbuf = torch.empty_like(hbuf, device=dev)
obuf = torch.empty(226_511_200, dtype=torch.uint8).pin_memory()
dout = torch.empty_like(obuf, device=dev)
a = torch.empty(1 << 30, dtype=torch.uint8, device=dev)
b = torch.empty_like(a)
s1, s2, s3 = torch.cuda.Stream(dev), torch.cuda.Stream(dev), torch.cuda.Stream(dev)


def rate(do_in, do_out, busy, n=10):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if busy:
        with torch.cuda.stream(s3):
            for _ in range(40 * n):  # 2 GB of traffic each, ~0.33 ms
                b.copy_(a, non_blocking=True)
    with torch.cuda.stream(s1):
        e0.record()
        for _ in range(n if do_in else 0):
            dbuf.copy_(hbuf, non_blocking=True)
        e1.record()
    with torch.cuda.stream(s2):
        f0.record()
        for _ in range(n if do_out else 0):
            obuf.copy_(dout, non_blocking=True)
        f1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n, f0.elapsed_time(f1) / n


rate(True, True, True, 2)
for busy in (False, True):
    for do_in, do_out in ((True, False), (False, True), (True, True)):
        i, o = rate(do_in, do_out, busy)
        print(f"SMs {'streaming HBM' if busy else 'idle         '}: H2D 265 MB {i:6.3f} ms   D2H 227 MB {o:6.3f} ms", flush=True)
