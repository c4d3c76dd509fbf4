"""Diagnostic for the host-buffer stream (bench.py e2e leg): copy-engine rates alone and together, then the
per-step device timeline of HostStream (when each step's H2D / stage 2 / D2H finished, when the host submitted)."""
import importlib
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
nsof = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200")
pm = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.pipeline")
syn = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.synthetic")
Cn, F, H, W = 8, 16, 1080, 1920
dev = torch.device("cuda:0")
frames = torch.stack([syn.make_clip(100 + c, H, W, F, dev) for c in range(Cn)], dim=1).contiguous()
clips_host = [frames[:, c].cpu().pin_memory() for c in range(Cn)]
K = int(os.environ.get("K", "12"))

# ---- copy engines alone
hbuf = torch.empty(265_420_800, dtype=torch.uint8).pin_memory()
dbuf = torch.empty_like(hbuf, device=dev)
obuf = torch.empty(226_511_200, dtype=torch.uint8).pin_memory()
dout = torch.empty_like(obuf, device=dev)
s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
def rate(do_in, do_out, n=10):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        if do_in:
            with torch.cuda.stream(s1):
                dbuf.copy_(hbuf, non_blocking=True)
        if do_out:
            with torch.cuda.stream(s2):
                obuf.copy_(dout, non_blocking=True)
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3
for _ in range(2):
    rate(True, True, 2)
print(f"H2D 265 MB alone {rate(True, False):.3f} ms; D2H 227 MB alone {rate(False, True):.3f} ms; both {rate(True, True):.3f} ms", flush=True)

depth = int(os.environ.get("DEPTH", "3"))
chunks = None
if os.environ.get("CHUNKS"):
    sizes = [int(v) for v in os.environ["CHUNKS"].split(",")]
    chunks, a = [], 0
    for sz in sizes:
        chunks.append((a, a + sz))
        a += sz
hs = pm.HostStream(0, Cn, F, H, W, depth=depth, chunks=chunks)
for _ in range(depth + 1):
    hs.submit(clips_host)
hs.drain()
torch.cuda.synchronize()
g0 = torch.cuda.Event(enable_timing=True)
g0.record()
torch.cuda.synchronize()
t0 = time.perf_counter()
rows = []
base = hs.k
def note(k):
    hp = hs.slots[(base + k) % depth]
    rows.append((k, g0.elapsed_time(hp.ev_start), g0.elapsed_time(hp.ev_in[-1]), g0.elapsed_time(hp.pipe.ev_s2[-1]),
                 g0.elapsed_time(hp.ev_d2h[-1])))
marks = []
for k in range(K):
    a = time.perf_counter()
    if k >= depth:
        hs.slots[(base + k) % depth].ev_out.synchronize()
        note(k - depth)
    hs.submit(clips_host)
    marks.append(((a - t0) * 1e3, (time.perf_counter() - t0) * 1e3))
hs._flow_pending()
torch.cuda.synchronize()
total = (time.perf_counter() - t0) * 1e3
for k in range(max(K - depth, 0), K):
    note(k)
hs.drain()
print(f"{K} steps in {total:.2f} ms = {total / K:.3f} ms/step")
print("step  h2d_start  h2d_end  stage2_end  d2h_end | host submit begin..end")
for (k, a, b, c, d), (m0, m1) in zip(rows, marks):
    print(f"{k:3d}  {a:8.2f} {b:8.2f} {c:10.2f} {d:9.2f} | {m0:7.2f} .. {m1:7.2f}")
