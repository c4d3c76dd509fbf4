"""Profiling driver (ncu target): a few full-frame 1080p Farneback batches, nothing else.
usage: python scripts/ff_step.py [pairs] [steps] [roi]"""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

pairs = int(sys.argv[1]) if len(sys.argv) > 1 else 8
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
roi = len(sys.argv) > 3 and sys.argv[3] == "roi"
nsof = importlib.import_module("nsof_b200")
syn = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.synthetic")
pm = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.pipeline")
eng = nsof.Engine(0)
H, W, F = 1080, 1920, pairs + 1
frames = syn.make_clip(1234, H, W, F, "cuda")[:, None].contiguous()
pipe = pm.SequencePipeline(eng, 1, F, H, W, full_frame=not roi)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
pipe.step(frames)
torch.cuda.synchronize()
e0.record()
for _ in range(steps):
    pipe.step(frames)
e1.record()
torch.cuda.synchronize()
print(f"{pairs} pairs x {steps} steps: {e0.elapsed_time(e1) / steps / pairs:.4f} ms/pair")
eng.profile(True)
pipe.step(frames)
eng.profile(False)
st = eng.profile_read()
agg = {}
for k, v in st.items():
    c = k.split(".")[0]
    agg[c] = agg.get(c, 0.0) + v["ms"] / pairs
print("  per pair us:", {k: round(v * 1e3, 1) for k, v in agg.items()}, "L0:", {k: round(v["ms"] / pairs * 1e3, 1) for k, v in st.items() if k.endswith("L0")})
