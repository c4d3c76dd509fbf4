"""Profiling driver (ncu target): stage 1 only (Lanczos compress + synapse chains + ROI) on resident 1080p clips."""
import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
nsof = importlib.import_module("nsof_b200")
syn = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.synthetic")
pm = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.pipeline")
Cn, F, H, W = 8, 16, 1080, 1920
frames = torch.empty((F, Cn, H, W), dtype=torch.uint8, device="cuda")
for c in range(Cn):
    frames[:, c] = syn.make_clip(1234 + c, H, W, F, "cuda")
eng = nsof.Engine(0)
pipe = pm.SequencePipeline(eng, Cn, F, H, W, overlap=False, chunks=[(0, F - 1)])
for _ in range(3):
    pipe.stage1(frames)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
pipe.stage1(frames)
e1.record()
torch.cuda.synchronize()
print("stage 1:", e0.elapsed_time(e1), "ms")
