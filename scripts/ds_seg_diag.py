import importlib, os, sys, time
import torch
sys.path.insert(0, "/root/repo")
nsof = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200")
pm = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.pipeline")
syn = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.synthetic")
lib_mod = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200._lib")
Cn, F, H, W = 8, 16, 1080, 1920
dev = torch.device("cuda:0")
frames = torch.stack([syn.make_clip(1234 + c, H, W, F, dev) for c in range(Cn)], dim=1).contiguous()
def run(seg, its=5, K=8):
    ds = pm.DeviceStream(nsof.Engine(0), Cn, F, H, W, segment=seg)
    if seg:
        for s in ds.slots:
            s.seg_params = lib_mod.SegParams(1.0, 2, 10, 10, its)
    for _ in range(3): ds.submit(frames)
    ds.drain(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    h0 = time.perf_counter()
    e0.record()
    hs = []
    for _ in range(K):
        a = time.perf_counter(); ds.submit(frames); hs.append(time.perf_counter() - a)
    ds.drain(); e1.record()
    h1 = time.perf_counter()
    torch.cuda.synchronize()
    print(f"segment={seg} iterations={its}: {e0.elapsed_time(e1)/K:.3f} ms/step device, host enqueue loop {(h1-h0)/K*1e3:.3f} ms/step, submit calls {[round(x*1e3,2) for x in hs]}", flush=True)
    del ds; torch.cuda.empty_cache()
run(False); run(True, 0); run(True, 1); run(True, 5)
