"""Where the device-resident step's time goes: DeviceStream as benched vs (b) stage 1 switched off (array states given),
(c) no fork of the pyramid launches onto internal streams, (d) both.  usage: python scripts/step_diag.py [steps]"""
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

def run(tag, given=False, fork=True):
    eng = nsof.Engine(0)
    ds = pm.DeviceStream(eng, Cn, F, H, W, depth=2)
    if not fork:
        for s in ds.slots:
            eng.lib.check(eng.lib.nsof_debug_fork(s.eng.handle, 0))
    for _ in range(3):
        ds.submit(frames)
    ds.drain()
    if given:  # keep the array states of the simulated run, then stop simulating
        for s in ds.slots:
            s.stage1_mode = "given"
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K):
        ds.submit(frames)
    ds.drain()
    e1.record()
    torch.cuda.synchronize()
    print(f"{tag:34s} {e0.elapsed_time(e1) / K:7.3f} ms per step", flush=True)
    for s in ds.slots:
        for e in [s.eng, s.s1_eng] + [e for e, _ in s.s2_extra]:
            e.close()

run("as benched")
run("stage 1 off (states given)", given=True)
run("no fork")
run("stage 1 off + no fork", given=True, fork=False)
run("as benched (repeat)")
