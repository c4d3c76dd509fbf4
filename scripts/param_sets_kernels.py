"""Per-kernel table (CUDA events around every launch, nsof_profile) of a full-frame 1080p step for the two parameter sets
that run on the generic kernels: (0.6, 3, 3, 3, 10, 1.05) and (0.6, 3, 4, 2, 1, 1.05).
usage: python scripts/param_sets_kernels.py"""
import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
nsof = importlib.import_module("nsof_b200")
syn = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.synthetic")
pm = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.pipeline")
H, W, F = 1080, 1920, 9
dev = torch.device("cuda:0")
fr = syn.make_clip(1234, H, W, F, dev)[:, None].contiguous()
eng = nsof.Engine(0)
for name, p in (("autodriving_txt", dict(pyr_scale=0.6, levels=3, winsize=3, iterations=3, poly_n=10, poly_sigma=1.05, flags=0)),
                ("tabletennis", dict(pyr_scale=0.6, levels=3, winsize=4, iterations=2, poly_n=1, poly_sigma=1.05, flags=0)),
                ("grasp", dict(nsof.FARNEBACK_PARAMS))):
    pf = pm.SequencePipeline(eng, 1, F, H, W, nsof.RoiConfig(farneback_params=p), full_frame=True)
    for _ in range(2):
        pf.step(fr)
    torch.cuda.synchronize()
    eng.profile(True)
    for _ in range(2):
        pf.step(fr)
    eng.profile(False)
    st = eng.profile_read()
    tot = sum(s["ms"] for s in st.values())
    print(f"{name}: {tot / 2:.3f} ms per step of {pf.n_pairs} pairs (serialised kernel sum)")
    for k, s in sorted(st.items(), key=lambda kv: -kv[1]["ms"])[:int(os.environ.get("NSOF_ROWS", "12"))]:
        gb = s["algorithmic_bytes"] / (s["ms"] * 1e-3) / 1e9 if s["ms"] > 0 else 0
        print(f"    {k:22s} {s['ms'] / 2:7.3f} ms {s['ms'] / tot * 100:5.1f} %  {gb:7.0f} GB/s algorithmic")
    del pf
