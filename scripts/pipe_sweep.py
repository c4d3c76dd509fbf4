"""Tuning helper: time SequencePipeline.step / HostPipeline.run for several chunk schedules and stream counts.
usage: python scripts/pipe_sweep.py [clips] [frames]"""
import importlib
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

Cn = int(sys.argv[1]) if len(sys.argv) > 1 else 8
F = int(sys.argv[2]) if len(sys.argv) > 2 else 16
H, W = 1080, 1920
nsof = importlib.import_module("nsof_b200")
syn = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.synthetic")
pm = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.pipeline")
frames = torch.empty((F, Cn, H, W), dtype=torch.uint8, device="cuda")
for c in range(Cn):
    frames[:, c] = syn.make_clip(1234 + c, H, W, F, "cuda")
torch.cuda.synchronize()


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


P = F - 1
scheds = {k: v for k, v in {"1": [(0, P)], "1,2,4,8": pm.chunk_schedule(P), "1,2,4,4,4": pm.chunk_schedule(P, 1, 2, 4),
          "2,4,9": [(0, 2), (2, 6), (6, P)], "1,3,11": [(0, 1), (1, 4), (4, P)], "3x5": [(0, 5), (5, 10), (10, P)],
          "1,2,3,3,3,3": [(0, 1), (1, 3), (3, 6), (6, 9), (9, 12), (12, P)], "2,13": [(0, 2), (2, P)]}.items() if not os.environ.get("NSOF_SCHEDS") or k in os.environ["NSOF_SCHEDS"].split(";")}
for spec in os.environ.get("NSOF_SCHEDS", "").split(";"):  # sizes not in the table: "1,2,4,4,2,1,1"
    if spec and spec not in scheds and spec != "1":
        sizes, pos, ch = [int(v) for v in spec.split(",")], 0, []
        for n_ in sizes:
            ch.append((pos, min(P, pos + n_)))
            pos += n_
        if pos == P:
            scheds[spec] = ch
step_streams = [int(v) for v in os.environ.get("NSOF_S2", "1,2,3").split(",") if v]
for name, ch in scheds.items():
    if ch[-1][1] != P:
        continue
    for ns in step_streams:
        eng = nsof.Engine(0)
        pipe = pm.SequencePipeline(eng, Cn, F, H, W, chunks=ch, s2_streams=ns)
        ms = t_step(pipe)
        print(f"step  chunks {name:12s} s2_streams {ns}: {ms:7.3f} ms  {pipe.n_pairs / ms * 1e3:9.0f} pairs/s", flush=True)
        if os.environ.get("NSOF_TIMELINE"):
            torch.cuda.synchronize()
            e0 = torch.cuda.Event(enable_timing=True)
            e0.record()
            h0 = time.perf_counter()
            pipe.step(frames)
            h1 = time.perf_counter()
            torch.cuda.synchronize()
            print("      host enqueue", round((h1 - h0) * 1e3, 3), "ms; stage1 done", [round(e0.elapsed_time(e), 3) for e in pipe.ev_chunk],
                  "stage2 done", [round(e0.elapsed_time(e), 3) for e in pipe.ev_s2], flush=True)
        del pipe, eng
        torch.cuda.empty_cache()
eng = nsof.Engine(0)
pipe = pm.SequencePipeline(eng, Cn, F, H, W, overlap=False, chunks=[(0, P)])
print(f"step  serial single chunk: {t_step(pipe):7.3f} ms", flush=True)
del pipe, eng
torch.cuda.empty_cache()

clips_host = [torch.empty((F, H, W), dtype=torch.uint8).pin_memory() for _ in range(Cn)]
for c in range(Cn):
    clips_host[c].copy_(frames[:, c])
for name, ch in scheds.items():
    if ch[-1][1] != P:
        continue
    hp = pm.HostPipeline(0, Cn, F, H, W, chunks=ch)
    for _ in range(2):
        hp.run(clips_host)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(5):
        hp.run(clips_host)
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) / 5 * 1e3
    print(f"host  chunks {name:12s}: {ms:7.3f} ms  {hp.pipe.n_pairs / ms * 1e3:9.0f} pairs/s", flush=True)
    if os.environ.get("NSOF_TIMELINE"):
        for r in hp.timeline():
            print("     ", r, flush=True)
    del hp
    torch.cuda.empty_cache()

# steps overlapped through HostStream (depth = steps in flight)
for depth in [int(v) for v in os.environ.get("NSOF_DEPTHS", "2,3,4").split(",") if v]:
    for name, ch in scheds.items():
        if ch[-1][1] != P:
            continue
        hs = pm.HostStream(0, Cn, F, H, W, depth=depth, chunks=ch)
        for _ in range(depth + 1):
            hs.submit(clips_host)
        hs.drain()
        torch.cuda.synchronize()
        for K in (5, 20):
            t0 = time.perf_counter()
            for _ in range(K):
                hs.submit(clips_host)
            hs.drain()
            torch.cuda.synchronize()
            ms = (time.perf_counter() - t0) / K * 1e3
            print(f"stream depth {depth} chunks {name:12s} K={K:2d}: {ms:7.3f} ms/step  {hs.slots[0].pipe.n_pairs / ms * 1e3:9.0f} pairs/s",
                  flush=True)
        del hs
        torch.cuda.empty_cache()
