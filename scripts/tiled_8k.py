"""BASELINE.json configs[4]: synthetic 8K (7680x4320) frame pairs, each split by rows over all ranks, halo rows over
NCCL (tiled.py).  Launch with torchrun (one rank per GPU); rank 0 prints one JSON line.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
      scripts/tiled_8k.py [--pairs 4] [--steps 3] [--check]

--check first runs a 1080x1920 pair tiled and untiled and requires the gathered bands to be bit-identical."""
import argparse
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--pairs", type=int, default=4)
ap.add_argument("--steps", type=int, default=3)
ap.add_argument("--warmup", type=int, default=2)
ap.add_argument("--height", type=int, default=4320)
ap.add_argument("--width", type=int, default=7680)
ap.add_argument("--halo", type=int, default=32)
ap.add_argument("--check", action="store_true")
ap.add_argument("--overlap", action="store_true", help="edges -> exchange on a second stream || interior (tiled.py)")
ap.add_argument("--linked", action="store_true", help="peer-memory path: the level loop in one library call (run_linked)")
args = ap.parse_args()

nsof = importlib.import_module("nsof_b200")
syn = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.synthetic")
par = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.parallel")
tiled = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.tiled")
rank, world, local = par.rank_info()
torch.cuda.set_device(local)
dev = torch.device(f"cuda:{local}")
os.environ.setdefault("NCCL_DEBUG", "WARN")
par.init("nccl", dev)
eng = nsof.Engine(local)
tf = tiled.TiledFarneback(eng, rank, world, halo_r1=args.halo, overlap=args.overlap)
P = nsof.FARNEBACK_PARAMS


def crops_of(frames, flow):
    F, H, W = frames.shape
    return [(frames[t].data_ptr(), frames[t + 1].data_ptr(), H, W, W, W, flow[t].data_ptr(), W * 8) for t in range(F - 1)]


if args.check:
    H, W = 1080, 1920
    frames = syn.make_clip(4321, H, W, 3, dev, velocity=(5.0, -4.0))  # same seed on every rank: the same frames
    flow = torch.full((2, H, W, 2), float("nan"), dtype=torch.float32, device=dev)
    tfc = tiled.TiledFarneback(eng, rank, world, halo_r1=min(args.halo, (H // 8) // world), overlap=args.overlap)  # 135 coarsest rows
    (tfc.run_linked if args.linked else tfc.run)(crops_of(frames, flow), **P)
    torch.cuda.synchronize()
    lo, hi = tiled.band_rows(H, rank, world)
    ref = torch.empty((2, H, W, 2), dtype=torch.float32, device=dev)
    eng.farneback_batch_device(crops_of(frames, ref), torch.cuda.current_stream().cuda_stream, **P)
    torch.cuda.synchronize()
    same = torch.equal(flow[:, lo:hi], ref[:, lo:hi])
    untouched = bool(torch.isnan(flow[:, :lo]).all() and torch.isnan(flow[:, hi:]).all())
    ok = torch.tensor([int(same and untouched)], device=dev)
    if world > 1:
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(f"# check 1080p x{world}: bands bit-identical to the untiled flow on every rank: {bool(ok.item())}; "
              f"{tfc.messages} p2p ops, {tfc.bytes_sent / 1e6:.1f} MB sent by rank 0", flush=True)
    assert ok.item() == 1
    del frames, flow, ref, tfc

H, W, NP = args.height, args.width, args.pairs
frames = syn.make_clip(999, H, W, NP + 1, dev, velocity=(6.0, -3.0))
flow = torch.zeros((NP, H, W, 2), dtype=torch.float32, device=dev)
crops = crops_of(frames, flow)


def barrier():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


step = (lambda: tf.run_linked(crops, check_guard=False, **P)) if args.linked else (lambda: tf.run(crops, **P))
for _ in range(args.warmup):
    step()
barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(args.steps):
    step()
e1.record()
barrier()
ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
if world > 1:
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
if rank == 0:
    ms_step = float(ms.item()) / args.steps
    bytes_pair = eng.algorithmic_bytes(H, W)
    print(json.dumps({"metric": "tiled_farneback_frame_pairs_per_sec_8k", "value": NP / (ms_step * 1e-3), "unit": "frame-pairs/s",
                      "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "scaling": "strong",
                      "config": {"workload": "configs[4]: synthetic 8K frames, row-tiled over the ranks, NCCL halo rows",
                                 "frame": [H, W], "pairs_per_step": NP, "halo_r1": args.halo,
                                 "schedule": "peer memory: store-to-neighbour + flag-wait kernels, one host call per batch" if args.linked else
                                 "whole-band stages, blocking exchanges" if not args.overlap else
                                 f"edges -> exchange on a second stream || interior on {tf.split_levels} of the levels"},
                      "p2p_ops_per_step_rank0": tf.messages, "MB_sent_per_step_rank0": tf.bytes_sent / 1e6,
                      "algorithmic_GBps_whole_job": NP * bytes_pair / (ms_step * 1e-3) / 1e9}), flush=True)
if dist.is_initialized():
    dist.destroy_process_group()
