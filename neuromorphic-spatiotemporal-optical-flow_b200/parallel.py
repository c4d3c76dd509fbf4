"""Multi-GPU plumbing: one process per GPU, clips sharded across ranks, NO data-path collective (SURVEY.md 8(e):
stage-2 pairs are independent, stage 1 is a recurrence only along t within a clip).  torch.distributed (NCCL on
GPUs, gloo in the CPU tests) is used for the barrier and for the max-over-ranks / sum-over-ranks of counters."""
from __future__ import annotations

import os


def rank_info():
    """(rank, world_size, local_rank) from the torchrun environment; (0, 1, 0) when launched plainly."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def owner_of_clip(clip: int, world: int) -> int:
    """clip c -> GPU c mod G."""
    return clip % world


def clips_of_rank(n_clips_total: int, rank: int, world: int) -> list:
    """Strong-scaling shard: the clips of a fixed set that `rank` owns."""
    return [c for c in range(n_clips_total) if owner_of_clip(c, world) == rank]


def weak_scaling_clips(clips_per_gpu: int, rank: int, world: int) -> list:
    """Weak-scaling shard (bench.py): clips_per_gpu * world clips in total, the same c mod G ownership."""
    return clips_of_rank(clips_per_gpu * world, rank, world)


def clip_seed(clip: int, seed0: int = 1234) -> int:
    return seed0 + clip


def init(backend: str, device=None):
    import torch.distributed as dist
    rank, world, _ = rank_info()
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        kw = {"device_id": device} if (device is not None and backend == "nccl") else {}
        dist.init_process_group(backend, rank=rank, world_size=world, **kw)
    return rank, world


def barrier():
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        dist.barrier()


def whole_job(pairs_local: float, ms_local: float, device="cpu"):
    """-> (pairs of all ranks, max-over-ranks milliseconds): the two numbers `value` is made of."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return float(pairs_local), float(ms_local)
    p = torch.tensor([float(pairs_local)], dtype=torch.float64, device=device)
    m = torch.tensor([float(ms_local)], dtype=torch.float64, device=device)
    dist.all_reduce(p, op=dist.ReduceOp.SUM)
    dist.all_reduce(m, op=dist.ReduceOp.MAX)
    return float(p.item()), float(m.item())
