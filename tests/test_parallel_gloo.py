"""N > 1 host logic on CPU: two gloo ranks shard clips (c mod G), generate their own clips and agree on the
whole-job numbers; no data-path collective is involved."""
import importlib
import os
import socket

import numpy as np
import torch
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_dir):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    par = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.parallel")
    syn = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.synthetic")
    r, w = par.init("gloo")
    assert (r, w) == (rank, world)
    mine = par.weak_scaling_clips(3, rank, world)
    # every rank generates only its own clips (small frames here); checksums prove they differ per clip
    sums = [int(syn.make_clip(par.clip_seed(c), 48, 64, 3, "cpu").sum()) for c in mine]
    pairs_local = len(mine) * 2
    ms_local = 10.0 * (rank + 1)
    par.barrier()
    pairs, ms = par.whole_job(pairs_local, ms_local)
    np.save(os.path.join(out_dir, f"r{rank}.npy"), np.array([pairs, ms] + mine + sums, dtype=np.float64))
    torch.distributed.destroy_process_group()


def test_two_ranks_shard_clips_without_collective(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    r0, r1 = (np.load(tmp_path / f"r{r}.npy") for r in range(world))
    assert r0[0] == r1[0] == 12 and r0[1] == r1[1] == 20.0  # 6 clips x 2 pairs; max over ranks of the time
    clips0, clips1 = r0[2:5].astype(int).tolist(), r1[2:5].astype(int).tolist()
    assert clips0 == [0, 2, 4] and clips1 == [1, 3, 5]
    assert len(set(r0[5:8].tolist() + r1[5:8].tolist())) == 6  # six different clips


def test_strong_scaling_shard_covers_every_clip_once():
    par = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.parallel")
    for world in (1, 2, 4, 8):
        owned = [par.clips_of_rank(64, r, world) for r in range(world)]
        assert sorted(c for o in owned for c in o) == list(range(64))
        assert max(len(o) for o in owned) - min(len(o) for o in owned) == 0
