"""Spatial tiling (BASELINE.json configs[4]) on CPU: 2 and 3 gloo ranks run the row-tiled Farneback driver
(tiled.py + nsof_fb_tiled_*) over the block-emulator build of the kernels, exchange halo rows, and together reproduce
the untiled result bit for bit."""
import importlib
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import helpers  # noqa: E402

H, W = 384, 328
PARAMS = dict(pyr_scale=0.5, levels=3, winsize=15, iterations=3, poly_n=5, poly_sigma=1.2, flags=0)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _pairs():
    return [helpers.textured_pair(7, H, W, dx=2.5, dy=-3.0), helpers.textured_pair(8, H, W, dx=-1.0, dy=4.5)]


def _worker(rank, world, port, out_dir, halo, overlap=True, params=None):
    params = params or PARAMS
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    nsof = importlib.import_module("nsof_b200")
    tiled = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.tiled")
    par = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.parallel")
    par.init("gloo")
    eng = nsof.Engine(0, lib=nsof.Library(helpers.build_emu()), _emulated=True)
    pairs = _pairs()
    flows = [np.full((H, W, 2), np.nan, dtype=np.float32) for _ in pairs]
    crops = [(a.ctypes.data, b.ctypes.data, H, W, a.strides[0], b.strides[0], f.ctypes.data, f.strides[0])
             for (a, b), f in zip(pairs, flows)]
    tf = tiled.TiledFarneback(eng, rank, world, halo_r1=halo, overlap=overlap)
    status = "ok"
    try:
        tf.run(crops, **params)
    except RuntimeError as e:
        status = "guard" if "halo" in str(e) else "error: " + str(e)
    lo, hi = tiled.band_rows(H, rank, world)
    np.savez(os.path.join(out_dir, f"r{rank}.npz"), lo=lo, hi=hi, status=status, messages=tf.messages, split=tf.split_levels,
             **{f"flow{i}": f for i, f in enumerate(flows)})
    torch.distributed.destroy_process_group()
    eng.close()


@pytest.mark.parametrize("world", [2, 3])
def test_tiled_ranks_reproduce_the_untiled_flow(tmp_path, emu_engine, world):
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path), 12), nprocs=world, join=True)
    want = [emu_engine.calcOpticalFlowFarneback(a, b, None, **PARAMS) for a, b in _pairs()]
    covered = 0
    for r in range(world):
        z = np.load(tmp_path / f"r{r}.npz")
        assert str(z["status"]) == "ok"
        lo, hi = int(z["lo"]), int(z["hi"])
        covered += hi - lo
        for i, wnt in enumerate(want):
            got = z[f"flow{i}"]
            assert np.array_equal(got[lo:hi], wnt[lo:hi]), (world, r, i, np.abs(got[lo:hi] - wnt[lo:hi]).max())
            outside = np.ones(H, bool)
            outside[lo:hi] = False
            assert np.isnan(got[outside]).all()  # a rank writes its band only
        assert int(z["messages"]) > 0
        # levels with a band of >= 4 halos ran as edges -> exchange -> interior (nsof_fb_tiled_stage_rows), the rest whole
        assert int(z["split"]) == (3 if world == 2 else 2)
    assert covered == H


@pytest.mark.parametrize("pname", ["autodriving_txt", "tabletennis"])
def test_tiled_other_parameter_sets(tmp_path, emu_engine, pname):
    """The two other shipped parameter sets (window radius 1 / 2 on the tuned blur+solve kernel's small-radius
    instantiations, poly_n 10 / 1 with the fused level-0 pyramid, bilinear 0.6 pyramid above it) row-tiled over 2 ranks:
    bands bit-identical to the untiled flow."""
    params = helpers.PARAM_SETS[pname]
    mp.spawn(_worker, args=(2, _free_port(), str(tmp_path), 12, True, params), nprocs=2, join=True)
    want = [emu_engine.calcOpticalFlowFarneback(a, b, None, **params) for a, b in _pairs()]
    for r in range(2):
        z = np.load(tmp_path / f"r{r}.npz")
        assert str(z["status"]) == "ok", str(z["status"])
        lo, hi = int(z["lo"]), int(z["hi"])
        for i, wnt in enumerate(want):
            got = z[f"flow{i}"]
            assert np.array_equal(got[lo:hi], wnt[lo:hi]), (pname, r, i, np.abs(got[lo:hi] - wnt[lo:hi]).max())


def test_tiled_guard_trips_when_the_halo_is_too_small(tmp_path):
    """dy = 4.5 px at full resolution needs R1 rows the 1-row halo does not bring: every rank must raise."""
    mp.spawn(_worker, args=(2, _free_port(), str(tmp_path), 1), nprocs=2, join=True)
    for r in range(2):
        assert str(np.load(tmp_path / f"r{r}.npz")["status"]) == "guard"


@pytest.mark.parametrize("world", [2, 3])
def test_tiled_peer_memory_run_reproduces_the_untiled_flow(emu_engine, nsof, world):
    """nsof_fb_tiled_run (the level loop in one call; halo rows stored straight into the neighbour's workspace, flags
    instead of messages) with the ranks as THREADS of this process over the emulator build: bands bit-identical to the
    untiled flow, twice in a row (the exchange sequence numbers carry over between batches)."""
    import threading
    tiled = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.tiled")
    lib = nsof.Library(helpers.build_emu())
    pairs = _pairs()
    engs = [nsof.Engine(0, lib=lib, _emulated=True) for _ in range(world)]
    tfs = {r: tiled.TiledFarneback(engs[r], r, world, halo_r1=12) for r in range(world)}
    flows = {r: [np.full((H, W, 2), np.nan, dtype=np.float32) for _ in pairs] for r in range(world)}
    crops = {r: [(a.ctypes.data, b.ctypes.data, H, W, a.strides[0], b.strides[0], f.ctypes.data, f.strides[0])
                 for (a, b), f in zip(pairs, flows[r])] for r in range(world)}
    want = [emu_engine.calcOpticalFlowFarneback(a, b, None, **PARAMS) for a, b in pairs]
    for rep in range(2):
        os.environ["NSOF_TILED_SPLIT"] = str(rep)   # second repetition: edges -> push -> interior -> wait
        for r in range(world):
            for f in flows[r]:
                f.fill(np.nan)
            tfs[r].begin_linked(crops[r], **PARAMS)
        for r in range(world):
            tfs[r]._link(peers=tfs)
        errs = []

        def go(r):
            try:
                tfs[r].go_linked()
            except Exception as e:  # pragma: no cover
                errs.append((r, e))

        th = [threading.Thread(target=go, args=(r,)) for r in range(world)]
        for t in th:
            t.start()
        for t in th:
            t.join(timeout=300)
        assert not errs and not any(t.is_alive() for t in th), errs
        for r in range(world):
            lo, hi = tiled.band_rows(H, r, world)
            tripped = __import__("ctypes").c_int(0)
            lib.check(lib.nsof_fb_tiled_guard(engs[r].handle, __import__("ctypes").byref(tripped), None))
            assert tripped.value == 0
            for i, wnt in enumerate(want):
                assert np.array_equal(flows[r][i][lo:hi], wnt[lo:hi]), (rep, r, i)
                outside = np.ones(H, bool)
                outside[lo:hi] = False
                assert np.isnan(flows[r][i][outside]).all()
    os.environ.pop("NSOF_TILED_SPLIT", None)
    assert tfs[0]._seq == tfs[world - 1]._seq > 0
    for e in engs:
        e.close()
