"""Host-side formats and NMS: hostio.nms against the reference's own py_cpu_nms outputs, the .mat
/ .npz round trips."""
import importlib
import os

import numpy as np

import helpers

hostio = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.hostio")


def test_nms_matches_reference():
    g = np.load(os.path.join(helpers.GOLDEN, "hostio_reference.npz"))
    for k in range(int(g["n_cases"])):
        for th in (0.2, 0.5):
            assert hostio.nms(g[f"nms__{k}__dets"], th) == [int(v) for v in g[f"nms__{k}__keep_{th}"]], (k, th)
    assert hostio.nms(np.zeros((0, 5)), 0.2) == []


def test_mat_and_npz_round_trip(tmp_path, golden):
    import scipy.io
    cur = golden.roi["autodriving__currents"]                       # (T, Hc, Wc) as the scripts index it
    scipy.io.savemat(tmp_path / "m.mat", {"constructed3DMatrix": np.moveaxis(cur, 0, 2)})
    back = hostio.load_mem_states(str(tmp_path / "m.mat"))
    assert back.dtype == np.float64 and np.array_equal(back, cur)
    w = np.linspace(0, 1, 12, dtype=np.float32).reshape(3, 4)
    hostio.save_state_npz(str(tmp_path / "s.npz"), w, [w * 2, w * 3])
    w2, r2 = hostio.load_state_npz(str(tmp_path / "s.npz"))
    assert np.array_equal(w2, w) and r2.dtype == np.float32 and r2.shape == (2, 3, 4)
