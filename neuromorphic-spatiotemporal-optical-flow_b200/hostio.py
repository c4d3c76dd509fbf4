"""Host-side formats and the tiny box post-processing either side of the hot path ("keep on CPU").

* `load_mem_states`   -- `scipy.io.loadmat(MEMMATPATH)['constructed3DMatrix']` (optical_flow_seg.py:398-399) as a
                         (T, Hc, Wc) float64 array, slice t = constructed3DMatrix[:, :, t] (what `SequencePipeline(...,
                         stage1="given").currents` and `Engine.roi_from_currents` take).
* `load_events`       -- `/CD/events` of a Prophesee HDF5 recording (eventsim/event_mem_sim.py:69-75); needs h5py.
* `save_state_npz` / `load_state_npz` -- the `.npz` the event simulator writes (event_mem_sim.py:289-303: `w_final`,
                         `resistances`).
* `nms`               -- `py_cpu_nms` of optical_flow_ob.py:96-135 (greedy IoU suppression with the +1 pixel convention).
Nothing here touches the GPU.
"""
from __future__ import annotations

import numpy as np


def load_mem_states(mat_path: str, key: str = "constructed3DMatrix") -> np.ndarray:
    import scipy.io
    m = scipy.io.loadmat(mat_path)[key]
    if m.ndim != 3:
        raise ValueError(f"{key} must be (Hc, Wc, T), got {m.shape}")
    return np.ascontiguousarray(np.moveaxis(m, 2, 0), dtype=np.float64)


def load_events(h5_path: str):
    """-> x, y, p (int), t, H, W exactly as event_mem_sim.load_events returns them."""
    try:
        import h5py
    except ImportError as e:  # the build image has no h5py; the format is the reference's, not ours
        raise ImportError("load_events needs h5py (Prophesee /CD/events compound dataset)") from e
    with h5py.File(h5_path, "r") as f:
        ev = f["CD"]["events"]
        x, y, p, t = ev["x"][:], ev["y"][:], ev["p"][:].astype(int), ev["t"][:]
    return x, y, p, t, int(y.max()) + 1, int(x.max()) + 1


def save_state_npz(path: str, w_final, resistances) -> None:
    np.savez_compressed(path, w_final=np.asarray(w_final), resistances=np.array(resistances, dtype=np.float32))


def load_state_npz(path: str):
    z = np.load(path)
    return z["w_final"], z["resistances"]


def nms(dets, thresh: float):
    """dets: (n, 5) rows x1, y1, x2, y2, score -> indices kept, highest score first.  Same decisions as py_cpu_nms:
    areas and overlaps with the +1 convention, a box survives when its IoU with the kept one is <= thresh."""
    d = np.asarray(dets, dtype=np.float64)
    if d.size == 0:
        return []
    x1, y1, x2, y2, score = d.T
    area = (y2 - y1 + 1) * (x2 - x1 + 1)
    order = list(np.argsort(score)[::-1])
    keep = []
    while order:
        i = order.pop(0)
        keep.append(int(i))
        if not order:
            break
        rest = np.array(order)
        w = np.maximum(0, np.minimum(x2[i], x2[rest]) - np.maximum(x1[i], x1[rest]) + 1)
        h = np.maximum(0, np.minimum(y2[i], y2[rest]) - np.maximum(y1[i], y1[rest]) + 1)
        inter = w * h
        iou = inter / (area[i] + area[rest] - inter)
        order = [int(j) for j, v in zip(rest, iou) if v <= thresh]
    return keep
