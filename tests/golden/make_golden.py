#!/usr/bin/env python3
"""Generates the golden fixtures under tests/golden/ by running the REFERENCE's own code in the
build container:

* stage 2: `cv2.calcOpticalFlowFarneback` (cv2 4.13.0 here; the reference pins opencv-python 4.9.0) on
  small synthetic pairs and on real data/grasp + data/autodriving crops, for the three parameter sets
  the reference ships (optical_flow_seg.py:61-112, data/*/Parameters.txt).
* stage-1 tail: the reference's functions `update_transition_pic`, `process_merged_region`,
  `process_separate_regions`, `opticalFlow3D` loaded by SOURCE SLICING from
  /root/reference/optical_flow_seg.py (the script runs its experiment at import time), driven with the
  shipped `constructed_3D_matrix.mat` of data/grasp and data/autodriving.
* event mode: `generate_synthetic_events`, `bincount_2d`, `update_state`, `resistance_exp`,
  `slice_indices` imported from /root/reference/eventsim/event_mem_sim.py with a stub `h5py`.

Run from the repo root:  python tests/golden/make_golden.py
/root/reference is only needed here; tests read the committed .npz files.
"""
import ast
import os
import sys
import types

import cv2
import numpy as np
import scipy.io

REF = "/root/reference"
OUT = os.path.dirname(os.path.abspath(__file__))

PARAM_SETS = {
    "grasp": dict(pyr_scale=0.5, levels=3, winsize=15, iterations=3, poly_n=5, poly_sigma=1.2, flags=0),
    "autodriving_txt": dict(pyr_scale=0.6, levels=3, winsize=3, iterations=3, poly_n=10, poly_sigma=1.05, flags=0),
    "tabletennis": dict(pyr_scale=0.6, levels=3, winsize=4, iterations=2, poly_n=1, poly_sigma=1.05, flags=0),
    # OPTFLOW_FARNEBACK_GAUSSIAN (SURVEY.md 8(a) S2-e'): no script selects it; pinned so the drop-in covers the flag
    "grasp_gaussian": dict(pyr_scale=0.5, levels=3, winsize=15, iterations=3, poly_n=5, poly_sigma=1.2,
                           flags=cv2.OPTFLOW_FARNEBACK_GAUSSIAN),
    "gaussian_w10": dict(pyr_scale=0.6, levels=2, winsize=10, iterations=2, poly_n=7, poly_sigma=1.5,
                         flags=cv2.OPTFLOW_FARNEBACK_GAUSSIAN),
}


def load_reference_functions(script, consts):
    """exec only the named top-level functions of a reference script, with the script's constants."""
    src = open(os.path.join(REF, script), encoding="utf-8").read()
    tree = ast.parse(src)
    want = {"update_transition_pic", "process_separate_regions", "process_merged_region", "opticalFlow3D"}
    body = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in want]
    mod = ast.Module(body=body, type_ignores=[])
    import time
    from numba import njit
    ns = dict(cv2=cv2, np=np, time=time, njit=njit, print=lambda *a, **k: None, mem_cal_times=[], mem_velocity_times=[],
              mem_opticalflow_times=[])
    ns.update(consts)
    exec(compile(mod, script, "exec"), ns)
    return ns


def synth_pair(rng, h, w, dx=3.0, dy=-1.5):
    base = rng.integers(0, 256, (h + 40, w + 40)).astype(np.float32)
    base = cv2.GaussianBlur(base, (0, 0), 3)
    base = (base - base.min()) / (base.max() - base.min()) * 255
    a = base[20:20 + h, 20:20 + w]
    M = np.float32([[1, 0, dx], [0, 1, dy]])
    b = cv2.warpAffine(base, M, (w + 40, h + 40), flags=cv2.INTER_LINEAR)[20:20 + h, 20:20 + w]
    return np.ascontiguousarray(a.astype(np.uint8)), np.ascontiguousarray(b.astype(np.uint8))


def grey(path):
    return cv2.cvtColor(cv2.imread(path), cv2.COLOR_RGB2GRAY)  # BGR read as RGB, as optical_flow_seg.py:439-443


def make_farneback():
    rng = np.random.default_rng(20260411)
    out = {}
    cases = []
    for (h, w) in [(96, 128), (75, 101), (40, 33), (17, 260)]:
        cases.append((f"synth_{h}x{w}",) + synth_pair(rng, h, w))
    g40, g41 = grey(f"{REF}/data/grasp/RGB/41.jpg"), grey(f"{REF}/data/grasp/RGB/42.jpg")
    cases.append(("grasp_crop", np.ascontiguousarray(g40[700:860, 400:590]), np.ascontiguousarray(g41[700:860, 400:590])))
    a0, a1 = grey(f"{REF}/data/autodriving/RGB/20.jpg"), grey(f"{REF}/data/autodriving/RGB/21.jpg")
    cases.append(("autodriving_crop", np.ascontiguousarray(a0[300:431, 250:420]), np.ascontiguousarray(a1[300:431, 250:420])))
    names = []
    for name, a, b in cases:
        names.append(name)
        out[f"{name}__prev"] = a
        out[f"{name}__next"] = b
        for pname, p in PARAM_SETS.items():
            out[f"{name}__{pname}"] = cv2.calcOpticalFlowFarneback(a, b, None, **p)
    out["cases"] = np.array(names)
    out["cv2_version"] = np.array(cv2.__version__)
    np.savez_compressed(os.path.join(OUT, "farneback_cv2.npz"), **out)
    print("farneback_cv2.npz", len(names), "cases")


def make_farneback_initflow():
    """OPTFLOW_USE_INITIAL_FLOW (cv2 flag 4; no script passes a flow): cv2 itself on the pairs of farneback_cv2.npz with
    a seeded smooth initial flow.  96x128 takes INTER_AREA's integer-ratio path (1/8), 75x101 the general one,
    40x33 has a single scale (the initial flow is used in place)."""
    fb = np.load(os.path.join(OUT, "farneback_cv2.npz"))
    rng = np.random.default_rng(20261105)
    out = {}
    sets = {
        "grasp_init": dict(PARAM_SETS["grasp"], flags=4, iterations=1),
        "gaussian_init": dict(PARAM_SETS["gaussian_w10"], flags=4 | 256),
    }
    names = ["synth_96x128", "synth_75x101", "synth_40x33"]
    for name in names:
        a, b = fb[f"{name}__prev"], fb[f"{name}__next"]
        f0 = rng.normal(size=a.shape + (2,)).astype(np.float32) * 2
        f0 = cv2.GaussianBlur(f0, (0, 0), 5) * 4 + np.float32([2.0, -1.0])
        out[f"{name}__flow0"] = f0.astype(np.float32)
        for pname, p in sets.items():
            out[f"{name}__{pname}"] = cv2.calcOpticalFlowFarneback(a, b, f0.astype(np.float32).copy(), **p)
    out["cases"] = np.array(names)
    out["cv2_version"] = np.array(cv2.__version__)
    np.savez_compressed(os.path.join(OUT, "farneback_initflow_cv2.npz"), **out)
    print("farneback_initflow_cv2.npz", len(names), "cases")


def make_farneback_f32():
    """cv2 accepts CV_32F frames as well as CV_8U (SURVEY.md 8(b) "Input conventions"); no reference call site passes
    them.  Non-integer grey values so that the float path cannot be served by rounding to uint8."""
    rng = np.random.default_rng(20261106)
    out, names = {}, []
    for (h, w) in [(96, 128), (75, 101)]:
        a, b = synth_pair(rng, h, w)
        af = (a.astype(np.float32) + rng.random((h, w)).astype(np.float32) * 0.9).astype(np.float32)
        bf = (b.astype(np.float32) + rng.random((h, w)).astype(np.float32) * 0.9).astype(np.float32)
        name = f"f32_{h}x{w}"
        names.append(name)
        out[f"{name}__prev"], out[f"{name}__next"] = af, bf
        for pname in ("grasp", "autodriving_txt"):
            out[f"{name}__{pname}"] = cv2.calcOpticalFlowFarneback(af, bf, None, **PARAM_SETS[pname])
    out["cases"] = np.array(names)
    out["cv2_version"] = np.array(cv2.__version__)
    np.savez_compressed(os.path.join(OUT, "farneback_f32_cv2.npz"), **out)
    print("farneback_f32_cv2.npz", names)


def make_hostio():
    """`py_cpu_nms` of optical_flow_ob.py:96-135 (source-sliced) on seeded box sets, for hostio.nms."""
    src = open(os.path.join(REF, "optical_flow_ob.py"), encoding="utf-8").read()
    tree = ast.parse(src)
    body = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "py_cpu_nms"]
    ns = dict(np=np)
    exec(compile(ast.Module(body=body, type_ignores=[]), "optical_flow_ob.py", "exec"), ns)
    rng = np.random.default_rng(5)
    out = {}
    for k, n in enumerate([1, 2, 7, 40, 120]):
        xy = rng.integers(0, 300, (n, 2))
        wh = rng.integers(5, 120, (n, 2))
        dets = np.concatenate([xy, xy + wh, rng.integers(500, 9000, (n, 1))], axis=1).astype(np.float64)
        if n >= 7:
            dets[1, :4] = dets[0, :4]            # an exact duplicate box, and a tie in the scores
            dets[2, 4] = dets[3, 4]
        for th in (0.2, 0.5):
            out[f"nms__{k}__dets"] = dets
            out[f"nms__{k}__keep_{th}"] = np.array(ns["py_cpu_nms"](dets, th), dtype=np.int64)
    out["n_cases"] = np.array(5)
    np.savez_compressed(os.path.join(OUT, "hostio_reference.npz"), **out)
    print("hostio_reference.npz")


def make_roi():
    out = {}
    for ds, script, consts, frame_hw in [
        ("grasp", "optical_flow_seg.py",
         dict(MEMSIZE=80, OFFSET=0, EXTEND_HEIGHT_UPPER=20, EXTEND_HEIGHT_LOWER=20, EXTEND_WIDTH_LEFT=20,
              EXTEND_WIDTH_RIGHT=20, THRES=250, CONNECT=4), (1920, 1080)),
        ("autodriving", "optical_flow_yolo.py",
         dict(MEMSIZE=200, OFFSET=15, EXTEND_HEIGHT_UPPER=60, EXTEND_HEIGHT_LOWER=60, EXTEND_WIDTH_LEFT=60,
              EXTEND_WIDTH_RIGHT=60, THRES=114, CONNECT=4), (801, 801)),
        # the three datasets no script is checked in for: constants of data/<ds>/Parameters.txt, functions of
        # optical_flow_seg.py (the scripts differ in their constants only for this part)
        ("uav", "optical_flow_seg.py",
         dict(MEMSIZE=40, OFFSET=0, EXTEND_HEIGHT_UPPER=30, EXTEND_HEIGHT_LOWER=30, EXTEND_WIDTH_LEFT=30,
              EXTEND_WIDTH_RIGHT=30, THRES=114, CONNECT=4), (161, 161)),
        ("uavnew2", "optical_flow_seg.py",
         dict(MEMSIZE=40, OFFSET=0, EXTEND_HEIGHT_UPPER=60, EXTEND_HEIGHT_LOWER=60, EXTEND_WIDTH_LEFT=60,
              EXTEND_WIDTH_RIGHT=60, THRES=245, CONNECT=4), (600, 600)),
        ("tabletennis", "optical_flow_seg.py",
         dict(MEMSIZE=10, OFFSET=0, EXTEND_HEIGHT_UPPER=20, EXTEND_HEIGHT_LOWER=20, EXTEND_WIDTH_LEFT=20,
              EXTEND_WIDTH_RIGHT=20, THRES=245, CONNECT=4), (160, 160)),
    ]:
        mat = scipy.io.loadmat(f"{REF}/data/{ds}/constructed_3D_matrix.mat")["constructed3DMatrix"]
        out[f"{ds}__currents"] = np.ascontiguousarray(np.moveaxis(mat, 2, 0)).astype(np.float64)  # (T, Hc, Wc)
        T = mat.shape[2]
        H, W = frame_hw
        blank = np.zeros((H, W), np.uint8)
        # memimg as the script computes it (optical_flow_seg.py:426-435)
        with np.errstate(divide="ignore", invalid="ignore"):
            u8 = np.clip(-3366 / np.log10(np.moveaxis(mat, 2, 0).astype(np.double)) - 306, 0, 255).astype(np.uint8)
        out[f"{ds}__memimg"] = u8
        for flag in (1, 2):
            # farneback on blank frames is irrelevant here: stub cv2.calcOpticalFlowFarneback out of the namespace
            ns = load_reference_functions(script, dict(consts, FLAG=flag, farneback_params={}))
            fake_cv2 = types.SimpleNamespace(**{k: getattr(cv2, k) for k in dir(cv2) if k.startswith("CC_STAT") or k == "connectedComponentsWithStats"})
            fake_cv2.calcOpticalFlowFarneback = lambda a, b, f, **kw: np.zeros(a.shape + (2,), np.float32)
            ns["cv2"] = fake_cv2
            rects = np.zeros((T, 64, 4), np.int32)
            nrect = np.zeros(T, np.int32)
            nlab = np.zeros(T, np.int32)
            for t in range(T):
                res = ns["opticalFlow3D"](u8[t], u8[t], blank, blank, consts["MEMSIZE"], consts["MEMSIZE"])
                if flag == 1:
                    nlab[t] = res[4]
                    info = res[5]
                    nrect[t] = len(info)
                    for j, r in enumerate(info):
                        rects[t, j] = r
                else:
                    r = res[4]
                    if r != (0, 0, 0, 0):
                        nrect[t] = 1
                        rects[t, 0] = r
            out[f"{ds}__flag{flag}__rects"] = rects
            out[f"{ds}__flag{flag}__nrect"] = nrect
            if flag == 1:
                out[f"{ds}__num_labels"] = nlab
        # cv2 CC labels + stats on the transition picture
        labels = np.zeros(u8.shape, np.int32)
        stats = np.zeros((T, 64, 5), np.int32)
        for t in range(T):
            tp = np.where(u8[t] >= consts["THRES"], 255, 0).astype(np.uint8)
            n, lab, st, _ = cv2.connectedComponentsWithStats(tp, connectivity=4)
            labels[t] = lab
            stats[t, :n - 1] = st[1:]
        out[f"{ds}__labels"] = labels
        out[f"{ds}__stats"] = stats
    # random masks: CC numbering/stats pin against cv2 on denser patterns than the datasets offer
    rng = np.random.default_rng(7)
    masks = (rng.random((40, 24, 13)) < rng.uniform(0.1, 0.7, (40, 1, 1))).astype(np.uint8) * 255
    lab = np.zeros(masks.shape, np.int32)
    st = np.zeros((40, 200, 5), np.int32)
    nl = np.zeros(40, np.int32)
    for i in range(40):
        n, l, s, _ = cv2.connectedComponentsWithStats(masks[i], connectivity=4)
        lab[i] = l
        st[i, :n - 1] = s[1:]
        nl[i] = n
    out["random__masks"], out["random__labels"], out["random__stats"], out["random__num_labels"] = masks, lab, st, nl
    np.savez_compressed(os.path.join(OUT, "roi_reference.npz"), **out)
    print("roi_reference.npz")


def make_events():
    sys.modules.setdefault("h5py", types.ModuleType("h5py"))
    sys.path.insert(0, f"{REF}/eventsim")
    import event_mem_sim as ems
    x, y, p, t = ems.generate_synthetic_events()
    H, W = 240, 320
    out = dict(x=x.astype(np.int16), y=y.astype(np.int16), p=p.astype(np.int8), t=t.astype(np.int64), H=H, W=W)
    slices = list(ems.slice_indices(t, 1000))
    out["slice_bounds"] = np.array([s.start for s in slices] + [slices[-1].stop], np.int64)
    # bincount of a few slices
    pick = [0, 1, 100, 500, len(slices) - 1]
    out["hist_slices"] = np.array(pick)
    out["hist"] = np.stack([ems.bincount_2d(x[slices[i]], y[slices[i]], H, W) for i in pick])
    # scheme 1 exactly as simulate() runs it (event_mem_sim.py:204-228), CLI defaults active_v=-6, silent_v=0
    w = ems.PARAMS["wini"] * np.ones((H, W), dtype=np.float32)
    snaps = {}
    for i, sl in enumerate(slices):
        V = np.full((H, W), 0.0, dtype=np.float32)
        if sl.stop > sl.start:
            ec = ems.bincount_2d(x[sl], y[sl], H, W)
            m = ec >= ems.THETA_EVENTS
            if m.any():
                V[m] = -6.0
        w = ems.update_state(w, V)
        if i in (0, 31, 32, 200, len(slices) - 1):
            snaps[i] = w.copy()
    out["w_snap_idx"] = np.array(sorted(snaps))
    out["w_snaps"] = np.stack([snaps[i] for i in sorted(snaps)])
    out["R_final"] = ems.resistance_exp(w)
    # a decaying variant: silent_v = 0.5 (> von) so every pixel updates every slice; first 64 slices
    w2 = ems.PARAMS["wini"] * np.ones((H, W), dtype=np.float32)
    for i, sl in enumerate(slices[:64]):
        V = np.full((H, W), 0.5, dtype=np.float32)
        if sl.stop > sl.start:
            ec = ems.bincount_2d(x[sl], y[sl], H, W)
            V[ec >= 1] = -6.0
        w2 = ems.update_state(w2, V)
    out["w_decay64"] = w2
    # single update_state call on a random state/voltage field
    rng = np.random.default_rng(3)
    w0 = rng.random((64, 96), dtype=np.float32)
    V0 = rng.uniform(-8, 2, (64, 96)).astype(np.float32)
    out["us_w0"], out["us_V0"], out["us_w1"] = w0, V0, ems.update_state(w0, V0)
    np.savez_compressed(os.path.join(OUT, "events_reference.npz"), **out)
    print("events_reference.npz", len(x), "events", len(slices), "slices")


def make_events_scheme2():
    """Scheme 2 (DC bias + event overlay with a per-pixel refractory period) by running the reference's own
    simulate() (event_mem_sim.py:164-322) with load_events patched to hand it the synthetic stream."""
    import tempfile
    from pathlib import Path
    sys.modules.setdefault("h5py", types.ModuleType("h5py"))
    sys.path.insert(0, f"{REF}/eventsim")
    import event_mem_sim as ems
    x, y, p, t = ems.generate_synthetic_events()
    H, W = 240, 320
    p01 = np.where(p > 0, 1, 0)  # a Prophesee file stores OFF as 0; the synthetic generator writes -1
    cases = {  # name: (polarity array, slice_us, active_v, silent_v (bias), polarity mode, number of events used)
        "split_pm1": (p, 1000, -6.0, 0.0, "split", len(x)),
        "split_01_decay": (p01, 500, -6.0, 0.3, "split", 12000),
        "magnitude": (p01, 700, -6.0, 0.3, "magnitude", 12000),
    }
    out = dict(x=x.astype(np.int16), y=y.astype(np.int16), t=t.astype(np.int64), p_pm1=p.astype(np.int8),
               p_01=p01.astype(np.int8), H=H, W=W, cases=np.array(list(cases)))
    for name, (pp, slice_us, av, sv, mode, n) in cases.items():
        ems.load_events = lambda path, pp=pp, n=n: (x[:n], y[:n], pp[:n].astype(int), t[:n], H, W)
        with tempfile.TemporaryDirectory() as d:
            h5 = Path(d) / "ev.hdf5"
            ems.simulate(h5, version=2, slice_us=slice_us, active_v=av, silent_v=sv, save_video=False, polarity=mode)
            out[f"{name}__w_a"] = np.load(h5.with_suffix(".V2.npz"))["w_final"]
            if mode == "split":
                out[f"{name}__w_b"] = np.load(h5.with_suffix(".V2_b.npz"))["w_final"]
        out[f"{name}__args"] = np.array([slice_us, av, sv, n, 0 if mode == "split" else 1], dtype=np.float64)
        print(name, "w_a range", out[f"{name}__w_a"].min(), out[f"{name}__w_a"].max())
    np.savez_compressed(os.path.join(OUT, "events_scheme2_reference.npz"), **out)
    print("events_scheme2_reference.npz")


def make_segment():
    """Masks of the reference's own `process_flow_region` (optical_flow_seg.py:322-357, source-sliced) on
    `cv2.cartToPolar` of float64 flow regions, exactly as `task_results` calls it (:308-311): the flow is a float64
    array holding negated float32 Farneback output (:215, :461)."""
    src = open(os.path.join(REF, "optical_flow_seg.py"), encoding="utf-8").read()
    tree = ast.parse(src)
    body = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "process_flow_region"]
    ns = dict(cv2=cv2, np=np, SEG_TH=1)  # SEG_TH: optical_flow_seg.py:49
    exec(compile(ast.Module(body=body, type_ignores=[]), "optical_flow_seg.py", "exec"), ns)
    ref = ns["process_flow_region"]
    rng = np.random.default_rng(77)
    cases = {}
    # (1) real Farneback flow of a moving textured patch: a compact blob with a ragged rim around |flow| = 1
    a, b = synth_pair(rng, 150, 200, dx=0.0, dy=0.0)
    patch = rng.integers(0, 256, (60, 80)).astype(np.float32)
    patch = cv2.GaussianBlur(patch, (0, 0), 1.5)
    patch = ((patch - patch.min()) / (patch.max() - patch.min()) * 255).astype(np.uint8)
    a = a.copy(); b = b.copy()
    a[40:100, 50:130] = patch
    b[42:102, 53:133] = patch
    cases["farneback_patch"] = cv2.calcOpticalFlowFarneback(a, b, None, **PARAM_SETS["grasp"])
    # (2) smooth random fields, about 12 % of the pixels above the threshold (components, holes, gaps narrower and
    # wider than the element), quantised to multiples of 1/32 (stored as int8; plenty of magnitudes exactly 1)
    for name, (h, w, sig) in {"noise_97x131": (97, 131, 4.0), "noise_64x64": (64, 64, 2.0), "noise_33x65": (33, 65, 3.0),
                              "noise_200x31": (200, 31, 2.5), "noise_7x9": (7, 9, 1.0), "noise_1x40": (1, 40, 1.0),
                              "noise_40x1": (40, 1, 1.0), "noise_1x1": (1, 1, 1.0), "noise_420x560": (420, 560, 6.0)}.items():
        f = rng.standard_normal((h + 40, w + 40, 2)).astype(np.float32)
        f = cv2.GaussianBlur(f, (0, 0), sig)[20:20 + h, 20:20 + w].reshape(h, w, 2)
        f = f / (np.abs(f).mean() + 1e-9) * 0.39
        q = np.clip(np.rint(f * 32), -128, 127).astype(np.int8)
        cases[name] = q
    # (3) magnitudes exactly at and next to the threshold, isolated pixels at the corners and edges, NaN / Inf
    f = np.zeros((50, 70, 2), np.float32)
    f[0, 0] = (1.0, 0.0)                                   # == 1: not set
    f[0, 69] = (np.nextafter(np.float32(1), np.float32(2)), 0.0)
    f[49, 0] = (0.6, 0.8)                                  # float32(0.6)^2 + float32(0.8)^2 in double
    f[49, 69] = (-3.0, 4.0)
    f[25, 35] = (0.0, -1.0000001)
    f[10, 20] = (np.nan, 5.0)
    f[12, 40] = (np.inf, 0.0)
    f[30, 5:64] = (0.70710677, 0.70710677)                 # float32(sqrt(.5)): squares sum to just below / above 1
    f[31, 5:64] = (0.7071068, 0.7071068)
    cases["edge_values"] = f
    out = {"cases": np.array(sorted(cases))}
    for name, stored in cases.items():
        f32 = stored.astype(np.float32) / 32 if stored.dtype == np.int8 else stored
        flow = -np.asarray(f32, dtype=np.float64)          # the reference's float64 flow after `flow = -flow`
        mag, ang = cv2.cartToPolar(flow[..., 0], flow[..., 1])
        m = ref(mag, ang)
        assert m.dtype == np.uint8 and m.shape == f32.shape[:2] and set(np.unique(m)) <= {0, 255}
        out[f"{name}__flow"] = stored                       # int8: flow = stored / 32
        out[f"{name}__mask"] = np.packbits(m > 0, axis=1)
        out[f"{name}__mag_gt"] = np.packbits(mag > 1, axis=1)
    ks = [(10, 10), (3, 3), (5, 7), (7, 5), (1, 1), (2, 2), (4, 9), (32, 32), (31, 17), (15, 15)]
    for shape in (0, 1, 2):
        for kw, kh in ks:
            out[f"element__{shape}_{kw}_{kh}"] = cv2.getStructuringElement(shape, (kw, kh))
    # dilate / erode alone (one pass each) with other elements, on the noise field
    base = (np.sqrt(((cases["noise_97x131"].astype(np.float64) / 32) ** 2).sum(-1)) > 1).astype(np.uint8) * 255
    for shape, (kw, kh) in [(2, (10, 10)), (0, (3, 5)), (1, (7, 7)), (2, (4, 9)), (2, (31, 17)), (0, (32, 32))]:
        el = cv2.getStructuringElement(shape, (kw, kh))
        out[f"dilate__{shape}_{kw}_{kh}"] = np.packbits(cv2.dilate(base, el) > 0, axis=1)
        out[f"erode__{shape}_{kw}_{kh}"] = np.packbits(cv2.erode(base, el) > 0, axis=1)
    out["cv2_version"] = np.array(cv2.__version__)
    np.savez_compressed(os.path.join(OUT, "segment_reference.npz"), **out)
    print("segment_reference.npz", {k: (int(np.unpackbits(out[f"{k}__mag_gt"]).sum()), int(np.unpackbits(out[f"{k}__mask"]).sum()),
                                          cases[k].shape[0] * cases[k].shape[1]) for k in cases})


def make_frames():
    """cv2.cvtColor(COLOR_RGB2GRAY) as the scripts call it (optical_flow_seg.py:439-443) on crops of real frames, and
    the reference's own prediction `task_results` (optical_flow_prediction.py:257-361, source-sliced) on flow fields."""
    import glob
    import time
    src = open(os.path.join(REF, "optical_flow_prediction.py"), encoding="utf-8").read()
    tree = ast.parse(src)
    body = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "task_results"]
    ns = dict(cv2=cv2, np=np, time=time, PADDING=10, mem_combination_times=[], mem_task_times=[])
    exec(compile(ast.Module(body=body, type_ignores=[]), "optical_flow_prediction.py", "exec"), ns)
    ref_task = ns["task_results"]
    rng = np.random.default_rng(99)
    out = {}
    files = sorted(glob.glob(os.path.join(REF, "data", "grasp", "RGB", "*.jpg")))[:2]
    frames = [cv2.imread(f) for f in files]
    # grey: two real crops + a random frame with odd width (every residue of the 4-pixel groups)
    greys = {"grasp_a": frames[0][300:420, 200:457], "grasp_b": frames[1][800:899, 500:763],
             "random": rng.integers(0, 256, (37, 131, 3), dtype=np.uint8)}
    for k, f in greys.items():
        f = np.ascontiguousarray(f)
        out[f"grey__{k}__in"] = f
        out[f"grey__{k}__out"] = cv2.cvtColor(f, cv2.COLOR_RGB2GRAY)
    # prediction: a real frame crop as next_frame, smooth + noisy flow, regions touching the borders (replicate)
    nf = np.ascontiguousarray(frames[1][600:840, 300:620])      # 240 x 320 x 3
    H, W = nf.shape[:2]
    smooth = cv2.GaussianBlur(rng.standard_normal((H, W, 2)).astype(np.float32) * 40, (0, 0), 9)
    flow = (smooth + rng.standard_normal((H, W, 2)).astype(np.float32) * 0.5).astype(np.float32)
    flow[5, 7] = (0.5, -0.5)          # ties of the 1/32 rounding
    flow[6, 8] = (1.0 / 64, 3.0 / 64)
    flow[0, 0] = (-30.0, -30.0)       # far outside: replicate
    flow[H - 1, W - 1] = (1e4, 1e4)
    out["pred__next"] = nf
    out["pred__flow"] = flow
    rects = {"whole": (0, 0, W, H), "inner": (37, 21, 250, 190), "corner": (200, 150, W, H), "thin": (10, 100, 11, 180)}
    for k, r in rects.items():
        got = ref_task(nf, nf, flow, 2, r, 2, True)
        assert got.shape == nf.shape and got.dtype == np.uint8
        out[f"pred__{k}__rect"] = np.array(r)
        out[f"pred__{k}__out"] = got
    # EST_FLAG 1, merged, PADDING 10 (optical_flow_prediction.py:269-300)
    regs = [(60, 40, 120, 90), (150, 100, 230, 170)]
    out["pred__merge__regions"] = np.array(regs)
    out["pred__merge__out"] = ref_task(nf, nf, flow, 3, regs, 1, True)
    out["cv2_version"] = np.array(cv2.__version__)
    np.savez_compressed(os.path.join(OUT, "frames_reference.npz"), **out)
    print("frames_reference.npz", {k: v.shape for k, v in out.items() if hasattr(v, "shape") and v.ndim >= 2})


def _script_namespace(script, names, consts):
    """Top-level functions `names` of a reference script (source-sliced, unmodified) in one namespace together with
    the script's module-level constants `consts` and its timing-list globals."""
    import time
    from numba import njit
    src = open(os.path.join(REF, script), encoding="utf-8").read()
    tree = ast.parse(src)
    body = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in names]
    assert {n.name for n in body} == set(names), (script, names)
    ns = dict(cv2=cv2, np=np, time=time, njit=njit, print=lambda *a, **k: None, mem_cal_times=[], mem_velocity_times=[],
              mem_opticalflow_times=[], mem_task_times=[], mem_combination_times=[])
    ns.update(consts)
    exec(compile(ast.Module(body=body, type_ignores=[]), script, "exec"), ns)
    return ns


REALDATA = {
    # dataset: (script, constants of that script, pairs kept with their crops, pairs kept with FULL frames)
    "grasp": ("optical_flow_seg.py",
              dict(MEMSIZE=80, OFFSET=0, EXTEND_HEIGHT_UPPER=20, EXTEND_HEIGHT_LOWER=20, EXTEND_WIDTH_LEFT=20,
                   EXTEND_WIDTH_RIGHT=20, THRES=250, CONNECT=4, FLAG=2, PADDING=20, SEG_TH=1,
                   farneback_params=PARAM_SETS["grasp"]), "12 consecutive", 1),
    "autodriving": ("optical_flow_yolo.py",
                    dict(MEMSIZE=200, OFFSET=15, EXTEND_HEIGHT_UPPER=60, EXTEND_HEIGHT_LOWER=60, EXTEND_WIDTH_LEFT=60,
                         EXTEND_WIDTH_RIGHT=60, THRES=114, CONNECT=4, FLAG=2, PADDING=60, SEG_TH=1,
                         farneback_params=PARAM_SETS["grasp"]), "all non-empty", 2),
}


def make_realdata(which=("grasp", "autodriving")):
    """configs[0] / configs[1] end to end: the body of the scripts' main loop (optical_flow_seg.py:413-545,
    optical_flow_yolo.py:752-797 + the segmentation task the yolo script shares with the seg script) driven over the
    shipped sequences with the reference's OWN functions (source-sliced: opticalFlow3D, process_merged_region,
    update_transition_pic, task_results, process_flow_region, calculate_pixel_accuracy).  The only changes against the
    scripts: the Windows-only sort key (`x.split('\\')`, optical_flow_seg.py:402) becomes os.path.basename, nothing is
    written to disk, and YOLO (absent weights) is skipped.  The `mem_state2 = mem_state1` line (:435) is kept, so pair
    i is gated by slice OFFSET + i.

    Per dataset -> tests/golden/realdata_<ds>.npz:
      all pairs   : rects, accuracy1 ("our method"), accuracy2 (full-frame baseline), dataset averages
      kept pairs  : memimg, rectangle, the grey frames cropped to the rectangles that read them (the flow inside a
                    rectangle depends on nothing else), cv2's flow of the crop at every 8th pixel of every 8th row
                    (+ |flow| checksum; the GPU box has cv2 and compares every pixel live), the reference's motion_binary
                    (bit-packed), the ground-truth mask (bit-packed), accuracy1
      full pairs  : both whole grey frames, the baseline's motion_binary1 and accuracy2 (full-frame Farneback path,
                    optical_flow_seg.py:492-545)
    """
    for ds in which:
        script, consts, _, n_full = REALDATA[ds]
        ns = _script_namespace(script, ["update_transition_pic", "process_separate_regions", "process_merged_region",
                                        "opticalFlow3D", "task_results", "process_flow_region",
                                        "calculate_pixel_accuracy"], consts)
        base = f"{REF}/data/{ds}"
        imgs = open(f"{base}/imgs.txt").read().splitlines()
        mem_state = scipy.io.loadmat(f"{base}/constructed_3D_matrix.mat")["constructed3DMatrix"]
        rgbimages = sorted([os.path.join(base, "RGB", i) for i in imgs],
                           key=lambda x: int(os.path.basename(x).split(".")[0]))
        M, OFFSET = consts["MEMSIZE"], consts["OFFSET"]
        n_pairs = len(imgs) - 2
        rects = np.zeros((n_pairs, 4), np.int32)
        acc1 = np.zeros(n_pairs)
        acc2 = np.zeros(n_pairs)
        per_pair = {}
        kernel = cv2.getStructuringElement(cv2.MORPH_ELLIPSE, (10, 10))
        for i in range(n_pairs):
            mem_state1 = mem_state[:, :, OFFSET + i].astype(np.double)
            with np.errstate(divide="ignore", invalid="ignore"):
                mem_state1 = -3366 / np.log10(mem_state1) - 306
            mem_state1 = np.clip(mem_state1, 0, 255).astype(np.uint8)
            memimg1 = memimg2 = mem_state1.astype(np.uint8)          # optical_flow_seg.py:435
            prev_frame, next_frame = cv2.imread(rgbimages[i]), cv2.imread(rgbimages[i + 1])
            prev_gray = cv2.cvtColor(prev_frame, cv2.COLOR_RGB2GRAY)
            next_gray = cv2.cvtColor(next_frame, cv2.COLOR_RGB2GRAY)
            image = cv2.imread(f"{base}/gtmask/{imgs[i + 1]}")
            gray_image = cv2.cvtColor(image, cv2.COLOR_BGR2GRAY)
            _, binary_image = cv2.threshold(gray_image, 127, 256, cv2.THRESH_BINARY)
            for k in ("mem_cal_times", "mem_velocity_times", "mem_opticalflow_times", "mem_task_times",
                      "mem_combination_times"):
                ns[k].clear()
            flow, _, _, _, (x0, y0, x1, y1) = ns["opticalFlow3D"](memimg1, memimg2, prev_gray, next_gray, M, M)
            flow = -flow
            num_labels = 1 if (x0, y0, x1, y1) == (0, 0, 0, 0) else 2
            motion_binary = ns["task_results"](prev_frame, next_frame, flow, num_labels, (x0, y0, x1, y1),
                                               EST_FLAG=2, MERGE_FLAG=True)
            # baseline block (optical_flow_seg.py:492-539): only what reaches motion_binary1
            flow1 = -cv2.calcOpticalFlowFarneback(prev_gray, next_gray, None, **consts["farneback_params"])
            mag, _ = cv2.cartToPolar(flow1[..., 0], flow1[..., 1])
            mm = np.zeros(prev_frame.shape, np.uint8)
            mm[mag > consts["SEG_TH"]] = 255
            for _ in range(5):
                mm = cv2.erode(cv2.dilate(mm, kernel), kernel)
            _, motion_binary1 = cv2.threshold(cv2.cvtColor(mm, cv2.COLOR_BGR2GRAY), 1, 255, cv2.THRESH_BINARY)
            rects[i] = (x0, y0, x1, y1)
            acc1[i] = ns["calculate_pixel_accuracy"](motion_binary, binary_image)
            acc2[i] = ns["calculate_pixel_accuracy"](motion_binary1, binary_image)
            per_pair[i] = dict(memimg=memimg2, prev_gray=prev_gray, next_gray=next_gray, gt=binary_image,
                               motion=motion_binary, motion1=motion_binary1, flow=-flow)
            if ds == "grasp" and i >= 60:   # keep memory bounded: only the candidates below are used
                per_pair[i] = None
            print(ds, i, rects[i], f"{acc1[i]:.6f} {acc2[i]:.6f}", flush=True)
        area = (rects[:, 2] - rects[:, 0]) * (rects[:, 3] - rects[:, 1])
        if ds == "grasp":
            keep = list(range(40, 52))                               # 12 consecutive pairs, 41.jpg .. 53.jpg
        else:
            keep = [i for i in range(n_pairs) if area[i] > 0]        # every pair with a non-empty ROI
        full = keep[:n_full]
        H, W = per_pair[keep[0]]["prev_gray"].shape
        out = dict(frame_hw=np.array([H, W]), rects=rects, acc1=acc1, acc2=acc2, acc1_mean=acc1.mean(),
                   acc2_mean=acc2.mean(), keep=np.array(keep), full=np.array(full), cv2_version=np.array(cv2.__version__))
        # grey frames: frame t serves pair t-1 (as next) and pair t (as prev); stored once, cropped to the union of the
        # rectangles that read it (outside them nothing reaches the ROI path)
        boxes = {}
        for i in keep:
            for t in (i, i + 1):
                b = boxes.get(t)
                r = rects[i]
                boxes[t] = tuple(r) if b is None else (min(b[0], r[0]), min(b[1], r[1]), max(b[2], r[2]), max(b[3], r[3]))
        for t, (bx0, by0, bx1, by1) in boxes.items():
            src = per_pair[t]["prev_gray"] if per_pair.get(t) else per_pair[t - 1]["next_gray"]
            out[f"frame__{t}"] = np.ascontiguousarray(src[by0:by1, bx0:bx1])
            out[f"fbox__{t}"] = np.array([bx0, by0, bx1, by1], np.int32)
        for i in keep:
            d = per_pair[i]
            x0, y0, x1, y1 = (int(v) for v in rects[i])
            fl = d["flow"][y0:y1, x0:x1].astype(np.float32)          # cv2's float32 crop flow (pasted into float64)
            assert np.array_equal(fl.astype(np.float64), d["flow"][y0:y1, x0:x1])
            out[f"memimg__{i}"] = d["memimg"]
            out[f"flow8__{i}"] = np.ascontiguousarray(fl[::8, ::8])
            out[f"flowsum__{i}"] = np.array([np.abs(fl.astype(np.float64)).sum(), float(np.abs(fl).max())])
            out[f"motion__{i}"] = np.packbits(d["motion"] > 0, axis=1)
            out[f"gt__{i}"] = np.packbits(d["gt"] > 0, axis=1)
            assert set(np.unique(d["gt"])) <= {0, 255} and set(np.unique(d["motion"])) <= {0, 255}
        for i in full:
            d = per_pair[i]
            out[f"fullprev__{i}"] = d["prev_gray"]
            out[f"fullnext__{i}"] = d["next_gray"]
            out[f"motion1__{i}"] = np.packbits(d["motion1"] > 0, axis=1)
        path = os.path.join(OUT, f"realdata_{ds}.npz")
        np.savez_compressed(path, **out)
        print(f"realdata_{ds}.npz: {len(keep)} pairs kept, {len(full)} with full frames, {os.path.getsize(path) / 1e6:.1f} MB; "
              f"our method average accuracy {acc1.mean()!r}, original {acc2.mean()!r}")


REALDATA_SMALL = {
    # The three sequences no script is checked in for.  Constants = data/<ds>/Parameters.txt (which names no OFFSET:
    # 0 is used, every slice index stays inside the .mat); functions = optical_flow_seg.py, whose main loop honours
    # FLAG 1 as well as FLAG 2.  Last entry: which pairs with a non-empty gate are kept (every k-th).
    "uav": (dict(MEMSIZE=40, OFFSET=0, EXTEND_HEIGHT_UPPER=30, EXTEND_HEIGHT_LOWER=30, EXTEND_WIDTH_LEFT=30,
                 EXTEND_WIDTH_RIGHT=30, THRES=114, CONNECT=4, FLAG=1, PADDING=30, SEG_TH=1,
                 farneback_params=PARAM_SETS["autodriving_txt"]), 4),
    "uavnew2": (dict(MEMSIZE=40, OFFSET=0, EXTEND_HEIGHT_UPPER=60, EXTEND_HEIGHT_LOWER=60, EXTEND_WIDTH_LEFT=60,
                     EXTEND_WIDTH_RIGHT=60, THRES=245, CONNECT=4, FLAG=1, PADDING=60, SEG_TH=1,
                     farneback_params=PARAM_SETS["grasp"]), 4),
    "tabletennis": (dict(MEMSIZE=10, OFFSET=0, EXTEND_HEIGHT_UPPER=20, EXTEND_HEIGHT_LOWER=20, EXTEND_WIDTH_LEFT=20,
                         EXTEND_WIDTH_RIGHT=20, THRES=245, CONNECT=4, FLAG=2, PADDING=20, SEG_TH=1,
                         farneback_params=PARAM_SETS["tabletennis"]), 1),
}


def make_realdata_small(which=("uav", "uavnew2", "tabletennis")):
    """The main-loop body of optical_flow_seg.py:413-490 over data/uav, data/uavnew2 (FLAG 1: one Farneback call per
    connected region, pasted in label order, then task_results' merged padding box) and data/tabletennis (FLAG 2,
    poly_n = 1, even winsize), with the reference's own functions and each dataset's Parameters.txt.

    -> tests/golden/realdata_<ds>.npz: for ALL pairs the region rectangles, num_labels and accuracy1; for the kept
    pairs memimg, both whole grey frames (the frames are small), the pasted float32 flow at every 4th pixel of
    every 4th row + |flow| checksum, motion_binary and the ground-truth mask (bit-packed)."""
    for ds in which:
        consts, every = REALDATA_SMALL[ds]
        ns = _script_namespace("optical_flow_seg.py",
                               ["update_transition_pic", "process_separate_regions", "process_merged_region",
                                "opticalFlow3D", "task_results", "process_flow_region", "calculate_pixel_accuracy"], consts)
        base = f"{REF}/data/{ds}"
        imgs = open(f"{base}/imgs.txt").read().splitlines()
        mem_state = scipy.io.loadmat(f"{base}/constructed_3D_matrix.mat")["constructed3DMatrix"]
        rgbimages = sorted([os.path.join(base, "RGB", i) for i in imgs],
                           key=lambda x: int(os.path.basename(x).split(".")[0]))
        M, OFFSET, FLAG = consts["MEMSIZE"], consts["OFFSET"], consts["FLAG"]
        n_pairs = min(len(imgs) - 1, mem_state.shape[2] - OFFSET)
        rects = np.zeros((n_pairs, 32, 4), np.int32)
        nrect = np.zeros(n_pairs, np.int32)
        nlab = np.ones(n_pairs, np.int32)
        acc1 = np.zeros(n_pairs)
        per_pair = {}
        for i in range(n_pairs):
            mem_state1 = mem_state[:, :, OFFSET + i].astype(np.double)
            with np.errstate(divide="ignore", invalid="ignore"):
                mem_state1 = -3366 / np.log10(mem_state1) - 306
            memimg = np.clip(mem_state1, 0, 255).astype(np.uint8)      # :426-435, mem_state2 = mem_state1
            prev_frame, next_frame = cv2.imread(rgbimages[i]), cv2.imread(rgbimages[i + 1])
            prev_gray = cv2.cvtColor(prev_frame, cv2.COLOR_RGB2GRAY)
            next_gray = cv2.cvtColor(next_frame, cv2.COLOR_RGB2GRAY)
            image = cv2.imread(f"{base}/gtmask/{imgs[i + 1]}")
            _, binary_image = cv2.threshold(cv2.cvtColor(image, cv2.COLOR_BGR2GRAY), 127, 256, cv2.THRESH_BINARY)
            for k in ("mem_cal_times", "mem_velocity_times", "mem_opticalflow_times", "mem_task_times",
                      "mem_combination_times"):
                ns[k].clear()
            res = ns["opticalFlow3D"](memimg, memimg, prev_gray, next_gray, M, M)
            flow = -res[0]
            if FLAG == 1:
                num_labels, info = res[4], [tuple(int(v) for v in r) for r in res[5]]
                motion = ns["task_results"](prev_frame, next_frame, flow, num_labels, info, EST_FLAG=1, MERGE_FLAG=True)
            else:
                info1 = tuple(int(v) for v in res[4])
                num_labels = 1 if info1 == (0, 0, 0, 0) else 2
                motion = ns["task_results"](prev_frame, next_frame, flow, num_labels, info1, EST_FLAG=2, MERGE_FLAG=True)
                info = [] if num_labels == 1 else [info1]
            nrect[i], nlab[i] = len(info), num_labels
            for j, r in enumerate(info):
                rects[i, j] = r
            acc1[i] = ns["calculate_pixel_accuracy"](motion, binary_image)
            per_pair[i] = dict(memimg=memimg, prev_gray=prev_gray, next_gray=next_gray, gt=binary_image, motion=motion,
                               flow=-flow)
            print(ds, i, nlab[i], info[:3], f"{acc1[i]:.6f}", flush=True)
        keep = [i for i in range(n_pairs) if nrect[i] > 0][::every]
        multi = [i for i in range(n_pairs) if nrect[i] > 1 and i not in keep][:4]    # FLAG 1: several regions
        keep = sorted(keep + multi)
        H, W = per_pair[0]["prev_gray"].shape
        out = dict(frame_hw=np.array([H, W]), rects=rects, nrect=nrect, num_labels=nlab, acc1=acc1, acc1_mean=acc1.mean(),
                   keep=np.array(keep), flag=np.array(FLAG), cv2_version=np.array(cv2.__version__))
        for i in keep:
            d = per_pair[i]
            fl = d["flow"].astype(np.float32)
            assert np.array_equal(fl.astype(np.float64), d["flow"])
            out[f"memimg__{i}"] = d["memimg"]
            out[f"gray__{i}"] = d["prev_gray"]
            out[f"gray__{i + 1}"] = d["next_gray"]
            out[f"flow4__{i}"] = np.ascontiguousarray(fl[::4, ::4])
            out[f"flowsum__{i}"] = np.array([np.abs(fl.astype(np.float64)).sum(), float(np.abs(fl).max())])
            out[f"motion__{i}"] = np.packbits(d["motion"] > 0, axis=1)
            out[f"gt__{i}"] = np.packbits(d["gt"] > 0, axis=1)
            assert set(np.unique(d["gt"])) <= {0, 255} and set(np.unique(d["motion"])) <= {0, 255}
        path = os.path.join(OUT, f"realdata_{ds}.npz")
        np.savez_compressed(path, **out)
        print(f"realdata_{ds}.npz: {len(keep)} of {n_pairs} pairs kept, {os.path.getsize(path) / 1e6:.2f} MB; "
              f"average accuracy {acc1.mean()!r}")


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "realdata_small":
        make_realdata_small(tuple(sys.argv[2:]) or ("uav", "uavnew2", "tabletennis"))
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "roi":
        make_roi()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "hostio":
        make_hostio()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "f32":
        make_farneback_f32()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "realdata":
        make_realdata(tuple(sys.argv[2:]) or ("grasp", "autodriving"))
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "farneback":
        make_farneback()
        make_farneback_initflow()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "initflow":
        make_farneback_initflow()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "frames":
        make_frames()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "scheme2":
        make_events_scheme2()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "segment":
        make_segment()
        sys.exit(0)
    make_farneback()
    make_farneback_initflow()
    make_farneback_f32()
    make_roi()
    make_events()
    make_events_scheme2()
    make_segment()
    make_frames()
    make_hostio()
    make_realdata()
