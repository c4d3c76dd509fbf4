"""Shared test plumbing: oracle loader (ctypes over oracle/_build/libnsof_oracle.so), emulator build,
golden fixtures, synthetic inputs, comparison metrics."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")

PARAM_SETS = {
    "grasp": dict(pyr_scale=0.5, levels=3, winsize=15, iterations=3, poly_n=5, poly_sigma=1.2, flags=0),
    "autodriving_txt": dict(pyr_scale=0.6, levels=3, winsize=3, iterations=3, poly_n=10, poly_sigma=1.05, flags=0),
    "tabletennis": dict(pyr_scale=0.6, levels=3, winsize=4, iterations=2, poly_n=1, poly_sigma=1.05, flags=0),
    # OPTFLOW_FARNEBACK_GAUSSIAN = 256 (SURVEY.md 8(a) S2-e'): never selected by the scripts, covered for the drop-in
    "grasp_gaussian": dict(pyr_scale=0.5, levels=3, winsize=15, iterations=3, poly_n=5, poly_sigma=1.2, flags=256),
    "gaussian_w10": dict(pyr_scale=0.6, levels=2, winsize=10, iterations=2, poly_n=7, poly_sigma=1.5, flags=256),
}
# OPTFLOW_USE_INITIAL_FLOW (flag 4) sets of tests/golden/farneback_initflow_cv2.npz: one iteration, so that the result
# still depends on the initial flow (three iterations converge to the flags=0 answer within 1e-5 px)
INIT_PARAM_SETS = {
    "grasp_init": dict(pyr_scale=0.5, levels=3, winsize=15, iterations=1, poly_n=5, poly_sigma=1.2, flags=4),
    "gaussian_init": dict(pyr_scale=0.6, levels=2, winsize=10, iterations=2, poly_n=7, poly_sigma=1.5, flags=4 | 256),
}


def check_initial_flow(golden, farneback, what):
    """Shared by the oracle, emulator and GPU tests: `farneback(prev, next, flow0, **params)` -> flow, against cv2's
    result with the same initial flow, and (sanity of the fixture) NOT equal to the result without it."""
    g = golden.fb_init
    for case in [str(c) for c in g["cases"]]:
        prev, nxt, f0 = golden.fb[f"{case}__prev"], golden.fb[f"{case}__next"], g[f"{case}__flow0"]
        for pname, p in INIT_PARAM_SETS.items():
            want = g[f"{case}__{pname}"]
            got = farneback(prev, nxt, f0.copy(), **p)
            d = flow_err(got, want)
            assert d.mean() <= 1e-4 and d.max() <= 1e-3, f"{what} {case} {pname}: EPE mean {d.mean():.3e} max {d.max():.3e}"
            zero = farneback(prev, nxt, np.zeros_like(f0), **p)
            assert flow_err(zero, want).max() > 1e-2, f"{what} {case} {pname}: fixture does not depend on the initial flow"


# Flow tolerances vs OpenCV's float Farneback (BASELINE.json bar: EPE <= 0.01 px).  Pixels whose 2x2 system is
# near-singular (aperture problem inside a tiny window: winsize 3/4) amplify 1-ulp differences of ANY
# implementation -- the C oracle itself differs from cv2 by up to 0.6 px on 0.003 % of such pixels of a 1080p
# frame, and by 5e-3 px at the 99.9th percentile of the autodriving crop -- so for those parameter sets the
# bound is taken at the 99th percentile; the mean EPE is always checked.
TOL = {
    "grasp": dict(epe_mean=1e-4, q=1.0, qmax=1e-3),
    "autodriving_txt": dict(epe_mean=1e-3, q=0.99, qmax=1e-2),
    "tabletennis": dict(epe_mean=1e-3, q=0.99, qmax=1e-2),
    "grasp_gaussian": dict(epe_mean=1e-4, q=1.0, qmax=1e-3),
    "gaussian_w10": dict(epe_mean=1e-4, q=1.0, qmax=1e-3),
}


def build_oracle():
    path = os.path.join(ROOT, "oracle", "_build", "libnsof_oracle.so")
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle")], check=True)
    return path


def build_emu():
    if os.environ.get("NSOF_EMU_LIB"):  # e.g. the AddressSanitizer build (make -C .../csrc emu-asan)
        return os.path.abspath(os.environ["NSOF_EMU_LIB"])
    path = os.path.join(ROOT, "tests", "emu", "_build", "libnsof_emu.so")
    csrc = os.path.join(ROOT, "neuromorphic-spatiotemporal-optical-flow_b200", "csrc")
    subprocess.run(["make", "-s", "-C", csrc, "emu"], check=True)
    return path


DUMP_FN = C.CFUNCTYPE(None, C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_float), C.c_int, C.c_int, C.c_int)


class Oracle:
    def __init__(self, path):
        self.lib = C.CDLL(path)
        self.lib.nsof_oracle_farneback_dump.argtypes = [
            C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_ssize_t, C.c_ssize_t, C.c_double, C.c_int, C.c_int, C.c_int,
            C.c_int, C.c_double, C.c_int, C.c_void_p, DUMP_FN, C.c_void_p]
        self.lib.nsof_oracle_resize_area_c2.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
        self.lib.nsof_oracle_level_geometry.argtypes = [C.c_int, C.c_int, C.c_double, C.c_int] + [C.c_void_p] * 4

    def farneback(self, prev, nxt, pyr_scale, levels, winsize, iterations, poly_n, poly_sigma, flags=0, dump=False,
                  flow=None):
        prev = np.ascontiguousarray(prev)
        nxt = np.ascontiguousarray(nxt)
        h, w = prev.shape
        if prev.dtype == np.float32:   # cv2 accepts CV_32F frames; the restatement takes them through flag 0x10000
            assert nxt.dtype == np.float32
            flags |= 0x10000
        else:
            assert prev.dtype == np.uint8 and nxt.dtype == np.uint8
        out = np.zeros((h, w, 2), np.float32) if flow is None else np.ascontiguousarray(flow, np.float32).copy()
        stages = {}

        def cb(user, tag, k, it, data, hh, ww, ch):
            arr = np.ctypeslib.as_array(data, shape=(hh, ww, ch)).copy()
            stages[(tag, k, it)] = arr

        fn = DUMP_FN(cb) if dump else C.cast(None, DUMP_FN)
        rc = self.lib.nsof_oracle_farneback_dump(prev.ctypes.data, nxt.ctypes.data, h, w, prev.strides[0], nxt.strides[0], pyr_scale, levels,
                                                 winsize, iterations, poly_n, poly_sigma, flags, out.ctypes.data, fn,
                                                 None)
        assert rc == 0, rc
        return (out, stages) if dump else out

    def resize_area(self, src, hd, wd):
        src = np.ascontiguousarray(src, np.float32)
        out = np.empty((hd, wd, 2), np.float32)
        assert self.lib.nsof_oracle_resize_area_c2(src.ctypes.data, src.shape[0], src.shape[1], hd, wd, out.ctypes.data) == 0
        return out

    def level_geometry(self, h, w, pyr_scale, levels):
        hk = (C.c_int * 64)()
        wk = (C.c_int * 64)()
        ks = (C.c_int * 64)()
        sg = (C.c_double * 64)()
        n = self.lib.nsof_oracle_level_geometry(h, w, pyr_scale, levels, hk, wk, ks, sg)
        return [(hk[i], wk[i], ks[i], sg[i]) for i in range(n)]


def load_oracle():
    return Oracle(build_oracle())


class Golden:
    def __init__(self):
        self.fb = np.load(os.path.join(GOLDEN, "farneback_cv2.npz"))
        self.fb_init = np.load(os.path.join(GOLDEN, "farneback_initflow_cv2.npz"))
        self.roi = np.load(os.path.join(GOLDEN, "roi_reference.npz"))
        self.ev = np.load(os.path.join(GOLDEN, "events_reference.npz"))
        self.ev2 = np.load(os.path.join(GOLDEN, "events_scheme2_reference.npz"))
        self.seg = np.load(os.path.join(GOLDEN, "segment_reference.npz"))
        self.frames = np.load(os.path.join(GOLDEN, "frames_reference.npz"))
        self.fb_f32 = np.load(os.path.join(GOLDEN, "farneback_f32_cv2.npz"))


def segment_cases(seg):
    """(name, float32 flow region, golden mask bool, golden `mag > SEG_TH` bool) of segment_reference.npz."""
    for name in [str(c) for c in seg["cases"]]:
        stored = seg[f"{name}__flow"]
        flow = stored.astype(np.float32) / 32 if stored.dtype == np.int8 else stored
        h, w = flow.shape[:2]
        unpack = lambda a: np.unpackbits(a, axis=1)[:, :w].astype(bool)
        yield name, flow, unpack(seg[f"{name}__mask"]), unpack(seg[f"{name}__mag_gt"])


def scheme2_cases(ev2):
    """(name, kwargs for scheme2 / simulate_scheme2, golden w_a, golden w_b or None) of events_scheme2_reference.npz."""
    for name in [str(c) for c in ev2["cases"]]:
        slice_us, av, sv, n, mode = ev2[f"{name}__args"]
        n = int(n)
        pol = ev2["p_pm1"] if name == "split_pm1" else ev2["p_01"]
        kw = dict(x=ev2["x"][:n], y=ev2["y"][:n], pol=pol[:n], t=ev2["t"][:n], H=int(ev2["H"]), W=int(ev2["W"]),
                  slice_us=int(slice_us), active_v=float(av), silent_v=float(sv),
                  polarity="split" if mode == 0 else "magnitude")
        yield name, kw, ev2[f"{name}__w_a"], (ev2[f"{name}__w_b"] if mode == 0 else None)


def textured_pair(seed, h, w, dx=3.0, dy=-1.5):
    """Smooth random texture and its sub-pixel translation (numpy only: must run on the GPU box)."""
    rng = np.random.default_rng(seed)
    H, W = h + 32, w + 32
    base = rng.random((H, W)).astype(np.float64)
    k = np.exp(-0.5 * (np.arange(-9, 10) / 3.0) ** 2)
    k /= k.sum()
    base = np.apply_along_axis(lambda r: np.convolve(r, k, mode="same"), 1, base)
    base = np.apply_along_axis(lambda r: np.convolve(r, k, mode="same"), 0, base)
    base = (base - base.min()) / (base.max() - base.min()) * 255.0
    ys, xs = np.mgrid[0:h, 0:w].astype(np.float64)
    a = base[16:16 + h, 16:16 + w]
    sy, sx = ys + 16 - dy, xs + 16 - dx  # b(y, x) = a(y - dy, x - dx): content moves by (+dx, +dy)
    y0, x0 = np.floor(sy).astype(int), np.floor(sx).astype(int)
    fy, fx = sy - y0, sx - x0
    b = (base[y0, x0] * (1 - fx) * (1 - fy) + base[y0, x0 + 1] * fx * (1 - fy)
         + base[y0 + 1, x0] * (1 - fx) * fy + base[y0 + 1, x0 + 1] * fx * fy)
    return np.ascontiguousarray(a.astype(np.uint8)), np.ascontiguousarray(b.astype(np.uint8))


def flow_err(a, b):
    d = np.sqrt(((a.astype(np.float64) - b.astype(np.float64)) ** 2).sum(-1))
    return d


def assert_flow_close(got, want, pname, what=""):
    tol = TOL[pname]
    d = flow_err(got, want)
    epe = d.mean()
    qv = np.quantile(d, tol["q"]) if tol["q"] < 1.0 else d.max()
    assert epe <= tol["epe_mean"], f"{what} {pname}: mean EPE {epe:.3e} > {tol['epe_mean']}"
    assert qv <= tol["qmax"], f"{what} {pname}: q{tol['q']} EPE {qv:.3e} > {tol['qmax']}"
    assert epe <= 0.01  # BASELINE.json bar
    return epe, qv


def planar_to_aos(p):
    return np.ascontiguousarray(np.moveaxis(p, 0, -1))


def check_frames_against_golden(fr, rgb2gray, remap_flow, predict_frame):
    """Shared by the oracle, emulator and GPU tests: `fr` = frames_reference.npz; the three callables are the
    implementation under test.  Everything bit-exact."""
    for k in ("grasp_a", "grasp_b", "random"):
        assert np.array_equal(rgb2gray(fr[f"grey__{k}__in"]), fr[f"grey__{k}__out"]), k
    nf, flow = fr["pred__next"], fr["pred__flow"]
    for k in ("whole", "inner", "corner", "thin"):
        x0, y0, x1, y1 = (int(v) for v in fr[f"pred__{k}__rect"])
        want = fr[f"pred__{k}__out"]
        got = remap_flow(nf, flow[y0:y1, x0:x1], x0, y0)
        assert np.array_equal(got, want[y0:y1, x0:x1]), k
        full = predict_frame(nf, nf, flow, 2, (x0, y0, x1, y1), 2, True)
        assert np.array_equal(full, want), k
    if predict_frame is not None:
        regs = [tuple(int(v) for v in r) for r in fr["pred__merge__regions"]]
        assert np.array_equal(predict_frame(nf, nf, flow, 3, regs, 1, True, 10), fr["pred__merge__out"])
    # a single channel, the reference's float64 flow dtype
    x0, y0, x1, y1 = (int(v) for v in fr["pred__inner__rect"])
    got = remap_flow(np.ascontiguousarray(nf[:, :, 1]), flow[y0:y1, x0:x1].astype(np.float64), x0, y0)
    assert np.array_equal(got, fr["pred__inner__out"][y0:y1, x0:x1, 1])


class RealData:
    """tests/golden/realdata_<ds>.npz (make_golden.make_realdata): the scripts' main-loop body run with the reference's own
    functions over the shipped data/grasp and data/autodriving sequences."""

    def __init__(self, ds):
        self.ds = ds
        self.g = np.load(os.path.join(GOLDEN, f"realdata_{ds}.npz"))
        self.H, self.W = (int(v) for v in self.g["frame_hw"])
        self.keep = [int(i) for i in self.g["keep"]]
        self.full = [int(i) for i in self.g["full"]]

    def rect(self, i):
        return tuple(int(v) for v in self.g["rects"][i])

    def frame(self, t):
        """Grey frame t: zero outside the stored box (nothing outside the pair's rectangle reaches the ROI path)."""
        x0, y0, x1, y1 = (int(v) for v in self.g[f"fbox__{t}"])
        f = np.zeros((self.H, self.W), np.uint8)
        f[y0:y1, x0:x1] = self.g[f"frame__{t}"]
        return f

    def pair(self, i):
        return self.frame(i), self.frame(i + 1)

    def full_pair(self, i):
        return self.g[f"fullprev__{i}"], self.g[f"fullnext__{i}"]

    def unpack(self, key):
        return np.unpackbits(self.g[key], axis=1)[:, :self.W].astype(bool)

    def check_flow(self, i, crop_flow, what, mean_tol=1e-4, max_tol=1e-3):
        """crop_flow: float32 (h, w, 2) of pair i's rectangle vs cv2's flow at every 8th pixel + the |flow| checksum."""
        want = self.g[f"flow8__{i}"]
        d = flow_err(crop_flow[::8, ::8], want)
        assert d.mean() <= mean_tol and d.max() <= max_tol, f"{what} {self.ds} pair {i}: EPE mean {d.mean():.3e} max {d.max():.3e}"
        s, m = self.g[f"flowsum__{i}"]
        got = np.abs(crop_flow.astype(np.float64)).sum()
        assert abs(got - s) <= mean_tol * 2 * crop_flow.shape[0] * crop_flow.shape[1], (what, i, got, s)
        return d.mean(), d.max()


class RealDataSmall:
    """tests/golden/realdata_{uav,uavnew2,tabletennis}.npz (make_golden.make_realdata_small): the main-loop body of
    optical_flow_seg.py:413-490 with the reference's own functions and each dataset's Parameters.txt -- FLAG 1 (one
    Farneback call per connected region) for the two uav sequences, FLAG 2 with poly_n = 1 for tabletennis."""

    def __init__(self, ds):
        self.ds = ds
        self.g = np.load(os.path.join(GOLDEN, f"realdata_{ds}.npz"))
        self.H, self.W = (int(v) for v in self.g["frame_hw"])
        self.keep = [int(i) for i in self.g["keep"]]
        self.flag = int(self.g["flag"])

    def rects(self, i):
        return [tuple(int(v) for v in r) for r in self.g["rects"][i, :int(self.g["nrect"][i])]]

    def unpack(self, key):
        return np.unpackbits(self.g[key], axis=1)[:, :self.W].astype(bool)


# Per-dataset flow bounds against cv2 (px): mean EPE, share of pixels above 1e-3, maximum.  uavnew2 runs the grasp
# parameter set and meets the standard bounds.  The other two sets are ill-conditioned on these frames -- winsize 3 / 4
# leaves the 2x2 systems of flat image regions near-singular (|flow| reaches 37 px on a 161 px frame) -- and OpenCV
# does not agree with ITSELF there: cv2.setUseOptimized(False) against the default build differs by up to 1.3e-2 px on
# the uav crops and 7.6e-2 px on the tabletennis crops (measured with the committed frames, cv2 4.x of this image),
# and the C oracle differs from cv2 by 1.2e-2 / 4.6e-2.  Their maxima are bounded a small factor above that; the mean
# and the share of pixels above 1e-3 px stay tight, and what the scripts consume (motion mask, accuracy) is exact.
REALDATA_SMALL_TOL = {
    "uavnew2": dict(mean=1e-4, share=0.0, max=1e-3, live_share=1e-4, live_max=5e-3),
    "uav": dict(mean=1e-4, share=5e-3, max=0.15, live_share=5e-3, live_max=0.15),
    "tabletennis": dict(mean=2e-4, share=3e-2, max=0.3, live_share=3e-2, live_max=0.3),
}


def check_realdata_small(engine, nsof, ds, padding, live_cv2=True, pairs=None):
    """Engine.opticalFlow3D -> `flow = -flow` -> task_results -> calculate_pixel_accuracy on the kept pairs of one of
    the three small sequences.  Region rectangles and num_labels equal the reference's; the pasted flow within
    REALDATA_SMALL_TOL[ds] of cv2's at every 4th pixel, and (when cv2 imports) of cv2 run live on every
    region crop at EVERY pixel (REALDATA_SMALL_TOL); motion mask bit for bit; accuracy to the last digit.
    -> (worst mean EPE, worst max EPE, pairs checked)."""
    cv2 = None
    if live_cv2:
        try:
            import cv2
        except Exception:
            cv2 = None
    rd = RealDataSmall(ds)
    cfg = getattr(nsof.RoiConfig, ds)()
    assert cfg.FLAG == rd.flag
    M, tol = cfg.MEMSIZE, REALDATA_SMALL_TOL[ds]
    worst = [0.0, 0.0]
    todo = rd.keep if pairs is None else rd.keep[:pairs]
    for i in todo:
        a, b = rd.g[f"gray__{i}"], rd.g[f"gray__{i + 1}"]
        memimg = rd.g[f"memimg__{i}"]
        res = engine.opticalFlow3D(memimg, memimg, a, b, M, M, cfg)
        flow = res[0]
        if rd.flag == 1:
            num_labels, info = res[4], [tuple(int(v) for v in r) for r in res[5]]
            assert num_labels == int(rd.g["num_labels"][i]) and info == rd.rects(i), (ds, i, info, rd.rects(i))
            regions = info
        else:
            info = tuple(int(v) for v in res[4])
            assert [info] == rd.rects(i), (ds, i, info)
            num_labels, regions = 2, [info]
        d = flow_err(flow[::4, ::4], rd.g[f"flow4__{i}"])
        assert d.mean() <= tol["mean"] and (d > 1e-3).mean() <= tol["share"] and d.max() <= tol["max"], \
            f"{ds} pair {i}: EPE mean {d.mean():.3e} share>1e-3 {(d > 1e-3).mean():.3e} max {d.max():.3e}"
        worst = [max(worst[0], d.mean()), max(worst[1], d.max())]
        ssum, _ = rd.g[f"flowsum__{i}"]
        assert abs(np.abs(flow).sum() - ssum) <= tol["mean"] * 2 * rd.H * rd.W, (ds, i)
        if cv2 is not None:
            want = np.zeros((rd.H, rd.W, 2))
            for (x0, y0, x1, y1) in regions:      # optical_flow_seg.py:155-159: later regions overwrite earlier ones
                want[y0:y1, x0:x1] = cv2.calcOpticalFlowFarneback(a[y0:y1, x0:x1], b[y0:y1, x0:x1], None, **cfg.farneback_params)
            d = flow_err(flow, want)
            assert d.mean() <= tol["mean"] and (d > 1e-3).mean() <= tol["live_share"] and d.max() <= tol["live_max"], \
                (ds, i, d.mean(), (d > 1e-3).mean(), d.max())
            worst[1] = max(worst[1], d.max())
        motion = engine.task_results(a, b, -flow, num_labels, info, EST_FLAG=rd.flag, MERGE_FLAG=True, padding=padding)
        assert np.array_equal(motion > 0, rd.unpack(f"motion__{i}")), (ds, i, int(((motion > 0) != rd.unpack(f"motion__{i}")).sum()))
        assert pixel_accuracy(motion > 0, rd.unpack(f"gt__{i}")) == rd.g["acc1"][i], (ds, i)
    return worst[0], worst[1], len(todo)


def pixel_accuracy(a, b):
    """calculate_pixel_accuracy (optical_flow_seg.py:384-388)."""
    return np.sum(a == b) / a.size * 100


def external_image(H, W, seed=20261105):
    """Seeded uint8 test image of tests/golden/make_stage1_external.py (noise with a vertical brightness ramp, so that
    the cells of a Lanczos reduction differ by far more than the comparison tolerance)."""
    rng = np.random.default_rng(seed + H * 10000 + W)
    img = rng.integers(0, 256, (H, W), dtype=np.uint8)
    return (0.5 * img + 0.5 * np.roll(img, 37, 1) * (np.arange(H)[:, None] / H)).astype(np.uint8)


class Stage1External:
    """tests/golden/stage1_external.npz (make_stage1_external.py): frame-mode fixtures from an independent 50-digit
    derivation -- Lanczos-3 contribution matrices, Pillow's resize as a loose cross-check, exact-arithmetic Euler
    trajectories of update_state."""

    def __init__(self):
        self.g = np.load(os.path.join(GOLDEN, "stage1_external.npz"))

    def check_compress(self, compress, what, tol=1e-13):
        """compress(img uint8 (H, W), m, n) -> cells; vs  W_y (img / 255) W_x^T  with the external matrices."""
        for H, W, M in [(1080, 1920, 80), (801, 801, 200)]:
            img = external_image(H, W)
            assert int(img.astype(np.int64).sum()) == int(self.g[f"pillow__{H}x{W}__imgsum"])
            want = self.g[f"lanczos__{H}_{H // M}"] @ (img / 255.0) @ self.g[f"lanczos__{W}_{W // M}"].T
            got = compress(img, M, M)
            assert got.shape == want.shape and np.abs(got - want).max() <= tol, (what, H, W, np.abs(got - want).max())
            if H == 1080:  # cells whose 6-cell support stays inside the image: Pillow's edge rule does not matter
                pil = self.g[f"pillow__{H}x{W}__cells"]
                assert np.abs(got - pil)[3:-3, 3:-3].max() <= 1e-6, what
        for n_in, n_out, m in [(161, 4, 40), (160, 4, 40), (600, 15, 40)]:   # uav / tabletennis / uavnew2 geometry
            img = external_image(n_in, n_in)
            Wm = self.g[f"lanczos__{n_in}_{n_out}"]
            got = compress(img, m, m)
            assert np.abs(got - Wm @ (img / 255.0) @ Wm.T).max() <= tol, (what, n_in)

    def euler_groups(self):
        """-> [(th1, th2, w0 (n,), delta (n,), want (n, 3))] -- one simulator call per threshold pair."""
        pts, W = self.g["euler__points"], self.g["euler__w"]
        out = []
        for th in sorted({(a, b) for _, _, a, b in pts}):
            sel = [i for i, q in enumerate(pts) if (q[2], q[3]) == th]
            out.append((th[0], th[1], pts[sel, 0], pts[sel, 1], W[sel]))
        return out

    def check_euler(self, simulate, what, tol):
        """simulate(cells (4, 1, n), w0 (1, n), th1, th2) -> currents or (w, currents); the state after each of the 3
        frames is recovered from the current (vds / R(w)) and compared with the 50-digit Euler recurrence."""
        lam = np.log(2104377 / 163305)
        worst = 0.0
        for th1, th2, w0, d, want in self.euler_groups():
            cells = np.zeros((4, 1, len(d)))
            cells[1::2, 0, :] = d / 256.0           # |256 a - 256 b| = d for every consecutive pair
            w_fin, cur = simulate(cells, w0[None, :], th1, th2)
            w_of = 1.0 + np.log(cur[1:, 0, :] * 163305.0) / lam       # invert I = exp(-lam (1 - w)) / Ron
            err = max(np.abs(w_fin[0] - want[:, 2]).max(), np.abs(w_of.T - want).max())
            worst = max(worst, err)
            assert np.array_equal(w_fin[0] == 0, want[:, 2] == 0) and np.array_equal(w_fin[0] == 1, want[:, 2] == 1), what
        assert worst <= tol, (what, worst)
        return worst


def check_ssim(ssim, frames, what, tol=1e-12):
    """Shared by the oracle, emulator and GPU tests: `ssim(a, b)` for 2-D uint8 arrays (channel views allowed) against
    known answers and the scipy-based restatement (oracle/frames_oracle.py; skimage itself is not installed)."""
    from oracle import frames_oracle as fo
    nf, pred = frames["pred__next"], frames["pred__whole__out"]
    assert ssim(nf[:, :, 2], nf[:, :, 2]) == 1.0, what                                 # identical images
    for c1, c2 in ((10, 200), (0, 255), (77, 77)):                                     # constant images: closed form
        a, b = np.full((40, 50), c1, np.uint8), np.full((40, 50), c2, np.uint8)
        C1 = (0.01 * 255.0) ** 2
        assert abs(ssim(a, b) - (2.0 * c1 * c2 + C1) / (c1 * c1 + c2 * c2 + C1)) <= tol, (what, c1, c2)
    rng = np.random.default_rng(0)
    cases = [(nf[:, :, 2], pred[:, :, 2]), (nf[:, :, 0], pred[:, :, 1]),               # strided channel views of real frames
             (rng.integers(0, 256, (37, 45), dtype=np.uint8), rng.integers(0, 256, (37, 45), dtype=np.uint8)),
             (rng.integers(0, 256, (7, 7), dtype=np.uint8), rng.integers(0, 256, (7, 7), dtype=np.uint8)),
             (rng.integers(0, 256, (100, 7), dtype=np.uint8), rng.integers(0, 256, (100, 7), dtype=np.uint8))]
    for a, b in cases:
        assert abs(ssim(a, b) - fo.ssim_u8(a, b)) <= tol, (what, a.shape)


def check_shared_frames(eng, device_batch=None):
    """A batch of consecutive pairs of one clip (frames f0..f3, the same crop) shares the expansion of the middle frames
    between neighbouring items; the flows must equal, BIT FOR BIT, those of single calls and of the batch with the
    sharing switched off.  Different crops of the same frames must not be shared."""
    base, _ = textured_pair(51, 260, 330)
    frames = [np.ascontiguousarray(np.roll(base, (2 * t, -3 * t), axis=(0, 1))) for t in range(4)]
    p = PARAM_SETS["grasp"]
    crop = (slice(20, 240), slice(16, 316))
    other = (slice(20, 240), slice(24, 316))   # another rectangle: its frames start at other addresses
    pairs = [(frames[t][crop], frames[t + 1][crop]) for t in range(3)] + [(frames[1][other], frames[2][other])]
    singles = [eng.calcOpticalFlowFarneback(a, b, None, **p) for a, b in pairs]
    if device_batch is None:   # host batch: the crops are copied to separate device buffers, nothing can be shared
        return singles
    got = device_batch(frames, [crop, crop, crop, other], [0, 1, 2, 1], True)
    off = device_batch(frames, [crop, crop, crop, other], [0, 1, 2, 1], False)
    for i, want in enumerate(singles):
        assert np.array_equal(got[i], want), i
        assert np.array_equal(off[i], want), i
    return singles


def check_float32_frames(golden, farneback, what):
    """cv2.calcOpticalFlowFarneback on FLOAT32 frames (non-integer grey values; tests/golden/farneback_f32_cv2.npz):
    `farneback(prev, next, **params)` within the same tolerances as the uint8 cases."""
    g = golden.fb_f32
    for case in [str(c) for c in g["cases"]]:
        prev, nxt = g[f"{case}__prev"], g[f"{case}__next"]
        assert prev.dtype == np.float32
        for pname in ("grasp", "autodriving_txt"):
            assert_flow_close(farneback(prev, nxt, **PARAM_SETS[pname]), g[f"{case}__{pname}"], pname, f"{what} f32 {case}")


def warped_pair(seed, h, w, kind):
    """Texture and a warped copy: kind "shift20" = translation by (+20.3, -12.6) px (large but uniform: the R1 window
    follows it), "zoom" = 6 % zoom about the centre (|flow| grows to ~12 px, varies across every tile), "shear" =
    dy proportional to x (vertical variation of > 3 px inside most tiles: the L2 path)."""
    if kind == "shift20":
        return _warp(seed, h, w, lambda y, x: (y + 12.6, x - 20.3))
    if kind == "zoom":
        return _warp(seed, h, w, lambda y, x: (h / 2 + (y - h / 2) / 1.06, w / 2 + (x - w / 2) / 1.06))
    if kind == "shear":
        return _warp(seed, h, w, lambda y, x: (y - 0.08 * (x - w / 2), x))
    raise ValueError(kind)


def _warp(seed, h, w, src_of):
    """b(y, x) = base(src_of(y, x)) by bilinear sampling of a larger seeded texture (numpy only)."""
    rng = np.random.default_rng(seed)
    pad = 48
    H, W = h + 2 * pad, w + 2 * pad
    base = rng.random((H, W)).astype(np.float64)
    k = np.exp(-0.5 * (np.arange(-9, 10) / 3.0) ** 2)
    k /= k.sum()
    base = np.apply_along_axis(lambda r: np.convolve(r, k, mode="same"), 1, base)
    base = np.apply_along_axis(lambda r: np.convolve(r, k, mode="same"), 0, base)
    base = (base - base.min()) / (base.max() - base.min()) * 255.0
    ys, xs = np.mgrid[0:h, 0:w].astype(np.float64)
    a = base[pad:pad + h, pad:pad + w]
    sy, sx = src_of(ys, xs)
    sy = np.clip(sy + pad, 0, H - 2)
    sx = np.clip(sx + pad, 0, W - 2)
    y0, x0 = np.floor(sy).astype(int), np.floor(sx).astype(int)
    fy, fx = sy - y0, sx - x0
    b = (base[y0, x0] * (1 - fx) * (1 - fy) + base[y0, x0 + 1] * fx * (1 - fy)
         + base[y0 + 1, x0] * (1 - fx) * fy + base[y0 + 1, x0 + 1] * fx * fy)
    return np.ascontiguousarray(a.astype(np.uint8)), np.ascontiguousarray(b.astype(np.uint8))


def r1_window_stats(eng):
    import ctypes as C
    out = (C.c_ulonglong * 4)()
    eng.lib.check(eng.lib.nsof_debug_r1_window_stats(eng.handle, out, eng._stream()), "r1_window_stats")
    return int(out[0]), int(out[1]), int(out[2]), int(out[3])   # blur+solve (window, L2), first UpdateMatrices (window, L2)


def check_r1_window(eng, hw=(300, 400), pname="grasp"):
    """The tuned blur+solve kernel gathers R1 from a shared-memory window that follows the tile's flow, or from L2 when
    the flow varies too much inside a tile: both paths load the same values, so the flow must be BIT-IDENTICAL with the
    window switched off -- for a uniform small shift, a uniform 20 px shift (window far from the tile), a zoom and a
    shear (tiles of both kinds).  The first UpdateMatrices of a level takes its own window kernel (its flow is the upsampled
    coarser flow) or, with the switch off, the round-1 kernel.  -> {(kind, big tiles): (blur+solve window tiles, L2 tiles,
    first-UpdateMatrices window tiles, L2 tiles)} of the window-on runs."""
    p = PARAM_SETS[pname]   # grasp: the M = 7 kernel; autodriving_txt / tabletennis: its small-radius instantiations
    r1_window_stats(eng)   # first call: counting on
    stats = {}
    try:
        for big in (0, 1):   # the 32 x 16 tile of latency-bound levels (default for one small pair) and the 64 x 32 tile
            eng.lib.check(eng.lib.nsof_debug_big_tiles(eng.handle, big))
            for kind in ("shift3", "shift20", "zoom", "shear"):
                a, b = textured_pair(61, *hw) if kind == "shift3" else warped_pair(61, *hw, kind)
                eng.lib.check(eng.lib.nsof_debug_r1_window(eng.handle, 1))
                on = eng.calcOpticalFlowFarneback(a, b, None, **p)
                stats[kind, big] = r1_window_stats(eng)
                eng.lib.check(eng.lib.nsof_debug_r1_window(eng.handle, 0))
                off = eng.calcOpticalFlowFarneback(a, b, None, **p)
                n_on, n_off, u_on, u_off = r1_window_stats(eng)   # window off: the round-1 um_first kernel counts nothing
                assert n_on == 0 and n_off == sum(stats[kind, big][:2]) and u_on == 0 and u_off == 0, (kind, big, n_on, n_off)
                assert np.array_equal(on, off), (kind, big, np.abs(on - off).max())
                assert np.isfinite(on).all()
            assert stats["shift3", big][0] > 0 and stats["shift20", big][0] > 0, stats
            assert stats["shear", big][1] > 0, stats
            assert stats["shift3", big][2] > 0 and stats["shear", big][2] > 0, stats   # first UpdateMatrices: window taken
    finally:
        eng.lib.check(eng.lib.nsof_debug_r1_window(eng.handle, 1))
        eng.lib.check(eng.lib.nsof_debug_big_tiles(eng.handle, 0))
    return stats


def check_small_radius_blur(eng, golden, hw=(300, 420)):
    """winsize 3 and 4 (the two other parameter sets the reference ships) run the tuned blur+solve kernel's small-radius
    instantiations by default; `nsof_debug_generic_blur(1)` puts them back on the generic kernel (double running sums).
    Both paths must meet the cv2 tolerances on the golden cases, agree with each other to 2e-3 px (the sums differ in the
    last bits; the near-singular systems of these sets amplify that), and the tile counters must show which one ran."""
    a, b = textured_pair(7, *hw)
    r1_window_stats(eng)
    out = {}
    try:
        for pname in ("autodriving_txt", "tabletennis"):
            p = PARAM_SETS[pname]
            flows = {}
            for generic in (0, 1, 2):   # 0: tuned kernel, small tiles where the level is small; 2: its 64 x 32 tiles everywhere
                eng.lib.check(eng.lib.nsof_debug_generic_blur(eng.handle, generic == 1))
                eng.lib.check(eng.lib.nsof_debug_big_tiles(eng.handle, generic == 2))
                flows[generic] = eng.calcOpticalFlowFarneback(a, b, None, **p)
                st = r1_window_stats(eng)
                assert (st[0] + st[1] > 0) == (generic != 1), (pname, generic, st)
                for case in ("synth_96x128", "synth_75x101"):
                    got = eng.calcOpticalFlowFarneback(golden.fb[f"{case}__prev"], golden.fb[f"{case}__next"], None, **p)
                    assert_flow_close(got, golden.fb[f"{case}__{pname}"], pname, f"generic={generic} {case}")
                r1_window_stats(eng)
            d = flow_err(flows[0], flows[1])
            assert d.mean() <= 2e-5 and d.max() <= 2e-3, (pname, d.mean(), d.max())
            assert np.array_equal(flows[0], flows[2]), pname   # the tile shape does not change a bit
            out[pname] = (d.mean(), d.max())
    finally:
        eng.lib.check(eng.lib.nsof_debug_generic_blur(eng.handle, 0))
        eng.lib.check(eng.lib.nsof_debug_big_tiles(eng.handle, 0))
    return out


def check_fused_pyramid_all_n(eng, hw=(230, 310), exact=False):
    """Level 0 of every shipped parameter set (poly_n 5, 10, 1) runs the fused pyramid + PolyExp kernel; its operations are
    those of the unfused pair of kernels in the same order: flows bit-identical with `nsof_debug_fuse(0)` on the emulator
    build (exact=True); on the device the compiler contracts the two kernels' multiply-adds differently (6e-6 px on the
    grasp set), so there the two must agree to mean 2e-5 / max 2e-3 px.  The profile must name the kernel that ran."""
    a, b = textured_pair(9, *hw)
    try:
        for pname in ("grasp", "autodriving_txt", "tabletennis"):
            p = PARAM_SETS[pname]
            eng.lib.check(eng.lib.nsof_debug_fuse(eng.handle, 1))
            eng.profile(True)
            f1 = eng.calcOpticalFlowFarneback(a, b, None, **p)
            eng.profile(False)
            assert "fb_pyr_polyexp.L0" in eng.profile_read(), pname
            eng.lib.check(eng.lib.nsof_debug_fuse(eng.handle, 0))
            eng.profile(True)
            f0 = eng.calcOpticalFlowFarneback(a, b, None, **p)
            eng.profile(False)
            assert "fb_pyr_polyexp.L0" not in eng.profile_read(), pname
            d = flow_err(f1, f0)
            assert d.mean() <= 2e-5 and d.max() <= 2e-3, (pname, d.mean(), d.max())
            if exact:
                assert np.array_equal(f1, f0), (pname, np.abs(f1 - f0).max())
    finally:
        eng.lib.check(eng.lib.nsof_debug_fuse(eng.handle, 1))
