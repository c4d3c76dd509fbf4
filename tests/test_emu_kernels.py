"""Kernel-logic checks WITHOUT a GPU: the same .cu sources compiled by g++ -DNSOF_EMU into a sequential
block emulator (tests/emu, nsof_rt.h) are driven through the same C ABI and compared with the oracle and the
golden vectors.  This exercises tiling, halos, borders, batching and the host driver; it says nothing about
races or performance -- the `-m gpu` tests are the parity tests proper."""
import numpy as np
import pytest

import helpers
from oracle import stage1_oracle as s1


@pytest.mark.parametrize("pname", list(helpers.PARAM_SETS))
def test_emu_farneback_vs_golden(emu_engine, golden, pname):
    p = helpers.PARAM_SETS[pname]
    for case in golden.fb["cases"]:
        prev, nxt = golden.fb[f"{case}__prev"], golden.fb[f"{case}__next"]
        got = emu_engine.calcOpticalFlowFarneback(prev, nxt, None, **p)
        helpers.assert_flow_close(got, golden.fb[f"{case}__{pname}"], pname, f"emu vs cv2 {case}")
    if pname == "grasp":  # both tiles of the tuned blur+solve kernel: small inputs take the 32x16 one by default
        emu_engine.lib.check(emu_engine.lib.nsof_debug_big_tiles(emu_engine.handle, 1))
        try:
            for case in golden.fb["cases"]:
                got = emu_engine.calcOpticalFlowFarneback(golden.fb[f"{case}__prev"], golden.fb[f"{case}__next"], None, **p)
                helpers.assert_flow_close(got, golden.fb[f"{case}__{pname}"], pname, f"emu (64x32 tile) vs cv2 {case}")
        finally:
            emu_engine.lib.check(emu_engine.lib.nsof_debug_big_tiles(emu_engine.handle, 0))


def test_emu_farneback_initial_flow_vs_golden(emu_engine, golden, nsof):
    """OPTFLOW_USE_INITIAL_FLOW through the C ABI (flow is in/out): cv2 goldens; flow=None with the flag raises."""
    helpers.check_initial_flow(golden, lambda a, b, f0, **p: emu_engine.calcOpticalFlowFarneback(a, b, f0, **p), "emu vs cv2")
    a = golden.fb["synth_40x33__prev"]
    with pytest.raises(nsof.error):
        emu_engine.calcOpticalFlowFarneback(a, a, None, **helpers.INIT_PARAM_SETS["grasp_init"])


@pytest.mark.parametrize("hw,fuse", [((150, 203), False), ((152, 204), True), ((150, 203), True)])
def test_emu_stage_dumps_vs_oracle(emu_engine, oracle, hw, fuse):
    """Every intermediate (I, R0, R1, M, per-iteration flow) of every level against the oracle's dump, through the
    generic pyramid + PolyExp kernels (fuse off: I is materialised) and through the fused ones (152 x 204 has an
    exact-2x level 1)."""
    prev, nxt = helpers.textured_pair(5, *hw)
    p = helpers.PARAM_SETS["grasp"]
    _, st = oracle.farneback(prev, nxt, **p, dump=True)
    geo = oracle.level_geometry(hw[0], hw[1], p["pyr_scale"], p["levels"])
    assert len(geo) == 3
    try:
        emu_engine.debug_fuse(fuse)
        for k, (hk, wk, _, _) in enumerate(geo):
            for it in (0, 1, p["iterations"]):
                emu_engine.debug_truncate(k, it)
                emu_engine.calcOpticalFlowFarneback(prev, nxt, None, **p)
                if it == 0:
                    for what, tag in ((0, 0), (1, 1), (2, 2), (3, 3), (4, 4)):
                        if what <= 1 and fuse:
                            continue
                        got = helpers.planar_to_aos(emu_engine.debug_fetch(0, what, hk, wk))
                        want = st[(tag, k, 0)]
                        scale = max(1.0, float(np.abs(want).max()))
                        assert np.abs(got - want).max() <= 2e-5 * scale, (k, what)
                else:
                    got = emu_engine.debug_fetch(0, 5, hk, wk)
                    want = st[(5, k, it - 1)]
                    assert np.abs(got - want).max() <= 1e-4, (k, it)
    finally:
        emu_engine.debug_truncate(-1, -1)
        emu_engine.debug_fuse(True)


def test_emu_batch_ragged_and_strided(emu_engine, oracle):
    p = helpers.PARAM_SETS["grasp"]
    big_a, big_b = helpers.textured_pair(11, 240, 320)
    views = [(slice(0, 240), slice(0, 320)), (slice(13, 140), slice(7, 96)), (slice(100, 133), slice(20, 300)),
             (slice(200, 208), slice(300, 308))]
    pairs = [(big_a[v], big_b[v]) for v in views]  # non-contiguous views, like the reference's crops
    flows = emu_engine.farneback_batch(pairs, **p)
    for (a, b), f in zip(pairs, flows):
        want = oracle.farneback(np.ascontiguousarray(a), np.ascontiguousarray(b), **p)
        helpers.assert_flow_close(f, want, "grasp", "emu batch")
        single = emu_engine.calcOpticalFlowFarneback(a, b, None, **p)
        assert np.array_equal(single, f)  # batching never changes a crop's result


def test_emu_errors_mirror_cv2(emu_engine, nsof):
    a = np.zeros((40, 40), np.uint8)
    p = dict(helpers.PARAM_SETS["grasp"])
    with pytest.raises(nsof.error):
        emu_engine.calcOpticalFlowFarneback(a[:0], a[:0], None, **p)
    with pytest.raises(nsof.error):
        emu_engine.calcOpticalFlowFarneback(a, a[:, :20], None, **p)
    with pytest.raises(nsof.error):
        emu_engine.calcOpticalFlowFarneback(a, a, None, **dict(p, pyr_scale=1.0))
    with pytest.raises(NotImplementedError):
        emu_engine.calcOpticalFlowFarneback(a, a, None, **dict(p, flags=8))  # not a cv2 flag
    with pytest.raises(NotImplementedError):
        emu_engine.calcOpticalFlowFarneback(a, a, None, **dict(p, winsize=1))
    z = emu_engine.calcOpticalFlowFarneback(a, a, None, **p)
    assert z.shape == (40, 40, 2) and z.dtype == np.float32 and not z.any()


@pytest.mark.parametrize("ds,cfgname,hw", [("grasp", "grasp", (1920, 1080)), ("autodriving", "autodriving", (801, 801)),
                                           ("uav", "uav", (161, 161)), ("uavnew2", "uavnew2", (600, 600)),
                                           ("tabletennis", "tabletennis", (160, 160))])
def test_emu_roi_bit_exact_vs_reference(emu_engine, nsof, golden, ds, cfgname, hw):
    cfg = getattr(nsof.RoiConfig, cfgname)()
    cur = golden.roi[f"{ds}__currents"]
    for flag in (1, 2):
        cfg.FLAG = flag
        r = emu_engine.roi_from_currents(cur, hw, cfg=cfg)
        assert np.array_equal(r["memimg"], golden.roi[f"{ds}__memimg"])
        assert np.array_equal(r["labels"], golden.roi[f"{ds}__labels"])
        for t in range(cur.shape[0]):
            n = int(golden.roi[f"{ds}__flag{flag}__nrect"][t])
            assert r["rects"][t] == [tuple(v) for v in golden.roi[f"{ds}__flag{flag}__rects"][t, :n]]
            assert np.array_equal(r["stats"][t], golden.roi[f"{ds}__stats"][t, :len(r["stats"][t])])
        if flag == 1:
            assert np.array_equal(r["num_labels"], golden.roi[f"{ds}__num_labels"])


def test_emu_cc_random_masks_vs_cv2(emu_engine, nsof, golden):
    masks = golden.roi["random__masks"]
    cfg = nsof.RoiConfig(THRES=255, MEMSIZE=1, EXTEND_WIDTH_LEFT=0, EXTEND_WIDTH_RIGHT=0, EXTEND_HEIGHT_UPPER=0,
                         EXTEND_HEIGHT_LOWER=0, FLAG=1)
    r = emu_engine.roi_from_memimg(masks, masks.shape[1:], 1, 1, cfg)
    assert np.array_equal(r["labels"], golden.roi["random__labels"])
    assert np.array_equal(r["num_labels"], golden.roi["random__num_labels"])
    for i in range(len(masks)):
        n = int(golden.roi["random__num_labels"][i]) - 1
        assert np.array_equal(r["stats"][i], golden.roi["random__stats"][i, :n])
        # FLAG 1 rectangles with MEMSIZE 1 and no extension are the stats boxes
        assert r["rects"][i] == [(int(s[0]), int(s[1]), int(s[0] + s[2]), int(s[1] + s[3])) for s in r["stats"][i]]


def test_emu_opticalflow3d_matches_reference_structure(emu_engine, nsof, golden, oracle):
    cfg = nsof.RoiConfig.autodriving()
    emu_engine.config = cfg
    memimg = golden.roi["autodriving__memimg"]
    t = int(np.argmax(golden.roi["autodriving__flag2__nrect"]))  # a slice with a non-empty ROI
    a, b = helpers.textured_pair(3, 801, 801)
    flow, cal, vel, region_list, info = emu_engine.opticalFlow3D(memimg[t], memimg[t], a, b, cfg.MEMSIZE, cfg.MEMSIZE)
    x0, y0, x1, y1 = info
    assert info == tuple(golden.roi["autodriving__flag2__rects"][t, 0])
    assert flow.dtype == np.float64 and flow.shape == (801, 801, 2)
    want = oracle.farneback(np.ascontiguousarray(a[y0:y1, x0:x1]), np.ascontiguousarray(b[y0:y1, x0:x1]), **cfg.farneback_params)
    helpers.assert_flow_close(flow[y0:y1, x0:x1].astype(np.float32), want, "grasp", "opticalFlow3D crop")
    outside = np.ones((801, 801), bool)
    outside[y0:y1, x0:x1] = False
    assert not flow[outside].any()
    # empty ROI -> zeros and (0,0,0,0), optical_flow_seg.py:227-234
    t0 = int(np.argmin(golden.roi["autodriving__flag2__nrect"]))
    res = emu_engine.opticalFlow3D(memimg[t0], memimg[t0], a, b, cfg.MEMSIZE, cfg.MEMSIZE)
    assert not res[0].any() and res[4] == (0, 0, 0, 0)
    emu_engine.config = nsof.RoiConfig()


def test_emu_events_vs_reference(emu_engine, golden):
    ev = golden.ev
    x, y, t = ev["x"], ev["y"], ev["t"]
    H, W = int(ev["H"]), int(ev["W"])
    b = ev["slice_bounds"]
    for k, i in enumerate(ev["hist_slices"]):
        sl = slice(b[i], b[i + 1])
        assert np.array_equal(emu_engine.bincount_2d(x[sl], y[sl], H, W), ev["hist"][k])  # integer: bit-exact
    got = emu_engine.update_state(ev["us_w0"], ev["us_V0"])
    assert np.abs(got - ev["us_w1"]).max() <= 4 * np.finfo(np.float32).eps
    w = emu_engine.simulate_scheme1(x, y, t, H, W, 1000)
    assert np.abs(w - ev["w_snaps"][-1]).max() <= 2e-6
    assert np.array_equal(w != np.float32(0.5), ev["w_snaps"][-1] != np.float32(0.5))  # same set of touched pixels
    w64 = emu_engine.simulate_scheme1(x[:b[64]], y[:b[64]], t[:b[64]], H, W, 1000, silent_v=0.5)
    # t[:b[64]] ends inside slice 63, so the stream has 64 windows
    assert np.abs(w64 - ev["w_decay64"]).max() <= 2e-5
    R = emu_engine.resistance_exp(ev["w_snaps"][-1])
    assert R.dtype == np.float64 and np.allclose(R, ev["R_final"], rtol=1e-12)


def test_emu_frame_mode_vs_oracle(emu_engine):
    rng = np.random.default_rng(0)
    frames = rng.integers(0, 256, (3, 161, 203), dtype=np.uint8)
    cells = emu_engine.compress_images(frames, 40, 40)
    for i in range(3):
        assert np.allclose(cells[i], s1.compress_image(frames[i], 40, 40), rtol=0, atol=1e-13)
    # width a multiple of 16: the row pass takes its bulk-copied shared-memory window (two column tiles: 512 + 48)
    frames = rng.integers(0, 256, (2, 170, 560), dtype=np.uint8)
    cells = emu_engine.compress_images(frames, 40, 40)
    for i in range(2):
        assert np.allclose(cells[i], s1.compress_image(frames[i], 40, 40), rtol=0, atol=1e-13)
    seq = np.clip(0.5 + 0.02 * rng.standard_normal((4, 3, 5)), 0, 1)
    w, cur = emu_engine.simulate_memristor_array(seq, nSubSteps=20)
    w_o, cur_o = s1.simulate_memristor_array(seq, nSubSteps=20)
    assert np.allclose(w, w_o, rtol=0, atol=1e-12) and np.allclose(cur, cur_o, rtol=1e-11)
    w2, cur2 = emu_engine.simulate_memristor_array_stepwise(seq, nSubSteps=20)
    assert np.allclose(w2, w_o, rtol=0, atol=1e-12) and np.allclose(cur2, cur_o, rtol=1e-11)
    # the carried-power chain at the reference's nSubSteps, through decay, potentiation and both clips of w
    T = 24
    seq = np.zeros((T, 2, 3))
    seq[::2, 0, :] = 1.0
    seq[:, 1, :] = np.linspace(0, 1, T)[:, None] * np.array([0.0, 0.6, 1.0])
    w, cur = emu_engine.simulate_memristor_array(seq, nSubSteps=1000)
    w_o, cur_o = s1.simulate_memristor_array(seq, nSubSteps=1000)
    assert (w_o == 1.0).any() and (w_o == 0.0).any()
    assert np.array_equal(w == 1.0, w_o == 1.0) and np.array_equal(w == 0.0, w_o == 0.0)
    assert np.allclose(w, w_o, rtol=0, atol=1e-12) and np.allclose(cur, cur_o, rtol=1e-11)


def test_emu_event_roi_glue_vs_oracle(emu_engine, nsof, golden):
    """Config 3 glue: event state -> cell currents (block max) -> uint8 map -> mask -> rectangles, bit-exact."""
    ev = golden.ev
    H, W = int(ev["H"]), int(ev["W"])
    w = ev["w_decay64"]
    cur = emu_engine.event_cell_currents(w, 20)
    want = s1.event_cell_currents(w, 20)
    assert cur.shape == (12, 16) and np.array_equal(cur, want)
    cfg = nsof.RoiConfig(MEMSIZE=20, THRES=250, EXTEND_WIDTH_LEFT=8, EXTEND_WIDTH_RIGHT=8, EXTEND_HEIGHT_UPPER=8,
                         EXTEND_HEIGHT_LOWER=8, FLAG=2)
    r = emu_engine.roi_from_currents(cur, (H, W), cfg=cfg)
    n, labels, stats, rects = s1.roi_from_memimg(s1.current_to_u8(want), (H, W), 20, 20, 250, (8, 8, 8, 8), 2)
    assert r["rects"] == rects and len(rects) == 1 and np.array_equal(r["labels"], labels)


def test_emu_event_scheme2_vs_reference(emu_engine, golden):
    """Scheme 2 (DC bias + event overlay, per-pixel refractory period; eventsim/event_mem_sim.py:231-286) against
    w_final of the reference's own simulate(): the set of pixels at the clip values and every hit / refractory decision
    must agree (a wrong decision moves w by >= 1e-3); float32 powf differs from numpy's by ulps per step."""
    for name, kw, w_a, w_b in helpers.scheme2_cases(golden.ev2):
        kw = dict(kw)
        got_a, got_b = emu_engine.simulate_scheme2(kw.pop("x"), kw.pop("y"), kw.pop("pol"), kw.pop("t"), kw.pop("H"), kw.pop("W"),
                                               kw.pop("slice_us"), **kw)
        for got, want in ((got_a, w_a), (got_b, w_b)):
            assert (got is None) == (want is None), name
            if want is None:
                continue
            assert got.dtype == np.float32 and got.shape == want.shape
            assert np.array_equal(got == 0, want == 0) and np.array_equal(got == 1, want == 1), name
            assert np.abs(got - want).max() <= 2e-5, (name, np.abs(got - want).max())


def test_emu_segment_vs_reference_and_oracle(emu_engine, golden):
    """Bit-plane segmentation kernels (emulator build): masks of the reference's process_flow_region bit for bit;
    threshold-only mode; other elements / iteration counts vs the (pinned) numpy restatement; the cv2 elements."""
    from oracle import seg_oracle as so
    for name, flow, mask, mag_gt in helpers.segment_cases(golden.seg):
        if flow.shape[0] * flow.shape[1] > 40000:
            continue  # the emulator is slow; the GPU test runs the ROI-sized case
        assert np.array_equal(emu_engine.segment_flow(flow, iterations=0) > 0, mag_gt), name
        got = emu_engine.segment_flow(flow)
        assert got.dtype == np.uint8 and set(np.unique(got)) <= {0, 255}
        assert np.array_equal(got > 0, mask), name
        assert np.array_equal(emu_engine.segment_flow(-flow.astype(np.float64)) > 0, mask), name
        emu_engine.debug_fuse(False)  # the per-word pass instead of the tiled one
        try:
            assert np.array_equal(emu_engine.segment_flow(flow) > 0, mask), name
        finally:
            emu_engine.debug_fuse(True)
    flow = next(f for n, f, _, _ in helpers.segment_cases(golden.seg) if n == "noise_97x131")
    for shape, ksize, its in [(0, (3, 5), 1), (1, (7, 7), 2), (2, (4, 9), 3), (2, (31, 17), 1), (0, (32, 32), 1), (2, (1, 1), 2),
                              (2, (10, 10), 1), (2, (15, 15), 2), (0, (2, 2), 4)]:
        want = so.process_flow_region(flow, 0.7, ksize, its, shape)
        got = emu_engine.segment_flow(flow, 0.7, ksize, its, shape)
        assert np.array_equal(got, want), (shape, ksize, its)
        assert np.array_equal(emu_engine.structuring_element(shape, ksize), so.structuring_element(shape, ksize))
    for bad in (dict(ksize=(33, 3)), dict(ksize=(0, 3)), dict(shape=3), dict(iterations=-1)):
        with pytest.raises(Exception):
            emu_engine.segment_flow(flow, **bad)
    with pytest.raises(Exception):
        emu_engine.segment_flow(np.zeros((0, 4, 2), np.float32))
    # negative threshold: every pixel passes; NaN magnitude never does
    f = np.zeros((5, 40, 2), np.float32)
    f[2, 7] = np.nan
    m = emu_engine.segment_flow(f, seg_th=-1.0, iterations=0)
    assert m[2, 7] == 0 and (m > 0).sum() == 5 * 40 - 1


def test_emu_frames_vs_reference(emu_engine, golden):
    """Grey conversion and flow-extrapolation kernels (emulator build) vs cv2 / the reference's task_results."""
    helpers.check_frames_against_golden(golden.frames, emu_engine.cvtColor_rgb2gray, emu_engine.remap_flow,
                                        emu_engine.predict_frame)
    with pytest.raises(Exception):
        emu_engine.cvtColor_rgb2gray(np.zeros((4, 4), np.uint8))
    with pytest.raises(Exception):
        emu_engine.remap_flow(np.zeros((4, 4, 3), np.uint8), np.zeros((0, 3, 2), np.float32), 0, 0)
    # flow_sign -1 == the reference's `flow = -flow` applied first
    nf, flow = golden.frames["pred__next"], golden.frames["pred__flow"]
    assert np.array_equal(emu_engine.remap_flow(nf, -flow[20:60, 30:90], 30, 20, flow_sign=-1),
                          emu_engine.remap_flow(nf, flow[20:60, 30:90], 30, 20))


def test_emu_realdata_pair_end_to_end(emu_engine, nsof):
    """One real grasp pair (configs[0]) through the reference-shaped host seam on the emulator build:
    opticalFlow3D -> -flow -> task_results -> pixel accuracy, against the reference's own outputs."""
    rd = helpers.RealData("grasp")
    cfg = nsof.RoiConfig.grasp()
    i = rd.keep[1]
    a, b = rd.pair(i)
    memimg = rd.g[f"memimg__{i}"]
    flow, _, _, _, info = emu_engine.opticalFlow3D(memimg, memimg, a, b, cfg.MEMSIZE, cfg.MEMSIZE, cfg)
    assert info == rd.rect(i)
    x0, y0, x1, y1 = info
    rd.check_flow(i, flow[y0:y1, x0:x1].astype(np.float32), "emu vs cv2")
    motion = emu_engine.task_results(a, b, -flow, 2, info)
    assert np.array_equal(motion > 0, rd.unpack(f"motion__{i}"))
    assert helpers.pixel_accuracy(motion > 0, rd.unpack(f"gt__{i}")) == rd.g["acc1"][i]


@pytest.mark.parametrize("ds,padding", [("uav", 30), ("uavnew2", 60), ("tabletennis", 20)])
def test_emu_realdata_small_sequences(emu_engine, nsof, ds, padding):
    """The three sequences without a script of their own (data/<ds>/Parameters.txt: FLAG 1 with poly_n 10 / winsize 3,
    FLAG 1 with the grasp set, FLAG 2 with poly_n 1 / winsize 4) on the emulator build, first pairs of each."""
    helpers.check_realdata_small(emu_engine, nsof, ds, padding, pairs=3 if ds == "uavnew2" else 6)


def test_emu_frame_mode_vs_external_fixtures(emu_engine):
    """The stage-1 frame-mode kernels (dense Lanczos matrices; Euler chain with the binomial-series carry between exact
    re-evaluations) against the independent 50-digit fixtures of tests/golden/make_stage1_external.py."""
    ext = helpers.Stage1External()
    ext.check_compress(lambda img, m, n: emu_engine.compress_images(img, m, n)[0], "emu kernels")
    sim = lambda cells, w0, th1, th2: emu_engine.simulate_memristor_array(cells, th1=th1, th2=th2, w0=w0)
    ext.check_euler(sim, "emu kernels (run)", tol=1e-10)
    sim2 = lambda cells, w0, th1, th2: emu_engine.simulate_memristor_array_stepwise(cells, th1=th1, th2=th2, w0=w0)
    ext.check_euler(sim2, "emu kernels (stepwise)", tol=1e-10)


def test_emu_ssim(emu_engine, golden):
    helpers.check_ssim(emu_engine.structural_similarity, golden.frames, "emu")
    with pytest.raises(ValueError):
        emu_engine.structural_similarity(np.zeros((6, 20), np.uint8), np.zeros((6, 20), np.uint8))
    with pytest.raises(ValueError):
        emu_engine.structural_similarity(np.zeros((8, 20), np.uint8), np.zeros((8, 21), np.uint8))


def test_shared_frame_expansion_bit_identical(emu_engine):
    """Consecutive pairs of a clip under one ROI rectangle: R of the shared frame is computed once per batch."""
    import ctypes as C
    eng = emu_engine
    to_dev = lambda f: f
    new_out = lambda hw: np.zeros(hw + (2,), np.float32)
    ptr = lambda a: a.ctypes.data
    stream = lambda: 0
    to_host = lambda o: o

    def device_batch(frames, crops, t0s, share):
        eng.lib.check(eng.lib.nsof_debug_share_frames(eng.handle, 1 if share else 0))
        try:
            H, W = frames[0].shape
            dev = [to_dev(f) for f in frames]
            outs = [new_out((sl[0].stop - sl[0].start, sl[1].stop - sl[1].start)) for sl in crops]
            items = []
            for sl, t, o in zip(crops, t0s, outs):
                off = sl[0].start * W + sl[1].start
                h, w = sl[0].stop - sl[0].start, sl[1].stop - sl[1].start
                items.append((ptr(dev[t]) + off, ptr(dev[t + 1]) + off, h, w, W, W, ptr(o), w * 8))
            eng.farneback_batch_device(items, stream(), **helpers.PARAM_SETS["grasp"])
            return [to_host(o) for o in outs]
        finally:
            eng.lib.check(eng.lib.nsof_debug_share_frames(eng.handle, 1))

    helpers.check_shared_frames(eng, device_batch)


def test_float32_frames_vs_cv2_golden(emu_engine, golden):
    """cv2 accepts CV_32F frames; they take the typed direct pyramid kernel.  Strided views included."""
    eng = emu_engine
    helpers.check_float32_frames(golden, lambda a, b, **p: eng.calcOpticalFlowFarneback(a, b, None, **p), "emu vs cv2")
    g = golden.fb_f32
    a, b = g["f32_96x128__prev"], g["f32_96x128__next"]
    big_a, big_b = np.zeros((120, 160), np.float32), np.zeros((120, 160), np.float32)
    big_a[10:106, 20:148], big_b[10:106, 20:148] = a, b
    got = eng.calcOpticalFlowFarneback(big_a[10:106, 20:148], big_b[10:106, 20:148], None, **helpers.PARAM_SETS["grasp"])
    assert np.array_equal(got, eng.calcOpticalFlowFarneback(a, b, None, **helpers.PARAM_SETS["grasp"]))
    mixed = eng.calcOpticalFlowFarneback(a, np.rint(b).astype(np.uint8), None, **helpers.PARAM_SETS["grasp"])  # mixed: both as float32
    assert mixed.shape == got.shape


@pytest.mark.parametrize("pname", ["grasp", "autodriving_txt", "tabletennis"])
def test_emu_r1_window_bit_identical(emu_engine, pname):
    """Blur+solve with the R1 window in shared memory (default) against the L2 gathers, both tile shapes, all three
    window radii the shipped parameter sets use."""
    stats = helpers.check_r1_window(emu_engine, hw=(200, 280), pname=pname)
    if pname == "grasp":
        assert stats["shift3", 1][1] == 0     # a uniform 3 px shift never leaves the window


def test_pipeline_flag_checks(emu_engine, nsof):
    """RoiConfig.FLAG in the sequence pipelines: 1 and 2 are served (1 = one item per connected region); the packed host
    path hands out one crop per region and refuses masks for FLAG 1 instead of silently merging; anything else is an error."""
    import importlib
    pm = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.pipeline")
    cfg = nsof.RoiConfig.uavnew2()
    assert cfg.FLAG == 1 and cfg.PADDING == 60
    with pytest.raises(NotImplementedError):   # a mask needs the pasted frame
        pm.SequencePipeline(emu_engine, 1, 2, 600, 600, cfg, packed=True, segment=True)
    cfg.FLAG = 3
    with pytest.raises(ValueError):
        pm.SequencePipeline(emu_engine, 1, 2, 600, 600, cfg)


def test_emu_small_radius_blur_paths(emu_engine, golden):
    helpers.check_small_radius_blur(emu_engine, golden, hw=(150, 210))


def test_emu_fused_pyramid_all_poly_n(emu_engine):
    helpers.check_fused_pyramid_all_n(emu_engine, exact=True)
