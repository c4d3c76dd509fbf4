"""Pins the oracle (oracle/farneback_oracle.c, oracle/stage1_oracle.py) against the golden vectors the
reference's own code produced (tests/golden/make_golden.py).  CPU only."""
import numpy as np
import pytest

import helpers
from oracle import stage1_oracle as s1


@pytest.mark.parametrize("pname", list(helpers.PARAM_SETS))
def test_farneback_oracle_matches_cv2_golden(oracle, golden, pname):
    p = helpers.PARAM_SETS[pname]
    for case in golden.fb["cases"]:
        prev, nxt = golden.fb[f"{case}__prev"], golden.fb[f"{case}__next"]
        want = golden.fb[f"{case}__{pname}"]
        got = oracle.farneback(prev, nxt, **p)
        helpers.assert_flow_close(got, want, pname, f"oracle vs cv2 {case}")


def test_farneback_oracle_initial_flow_matches_cv2(oracle, golden):
    helpers.check_initial_flow(golden, lambda a, b, f0, **p: oracle.farneback(a, b, flow=f0, **p), "oracle vs cv2")


def test_resize_area_oracle_matches_cv2():
    """INTER_AREA of a 2-channel float image (the initial-flow downscale), bit for bit against cv2.resize: integer
    ratios (the unrolled fast path), fractional ratios (DecimateAlpha tables), identity."""
    import cv2
    o = helpers.load_oracle()
    rng = np.random.default_rng(7)
    for hs, ws, hd, wd in [(101, 75, 13, 9), (96, 128, 12, 16), (96, 128, 48, 64), (75, 101, 27, 36), (40, 33, 40, 33),
                           (100, 130, 36, 47), (90, 120, 30, 40), (421, 523, 53, 65), (64, 64, 8, 64)]:
        src = (rng.normal(size=(hs, ws, 2)) * 5).astype(np.float32)
        want = cv2.resize(src, (wd, hd), interpolation=cv2.INTER_AREA)
        assert np.array_equal(o.resize_area(src, hd, wd), want), (hs, ws, hd, wd)


def test_level_geometry_matches_spec(oracle):
    # 1080x1920 at (0.5, 3) -> 4 scales, ksize 3,3,9,19
    g = oracle.level_geometry(1080, 1920, 0.5, 3)
    assert [(a, b, c) for a, b, c, _ in g] == [(1080, 1920, 3), (540, 960, 3), (270, 480, 9), (135, 240, 19)]
    # truncation while min(h, w) * scale < 32
    assert len(oracle.level_geometry(40, 500, 0.5, 3)) == 1
    assert len(oracle.level_geometry(64, 64, 0.5, 3)) == 2


@pytest.mark.parametrize("ds,cfg", [("grasp", dict(MEMSIZE=80, THRES=250, ext=(20, 20, 20, 20), hw=(1920, 1080))),
                                    ("autodriving", dict(MEMSIZE=200, THRES=114, ext=(60, 60, 60, 60), hw=(801, 801))),
                                    ("uav", dict(MEMSIZE=40, THRES=114, ext=(30, 30, 30, 30), hw=(161, 161))),
                                    ("uavnew2", dict(MEMSIZE=40, THRES=245, ext=(60, 60, 60, 60), hw=(600, 600))),
                                    ("tabletennis", dict(MEMSIZE=10, THRES=245, ext=(20, 20, 20, 20), hw=(160, 160)))])
def test_roi_oracle_matches_reference(golden, ds, cfg):
    cur = golden.roi[f"{ds}__currents"]
    memimg = s1.current_to_u8(cur)
    assert np.array_equal(memimg, golden.roi[f"{ds}__memimg"])
    for t in range(cur.shape[0]):
        for flag in (1, 2):
            n, labels, stats, rects = s1.roi_from_memimg(memimg[t], cfg["hw"], cfg["MEMSIZE"], cfg["MEMSIZE"], cfg["THRES"],
                                                         cfg["ext"], flag)
            want_n = int(golden.roi[f"{ds}__flag{flag}__nrect"][t])
            assert len(rects) == want_n
            assert [tuple(r) for r in golden.roi[f"{ds}__flag{flag}__rects"][t, :want_n]] == rects
        assert np.array_equal(labels, golden.roi[f"{ds}__labels"][t])
        assert np.array_equal(stats[1:], golden.roi[f"{ds}__stats"][t, :n - 1])
        assert n == golden.roi[f"{ds}__num_labels"][t] or n == 1


def test_cc_oracle_matches_cv2_on_random_masks(golden):
    for i, m in enumerate(golden.roi["random__masks"]):
        n, labels, stats = s1.connected_components_4(m)
        assert n == golden.roi["random__num_labels"][i]
        assert np.array_equal(labels, golden.roi["random__labels"][i])
        assert np.array_equal(stats[1:], golden.roi["random__stats"][i, :n - 1])


def test_event_oracle_matches_reference(golden):
    ev = golden.ev
    x, y, t = ev["x"], ev["y"], ev["t"]
    H, W = int(ev["H"]), int(ev["W"])
    b = s1.slice_bounds(t, 1000)
    assert np.array_equal(b, ev["slice_bounds"])
    for k, i in enumerate(ev["hist_slices"]):
        sl = slice(b[i], b[i + 1])
        assert np.array_equal(s1.bincount_2d(x[sl], y[sl], H, W), ev["hist"][k])
    assert np.array_equal(s1.update_state(ev["us_w0"], ev["us_V0"]), ev["us_w1"])
    for k, i in enumerate(ev["w_snap_idx"][:3]):
        assert np.array_equal(s1.scheme1(x, y, t, H, W, 1000, n_slices=int(i) + 1), ev["w_snaps"][k])
    assert np.array_equal(s1.scheme1(x, y, t, H, W, 1000, silent_v=0.5, n_slices=64), ev["w_decay64"])


def test_frame_mode_oracle_sanity():
    """Frame mode is parity-unpinned (no MATLAB here): check the properties the .m text implies."""
    wts, idx = s1.lanczos3_contributions(800, 4)
    assert np.allclose(wts.sum(1), 1.0) and idx.min() >= 0 and idx.max() < 800
    flat = np.full((160, 160), 77, np.uint8)
    assert np.allclose(s1.compress_image(flat, 40, 40), 77 / 255.0, atol=1e-12)  # weights sum to 1
    # a static scene decays towards w = 0 (d = 0 <= th1 -> func1 < 0 -> v_mod > von -> kon < 0)
    cells = np.full((3, 2, 2), 0.5)
    w, cur = s1.simulate_memristor_array(cells, nSubSteps=50)
    assert (w < 0.5).all() and np.allclose(cur[0], 1 / (163305 / np.exp(-np.log(2104377 / 163305) * 0.5)))


def test_event_scheme2_oracle_matches_reference(golden):
    """The numpy restatement of the scheme-2 loop against w_final of the reference's own simulate() (split with the
    generator's +/-1 polarities, split with 0/1 polarities and a decaying bias, magnitude mode): bit for bit."""
    for name, kw, w_a, w_b in helpers.scheme2_cases(golden.ev2):
        got_a, got_b = s1.scheme2(**kw)
        assert np.array_equal(got_a, w_a), name
        assert (got_b is None) == (w_b is None)
        if w_b is not None:
            assert np.array_equal(got_b, w_b), name


def test_segment_oracle_matches_reference(golden):
    """oracle/seg_oracle.py vs masks of the reference's own process_flow_region (cv2 cartToPolar / dilate / erode),
    the cv2 structuring elements and single cv2 dilate / erode passes: bit for bit."""
    from oracle import seg_oracle as so
    seg = golden.seg
    for name, flow, mask, mag_gt in helpers.segment_cases(seg):
        assert np.array_equal(so.process_flow_region(flow, iterations=0) > 0, mag_gt), name
        assert np.array_equal(so.process_flow_region(flow) > 0, mask), name
        assert np.array_equal(so.process_flow_region(-flow.astype(np.float64)) > 0, mask), name  # sign, dtype of :461
    base = None
    for key in seg.files:
        if key.startswith("element__"):
            shape, kw, kh = (int(v) for v in key.split("__")[1].split("_"))
            assert np.array_equal(so.structuring_element(shape, (kw, kh)), seg[key]), key
        elif key.startswith(("dilate__", "erode__")):
            if base is None:
                base = next(m for n, _, _, m in helpers.segment_cases(seg) if n == "noise_97x131")
            shape, kw, kh = (int(v) for v in key.split("__")[1].split("_"))
            got = so._morph(base, so.structuring_element(shape, (kw, kh)), erode=key.startswith("erode"))
            assert np.array_equal(got, np.unpackbits(seg[key], axis=1)[:, :base.shape[1]].astype(bool)), key
    frame = np.zeros((60, 80, 2), np.float32)
    frame[10:43, 5:70] = next(f for n, f, _, _ in helpers.segment_cases(seg) if n == "noise_33x65")
    mb = so.task_results((60, 80), frame, (5, 10, 70, 43))
    assert mb[10:43, 5:70].any() and not mb[:10].any() and not mb[:, 70:].any()


def test_frames_oracle_matches_reference(golden):
    """oracle/frames_oracle.py vs cv2.cvtColor(RGB2GRAY) on real frames and the reference's own prediction
    task_results (cv2.remap INTER_LINEAR / BORDER_REPLICATE per channel): bit for bit."""
    from oracle import frames_oracle as fo

    def predict(prev, nxt, flow, num_labels, regions, est, merge, padding=0):
        if est == 2:
            return fo.task_results(prev, nxt, flow, regions)
        h, w = prev.shape[:2]
        r = (max(0, min(q[0] for q in regions) - padding), max(0, min(q[1] for q in regions) - padding),
             min(w, max(q[2] for q in regions) + padding), min(h, max(q[3] for q in regions) + padding))
        return fo.task_results(prev, nxt, flow, r)

    helpers.check_frames_against_golden(golden.frames, fo.rgb2gray, fo.remap_flow, predict)


REALDATA_CFG = {"grasp": dict(MEMSIZE=80, OFFSET=0, THRES=250, ext=(20, 20, 20, 20)),
                "autodriving": dict(MEMSIZE=200, OFFSET=15, THRES=114, ext=(60, 60, 60, 60))}


@pytest.mark.parametrize("ds", ["grasp", "autodriving"])
def test_realdata_oracle_end_to_end(oracle, golden, ds):
    """BASELINE configs[0] / configs[1]: the oracle chain (ROI tail -> C Farneback on the crop -> -flow -> motion mask ->
    pixel accuracy) against what the reference's own functions produced on the shipped sequences
    (tests/golden/realdata_<ds>.npz).  Rectangles of EVERY pair of the sequence (slice OFFSET + i: the scripts'
    `mem_state2 = mem_state1` form); flow, mask and accuracy on the kept pairs -- masks bit for bit, accuracy to the
    last digit."""
    from oracle import seg_oracle as so
    rd = helpers.RealData(ds)
    c = REALDATA_CFG[ds]
    M = c["MEMSIZE"]
    memimg_all = s1.current_to_u8(golden.roi[f"{ds}__currents"])
    rects_all = rd.g["rects"]
    for i in range(len(rects_all)):
        _, _, _, r = s1.roi_from_memimg(memimg_all[c["OFFSET"] + i], (rd.H, rd.W), M, M, c["THRES"], c["ext"], 2)
        assert (r[0] if r else (0, 0, 0, 0)) == tuple(int(v) for v in rects_all[i]), (ds, i)
    worst = 0.0
    for i in rd.keep[::2]:   # every other kept pair keeps the CPU suite short; the GPU tests run all of them
        assert np.array_equal(memimg_all[c["OFFSET"] + i], rd.g[f"memimg__{i}"])
        a, b = rd.pair(i)
        x0, y0, x1, y1 = rd.rect(i)
        f = oracle.farneback(np.ascontiguousarray(a[y0:y1, x0:x1]), np.ascontiguousarray(b[y0:y1, x0:x1]),
                             **helpers.PARAM_SETS["grasp"])
        _, mx = rd.check_flow(i, f, "oracle vs cv2")
        worst = max(worst, mx)
        flow = np.zeros((rd.H, rd.W, 2))
        flow[y0:y1, x0:x1] = f
        mb = so.task_results((rd.H, rd.W), -flow, (x0, y0, x1, y1)) > 0
        assert np.array_equal(mb, rd.unpack(f"motion__{i}")), (ds, i)
        assert helpers.pixel_accuracy(mb, rd.unpack(f"gt__{i}")) == rd.g["acc1"][i], (ds, i)
    assert worst <= 1e-3
    # the dataset averages the scripts print (optical_flow_seg.py:619-620; BASELINE.md section 2 for grasp)
    if ds == "grasp":
        assert abs(float(rd.g["acc1_mean"]) - 99.02812363605189) < 1e-9
        assert abs(float(rd.g["acc2_mean"]) - 98.71762912691729) < 1e-9


def test_frame_mode_oracle_vs_external_fixtures():
    """Frame mode (simulation/*.m) cannot be pinned to the reference (no MATLAB, no shipped outputs): the numpy restatement
    is held to an independent 50-digit derivation instead (tests/golden/make_stage1_external.py) -- the dense
    Lanczos-3 contribution matrices for every ratio the configs use, Pillow's LANCZOS resize on interior cells, and
    exact-arithmetic Euler trajectories of update_state over both branches, the dead zone and both clips."""
    ext = helpers.Stage1External()
    for n_in, n_out in ext.g["lanczos_ratios"]:
        wts, idx = s1.lanczos3_contributions(int(n_in), int(n_out))
        dense = np.zeros((n_out, n_in))
        for j in range(n_out):
            np.add.at(dense[j], idx[j], wts[j])
        assert np.abs(dense - ext.g[f"lanczos__{n_in}_{n_out}"]).max() <= 1e-15, (n_in, n_out)
    ext.check_compress(s1.compress_image, "numpy restatement", tol=5e-15)

    def simulate(cells, w0, th1, th2):   # the scalar restatement, per cell, from an arbitrary initial state
        w = w0.copy()
        cur = [1.0 / (163305 / np.exp(-np.log(2104377 / 163305) * (1 - w)))]
        for i in range(3):
            diff = s1.calculate_difference_matrix(cells[i] * 256, cells[i + 1] * 256, th1, th2)
            for x in range(w.shape[1]):
                vm, wv = s1.modulatefunc(diff[0, x]), w[0, x]
                for _ in range(1000):
                    wv = s1.update_state_f64(wv, vm, 5e-4 / 1000)
                w[0, x] = wv
            cur.append(1.0 / (163305 / np.exp(-np.log(2104377 / 163305) * (1 - w))))
        return w, np.stack(cur)

    assert ext.check_euler(simulate, "numpy restatement", tol=1e-13) <= 1e-13


def test_bench_clip_gate_has_margin():
    """The frame-mode gate of the bench workload (synthetic.make_clip -> compress -> Euler -> current -> u8 >= THRES): how
    far every cell's current is from the THRES boundary.  The kernels track the restatement to 1e-10 in w, i.e.
    2.6e-10 relative in the current; the smallest relative distance to the boundary on these clips is > 1e-6, so no
    mask bit can flip (the GPU test compares the masks themselves)."""
    import importlib
    syn = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200.synthetic")
    nsof = importlib.import_module("nsof_b200")
    thr = nsof.u8_threshold_table()[250]
    margin = np.inf
    for seed in (1234, 1235):
        clip = syn.make_clip(seed, 1080, 1920, 8, "cpu").numpy()
        cells = np.stack([s1.compress_image(f, 80, 80) for f in clip])
        _, cur = s1.simulate_memristor_array_vec(cells)
        margin = min(margin, np.abs(cur / thr - 1.0).min())
        assert np.array_equal(s1.current_to_u8(cur) >= 250, cur >= thr)
    assert margin > 1e-6, margin


def test_ssim_oracle_known_answers(golden):
    """structural_similarity restatement (scikit-image is absent here and on the GPU box: parity unpinned against
    skimage itself): identical images -> 1, constant images -> the closed form, symmetry."""
    from oracle import frames_oracle as fo
    helpers.check_ssim(fo.ssim_u8, golden.frames, "oracle")
    nf, pred = golden.frames["pred__next"], golden.frames["pred__whole__out"]
    assert fo.ssim_u8(nf[:, :, 2], pred[:, :, 2]) == fo.ssim_u8(pred[:, :, 2], nf[:, :, 2])
    assert 0.9 < fo.ssim_u8(nf[:, :, 2], pred[:, :, 2]) < 1.0


def test_farneback_oracle_float32_frames_match_cv2(oracle, golden):
    helpers.check_float32_frames(golden, lambda a, b, **p: oracle.farneback(a, b, **p), "oracle vs cv2")
