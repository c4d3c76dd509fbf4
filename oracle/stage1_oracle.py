"""stage1_oracle.py -- TEST INFRASTRUCTURE ONLY (the parity checker for stage 1).

CPU restatement (numpy + small Python loops) of the reference's synaptic temporal-memory stage and
its ROI tail.  Each function cites the reference file:line it follows.  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this module;
the product path never does.

Pinning:
* ROI tail and event mode are PINNED: tests/golden/roi_reference.npz and events_reference.npz were
  produced by the reference's own functions (tests/golden/make_golden.py) and
  tests/test_oracle_golden.py checks this file against them bit for bit (float32 `w`: exactly equal,
  the restatement performs the same numpy operations).
* Frame mode (simulation/*.m) is "parity unpinned": there is no MATLAB/Octave in the build
  container, the reference ships no output of these scripts, and SURVEY.md 3.4 shows the shipped
  .mat fixtures were produced with unpublished parameters.  The restatement follows the .m text and
  MATLAB's documented imresize algorithm.
"""
import numpy as np

# eventsim/event_mem_sim.py:20-27 == simulation/simulationcode_v4_transistor_vehicle.m:26-33
PARAMS = dict(alphaoff=1, alphaon=1, voff=-0.2, von=0.1, koff=51.03, kon=-2.91, son=0.2, soff=0.8, bon=-5.12,
              boff=3.10, Ron=163_305, Roff=2_104_377, won=1, woff=0, wini=0.5)
DT = 5e-4
THETA_EVENTS = 1
REFRACTORY_US = 800


# ----------------------------------------------------------------------------- ROI tail
def current_to_u8(I, a=-3366.0, b=-306.0):
    """optical_flow_seg.py:426-435 (process_mem_state, optical_flow_prediction.py:429-433)."""
    with np.errstate(divide="ignore", invalid="ignore"):
        v = a / np.log10(np.asarray(I, dtype=np.double)) + b
    return np.clip(v, 0, 255).astype(np.uint8)


def update_transition_pic(prev_memristor, transition_pic, thres):
    """optical_flow_seg.py:115-121."""
    for i in range(prev_memristor.shape[0]):
        for j in range(prev_memristor.shape[1]):
            if prev_memristor[i, j] >= thres:
                transition_pic[i, j] = 255
    return transition_pic


def connected_components_4(mask):
    """cv2.connectedComponentsWithStats(mask, connectivity=4) (optical_flow_seg.py:223-224): labels are
    numbered in raster order of each component's first pixel; stats rows = (left, top, width, height, area),
    row 0 the background."""
    m = np.asarray(mask) != 0
    H, W = m.shape
    labels = np.zeros((H, W), np.int32)
    stats = [[0, 0, W, H, 0]]
    n = 0
    for y in range(H):
        for x in range(W):
            if not m[y, x] or labels[y, x]:
                continue
            n += 1
            stack = [(y, x)]
            labels[y, x] = n
            x0 = x1 = x
            y0 = y1 = y
            area = 0
            while stack:
                cy, cx = stack.pop()
                area += 1
                x0, x1, y0, y1 = min(x0, cx), max(x1, cx), min(y0, cy), max(y1, cy)
                for ny, nx in ((cy - 1, cx), (cy + 1, cx), (cy, cx - 1), (cy, cx + 1)):
                    if 0 <= ny < H and 0 <= nx < W and m[ny, nx] and not labels[ny, nx]:
                        labels[ny, nx] = n
                        stack.append((ny, nx))
            stats.append([x0, y0, x1 - x0 + 1, y1 - y0 + 1, area])
    st = np.array(stats, np.int32)
    bg = ~m
    if bg.any():
        ys, xs = np.nonzero(bg)
        st[0] = [xs.min(), ys.min(), xs.max() - xs.min() + 1, ys.max() - ys.min() + 1, bg.sum()]
    else:
        st[0] = [0, 0, 0, 0, 0]
    return n + 1, labels, st


def rect_from_cells(x, y, a, b, pixel_width, pixel_height, ext, w, h):
    """optical_flow_seg.py:134-138 / 180-183; ext = (left, right, upper, lower)."""
    return (max(x * pixel_width - ext[0], 0), max(y * pixel_height - ext[2], 0),
            min((x + a) * pixel_width + ext[1], w), min((y + b) * pixel_height + ext[3], h))


def roi_from_memimg(memimg, frame_hw, pixel_width, pixel_height, thres, ext, flag):
    """opticalFlow3D's ROI part (optical_flow_seg.py:211-241): returns (num_labels, labels, stats, rects)."""
    h, w = frame_hw
    tp = np.zeros((int(h / pixel_height), int(w / pixel_width)))
    tp = update_transition_pic(memimg, tp, thres).astype(np.uint8)
    n, labels, stats = connected_components_4(tp)
    if n == 1:
        return n, labels, stats, []
    if flag == 1:  # process_separate_regions :129-138
        rects = [rect_from_cells(*stats[i, :4], pixel_width, pixel_height, ext, w, h) for i in range(1, n)]
    else:  # process_merged_region :174-183
        x_min = min(stats[i, 0] for i in range(1, n))
        y_min = min(stats[i, 1] for i in range(1, n))
        x_max = max(stats[i, 0] + stats[i, 2] for i in range(1, n))
        y_max = max(stats[i, 1] + stats[i, 3] for i in range(1, n))
        rects = [rect_from_cells(x_min, y_min, x_max - x_min, y_max - y_min, pixel_width, pixel_height, ext, w, h)]
    return n, labels, stats, [tuple(int(v) for v in r) for r in rects]


# ----------------------------------------------------------------------------- event mode
def update_state(w, V, p=PARAMS, dt=DT):
    """eventsim/event_mem_sim.py:40-57 (same numpy operations, same dtypes)."""
    dwdt = np.zeros_like(w, dtype=np.float32)
    mask_off = V < p["voff"]
    mask_on = V > p["von"]
    if mask_off.any():
        dwdt[mask_off] = (p["koff"] * (V[mask_off] / p["voff"] - 1) ** p["alphaoff"]
                          * (1 - w[mask_off] * p["soff"]) ** p["boff"])
    if mask_on.any():
        dwdt[mask_on] = (p["kon"] * (V[mask_on] / p["von"] - 1) ** p["alphaon"]
                         * (1 - w[mask_on] * p["son"]) ** p["bon"])
    return np.clip(w + dwdt * dt, 0.0, 1.0)


def resistance_exp(w, p=PARAMS):
    """eventsim/event_mem_sim.py:60-63."""
    lam = np.log(p["Roff"] / p["Ron"])
    return p["Ron"] / np.exp(-lam * (1.0 - w))


def slice_bounds(t, slice_us):
    """eventsim/event_mem_sim.py:78-83 (the searchsorted bounds slice_indices yields pairs of)."""
    bounds = np.arange(t[0], t[-1] + slice_us, slice_us, dtype=t.dtype)
    return np.searchsorted(t, bounds)


def bincount_2d(x, y, H, W):
    """eventsim/event_mem_sim.py:100-104."""
    lin = y.astype(np.int64) * W + x.astype(np.int64)
    return np.bincount(lin, minlength=H * W).reshape(H, W).astype(np.int32)


def scheme1(x, y, t, H, W, slice_us, active_v=-6.0, silent_v=0.0, n_slices=None, p=PARAMS, theta=THETA_EVENTS):
    """Scheme-1 loop of simulate() (eventsim/event_mem_sim.py:204-228); returns w after n_slices windows."""
    idx = slice_bounds(t, slice_us)
    w = p["wini"] * np.ones((H, W), dtype=np.float32)
    ns = len(idx) - 1 if n_slices is None else n_slices
    for i in range(ns):
        V = np.full((H, W), silent_v, dtype=np.float32)
        if idx[i + 1] > idx[i]:
            sl = slice(idx[i], idx[i + 1])
            m = bincount_2d(x[sl], y[sl], H, W) >= theta
            if m.any():
                V[m] = active_v
        w = update_state(w, V, p)
    return w


REFRACTORY_US = 800  # eventsim/event_mem_sim.py:34


def scheme2(x, y, pol, t, H, W, slice_us, active_v=-6.0, silent_v=0.0, polarity="split", p=PARAMS):
    """Scheme-2 loop of simulate() (eventsim/event_mem_sim.py:231-286): DC bias + event overlay with a per-pixel
    refractory period.  polarity 'split': ON events (pol == 1) drive array A, OFF events (pol == 0) array B;
    'magnitude': every event drives array A.  Returns (w_a, w_b or None)."""
    idx = slice_bounds(t, slice_us)
    two = polarity == "split"
    w_a = p["wini"] * np.ones((H, W), dtype=np.float32)
    w_b = w_a.copy() if two else None
    next_a = np.zeros((H, W), dtype=np.int64)
    next_b = np.zeros((H, W), dtype=np.int64)

    def overlay(V, next_ok, xs, ys, t_start, t_last):
        if xs.size == 0:
            return
        ok = next_ok[ys, xs] <= t_start
        if np.any(ok):
            V[ys[ok], xs[ok]] += active_v  # duplicates apply once (numpy fancy-index semantics)
            next_ok[ys[ok], xs[ok]] = t_last + REFRACTORY_US

    for i in range(len(idx) - 1):
        a, b = idx[i], idx[i + 1]
        Va = np.full((H, W), float(silent_v), dtype=np.float32)
        Vb = np.full((H, W), float(silent_v), dtype=np.float32) if two else None
        if b > a:
            xs, ys, ps = x[a:b], y[a:b], pol[a:b]
            if two:
                overlay(Va, next_a, xs[ps == 1], ys[ps == 1], t[a], t[b - 1])
                overlay(Vb, next_b, xs[ps == 0], ys[ps == 0], t[a], t[b - 1])
            else:
                overlay(Va, next_a, xs, ys, t[a], t[b - 1])
        w_a = update_state(w_a, Va, p)
        if two:
            w_b = update_state(w_b, Vb, p)
    return w_a, w_b


# ----------------------------------------------------------------------------- frame mode (unpinned)
def _lanczos3(x):
    """MATLAB imresize.m lanczos3 kernel."""
    eps = np.finfo(np.float64).eps
    f = (np.sin(np.pi * x) * np.sin(np.pi * x / 3) + eps) / ((np.pi ** 2 * x ** 2 / 3) + eps)
    return f * (np.abs(x) < 3)


def lanczos3_contributions(n_in, n_out):
    """MATLAB imresize.m `contributions` with antialiasing: weights (n_out, P) and 0-based indices."""
    scale = n_out / n_in
    kw = 6.0
    if scale < 1:
        h = lambda x: scale * _lanczos3(scale * x)
        kw = kw / scale
    else:
        h = _lanczos3
    x = np.arange(1, n_out + 1, dtype=np.float64)
    u = x / scale + 0.5 * (1 - 1 / scale)
    left = np.floor(u - kw / 2)
    P = int(np.ceil(kw)) + 2
    ind = left[:, None] + np.arange(P)[None, :]
    wts = h(u[:, None] - ind)
    wts = wts / wts.sum(axis=1, keepdims=True)
    aux = np.concatenate([np.arange(1, n_in + 1), np.arange(n_in, 0, -1)])
    ind = aux[np.mod(ind.astype(np.int64) - 1, len(aux))]
    return wts, ind - 1


def compress_image(image_u8, m, n):
    """compress_image (simulationcode_v4_transistor_vehicle.m:111-121): im2double + imresize lanczos3,
    rows first (MATLAB resizes the dimensions in order of increasing scale; ties keep dim 1 first)."""
    img = np.asarray(image_u8, dtype=np.float64) / 255.0
    H, W = img.shape
    Hc, Wc = H // n, W // m
    wy, iy = lanczos3_contributions(H, Hc)
    wx, ix = lanczos3_contributions(W, Wc)
    tmp = np.einsum("ck,ckx->cx", wy, img[iy, :])          # (Hc, W)
    return np.einsum("ck,yck->yc", wx, tmp[:, ix])          # (Hc, Wc)


def calculate_difference_matrix(a, b, th1, th2):
    """vehicle.m:146-171 with func1-3."""
    d = np.abs(np.asarray(a, np.float64) - np.asarray(b, np.float64))
    r = np.zeros_like(d)
    r[d > th2] = (d[d > th2] + 4) * 0.75
    mid = (d > th1) & (d <= th2)
    r[mid] = (d[mid] + 4) * 0.75
    r[d <= th1] = (d[d <= th1] - 5.5) * 0.6
    return r


def modulatefunc(V):
    """vehicle.m:306-321 (V == 0 leaves v_mod undefined in MATLAB; 0 here, as in libnsof)."""
    if V > 0:
        v = 0.3 * V + 0
    elif V < 0:
        v = 3 * V + -3
    else:
        v = 0.0
    return -v


def update_state_f64(w, V, dt, p=PARAMS):
    """vehicle.m:173-184 (scalar, double)."""
    if V < p["voff"]:
        dwdt = p["koff"] * (V / p["voff"] - 1) ** p["alphaoff"] * (1 - w * p["soff"]) ** p["boff"]
    elif V > p["von"]:
        dwdt = p["kon"] * (V / p["von"] - 1) ** p["alphaon"] * (1 - w * p["son"]) ** p["bon"]
    else:
        dwdt = 0.0
    return max(0.0, min(w + dwdt * dt, 1.0))


def simulate_memristor_array(compressed, dt=5e-4, nSubSteps=1000, th1=2.0, th2=1.5, p=PARAMS, vds=1.0):
    """vehicle.m:187-227; returns (w_array, currents (T, Hc, Wc)) with currents[0] the initial state."""
    compressed = np.asarray(compressed, np.float64)
    T, Hc, Wc = compressed.shape
    dt_sub = dt / nSubSteps
    w = p["wini"] * np.ones((Hc, Wc))
    lam = np.log(p["Roff"] / p["Ron"])
    R = lambda w_: p["Ron"] / np.exp(-lam * (1 - w_))
    cur = [vds / R(w)]
    for i in range(T - 1):
        diff = calculate_difference_matrix(compressed[i] * 256, compressed[i + 1] * 256, th1, th2)
        for y in range(Hc):
            for x in range(Wc):
                vm = modulatefunc(diff[y, x])
                wv = w[y, x]
                for _ in range(nSubSteps):
                    wv = update_state_f64(wv, vm, dt_sub, p)
                w[y, x] = wv
        cur.append(vds / R(w))
    return w, np.stack(cur)


def simulate_memristor_array_vec(compressed, dt=5e-4, nSubSteps=1000, th1=2.0, th2=1.5, p=PARAMS, vds=1.0):
    """Same recurrence as simulate_memristor_array, vectorised over the cells (the sub-step loop stays
    sequential).  Used as the timed CPU baseline of stage 1; tests check it equals the scalar version."""
    compressed = np.asarray(compressed, np.float64)
    T = compressed.shape[0]
    dt_sub = dt / nSubSteps
    w = p["wini"] * np.ones(compressed.shape[1:])
    lam = np.log(p["Roff"] / p["Ron"])
    R = lambda w_: p["Ron"] / np.exp(-lam * (1 - w_))
    cur = [vds / R(w)]
    for i in range(T - 1):
        diff = calculate_difference_matrix(compressed[i] * 256, compressed[i + 1] * 256, th1, th2)
        vm = -np.where(diff > 0, 0.3 * diff + 0, np.where(diff < 0, 3 * diff + -3, 0.0))
        off, on = vm < p["voff"], vm > p["von"]
        g_off = p["koff"] * (vm[off] / p["voff"] - 1) ** p["alphaoff"]
        g_on = p["kon"] * (vm[on] / p["von"] - 1) ** p["alphaon"]
        w_off, w_on = w[off], w[on]
        for _ in range(nSubSteps):
            w_off = np.clip(w_off + g_off * (1 - w_off * p["soff"]) ** p["boff"] * dt_sub, 0.0, 1.0)
            w_on = np.clip(w_on + g_on * (1 - w_on * p["son"]) ** p["bon"] * dt_sub, 0.0, 1.0)
        w = w.copy()
        w[off], w[on] = w_off, w_on
        cur.append(vds / R(w))
    return w, np.stack(cur)


def event_cell_currents(w, memsize, p=PARAMS):
    """Event-mode ROI glue (builder-defined): per-pixel current 1 / resistance_exp(w),
    block-max pooled to memsize x memsize cells."""
    H, W = w.shape
    Hc, Wc = H // memsize, W // memsize
    cur = 1.0 / resistance_exp(w, p)
    return cur[:Hc * memsize, :Wc * memsize].reshape(Hc, memsize, Wc, memsize).max(axis=(1, 3))
