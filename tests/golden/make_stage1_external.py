#!/usr/bin/env python3
"""External fixtures for the FRAME-MODE synapse simulator (simulation/simulationcode_v4_transistor_vehicle.m).

There is no MATLAB / Octave in the build container or on the GPU box, and the reference ships no output of its
`.m` scripts, so frame mode cannot be pinned to the reference itself ("parity unpinned").  This
script produces the next best thing: results of an INDEPENDENT derivation in extended precision (mpmath, 50
digits) that both the numpy restatement (oracle/stage1_oracle.py) and the CUDA kernels are held to.

* `lanczos__<n_in>_<n_out>`: the dense antialiased Lanczos-3 contribution matrix of `imresize(..., 'lanczos3')`
  (vehicle.m:111-121) for the ratios the configs use (1080->13, 1920->24, 801->4, 161->4, 160->4, 600->15).  Derived
  from the definition rather than from imresize.m's index arithmetic: output sample j sits at input coordinate
  c_j = (j + 1/2) / s - 1/2 (s = n_out / n_in < 1), every integer tap k in Z contributes sinc(s (c_j - k)) sinc(s (c_j - k) / 3)
  for |s (c_j - k)| < 3 to the input pixel that symmetric reflection (... 2 1 0 | 0 1 2 ...) maps k to, and rows are
  normalised to sum 1.  No tap window, no `left` / `P` / `aux(mod(...))` bookkeeping -- if the restatement gets
  that bookkeeping wrong, it disagrees with this matrix.
* `pillow__*`: Pillow's `Image.resize(LANCZOS)` on a mode-F image, a second independent implementation of the same
  antialiased kernel (its edge handling differs -- truncate + renormalise instead of reflect -- so only interior
  cells are comparable, and its coefficients are float32-ish: a loose 1e-6 cross-check).
* `euler__*`: the 1000-sub-step forward-Euler recurrence of `update_state` (vehicle.m:173-184, 217-219) in 50-digit
  arithmetic for 64 (w0, |delta|) points spanning the potentiation branch, the decay branch, the dead zone
  (th1 = 7) and both clips (w = 0, w = 1), over three consecutive frames.  Constants are the IEEE doubles MATLAB
  parses (51.03, -5.12, ... and dt / nSubSteps evaluated in double), so the only difference to a float64
  implementation is rounding inside the recurrence.

Run from the repo root:  python tests/golden/make_stage1_external.py   (needs mpmath and Pillow; ~1 min)
"""
import os

import mpmath as mp
import numpy as np

OUT = os.path.dirname(os.path.abspath(__file__))
mp.mp.dps = 50

PARAMS = dict(alphaoff=1, alphaon=1, voff=-0.2, von=0.1, koff=51.03, kon=-2.91, son=0.2, soff=0.8, bon=-5.12,
              boff=3.10, Ron=163_305, Roff=2_104_377, won=1, woff=0, wini=0.5)   # vehicle.m:26-33


def lanczos_matrix(n_in, n_out):
    s = mp.mpf(n_out) / mp.mpf(n_in)
    assert s < 1
    W = [[mp.mpf(0)] * n_in for _ in range(n_out)]
    reach = int(mp.ceil(3 / s)) + 2
    for j in range(n_out):
        c = (mp.mpf(j) + mp.mpf(1) / 2) / s - mp.mpf(1) / 2
        k0 = int(mp.floor(c))
        tot = mp.mpf(0)
        for k in range(k0 - reach, k0 + reach + 1):
            x = s * (c - k)
            if abs(x) >= 3:
                continue
            wt = mp.mpf(1) if x == 0 else mp.sin(mp.pi * x) * mp.sin(mp.pi * x / 3) / (mp.pi ** 2 * x ** 2 / 3)
            i = k
            while i < 0 or i >= n_in:      # symmetric reflection about the image edges
                i = -i - 1 if i < 0 else 2 * n_in - 1 - i
            W[j][i] += wt
            tot += wt
        W[j] = [v / tot for v in W[j]]
    return np.array([[float(v) for v in row] for row in W], dtype=np.float64)


def euler_mp(w0, d, th1, th2, n_sub=1000, frames=3, dt=5e-4):
    """-> w after each of `frames` frames with the same |delta| = d (vehicle.m:146-184, 204-221, 306-321)."""
    P = {k: mp.mpf(float(v)) for k, v in PARAMS.items()}
    d = mp.mpf(float(d))
    # calculate_difference_matrix: th2, then the middle band, then th1 (later assignments win)
    V = mp.mpf(0)
    if d > th2:
        V = (d + 4) * mp.mpf(0.75)
    if th1 < d <= th2:
        V = (d + 4) * mp.mpf(0.75)
    if d <= th1:
        V = (d - mp.mpf(5.5)) * mp.mpf(0.6)
    # modulatefunc
    if V > 0:
        v = -(mp.mpf(0.3) * V + 0)
    elif V < 0:
        v = -(3 * V + -3)
    else:
        v = mp.mpf(0)
    dt_sub = mp.mpf(float(dt) / n_sub)
    w = mp.mpf(float(w0))
    out = []
    for _ in range(frames):
        for _ in range(n_sub):
            if v < P["voff"]:
                dw = P["koff"] * (v / P["voff"] - 1) ** P["alphaoff"] * (1 - w * P["soff"]) ** P["boff"]
            elif v > P["von"]:
                dw = P["kon"] * (v / P["von"] - 1) ** P["alphaon"] * (1 - w * P["son"]) ** P["bon"]
            else:
                dw = mp.mpf(0)
            w = w + dw * dt_sub
            w = max(mp.mpf(0), min(w, mp.mpf(1)))
        out.append(float(w))
    return out, float(v)


def main():
    out = {}
    ratios = [(1080, 13), (1920, 24), (801, 4), (161, 4), (160, 4), (600, 15)]
    for n_in, n_out in ratios:
        out[f"lanczos__{n_in}_{n_out}"] = lanczos_matrix(n_in, n_out)
        print("lanczos", n_in, n_out, out[f"lanczos__{n_in}_{n_out}"].sum(1)[:3])
    out["lanczos_ratios"] = np.array(ratios)
    # Pillow cross-check on seeded images
    from PIL import Image
    import sys
    sys.path.insert(0, os.path.dirname(OUT))
    import helpers
    for H, W, M in [(1080, 1920, 80), (801, 801, 200)]:
        img = helpers.external_image(H, W)     # seeded; regenerated by the tests instead of being stored
        pil = Image.fromarray(img.astype(np.float32) / np.float32(255.0), mode="F")
        res = np.asarray(pil.resize((W // M, H // M), Image.LANCZOS), dtype=np.float64)
        out[f"pillow__{H}x{W}__imgsum"] = np.array(int(img.astype(np.int64).sum()))
        out[f"pillow__{H}x{W}__cells"] = res
        out[f"pillow__{H}x{W}__m"] = np.array(M)
    # Euler trajectories
    rng = np.random.default_rng(7)
    pts = []
    for w0 in (0.0, 0.02, 0.242, 0.5, 0.9, 1.0):
        for d in (0.0, 0.5, 1.9, 2.0, 2.000001, 3.0, 10.0, 60.0):
            pts.append((w0, d, 2.0, 1.5))               # vehicle thresholds
    for _ in range(8):
        pts.append((float(rng.random()), float(rng.uniform(0, 40)), 0.7, 1.5))   # uav thresholds (uav.m:40-41)
    for d in (5.4, 5.5, 5.6, 6.0, 6.6, 6.7, 7.0, 7.1):
        pts.append((0.5, d, 7.0, 1.5))                  # th1 = 7: the dead zone voff <= v <= von is reachable
    W3, Vm = [], []
    for (w0, d, th1, th2) in pts:
        ws, v = euler_mp(w0, d, th1, th2)
        W3.append(ws)
        Vm.append(v)
    out["euler__points"] = np.array(pts, dtype=np.float64)       # (w0, |delta|, th1, th2)
    out["euler__w"] = np.array(W3, dtype=np.float64)             # (n, 3): after frames 1, 2, 3
    out["euler__vmod"] = np.array(Vm, dtype=np.float64)
    out["mpmath_version"] = np.array(mp.__version__)
    np.savez_compressed(os.path.join(OUT, "stage1_external.npz"), **out)
    print("stage1_external.npz", len(pts), "Euler points; w range", out["euler__w"].min(), out["euler__w"].max())


if __name__ == "__main__":
    main()
