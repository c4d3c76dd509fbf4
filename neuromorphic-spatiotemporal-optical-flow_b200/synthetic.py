"""Synthetic inputs of the named shapes (BASELINE.json configs 3-5).

* `make_clip`: a grey uint8 clip -- static blurred-noise background, one brighter textured rectangle moving
  (+3.0, -1.5) px/frame with sub-pixel bilinear compositing.  Written with torch so it runs on the GPU box's
  device (bench data generation, outside any timed region) and on the CPU here (tests); results are the
  input of BOTH the CUDA path and the CPU baseline, so generator details never enter a parity claim.
* `synthetic_events`: an event stream shaped like eventsim/event_mem_sim.py:109-158 (`generate_synthetic_events`:
  a box moving left to right, ON events on the leading edge, OFF on the trailing edge), numpy, deterministic.
"""
from __future__ import annotations

import numpy as np


def _gauss_kernel1d(sigma: float, torch, device):
    r = max(1, int(3 * sigma + 0.5))
    x = torch.arange(-r, r + 1, dtype=torch.float32, device=device)
    k = torch.exp(-0.5 * (x / sigma) ** 2)
    return k / k.sum(), r


def _blur(img, sigma, torch):
    """Separable Gaussian blur with reflect padding; img (H, W) float32."""
    import torch.nn.functional as F
    k, r = _gauss_kernel1d(sigma, torch, img.device)
    x = img[None, None]
    x = F.conv2d(F.pad(x, (r, r, 0, 0), mode="reflect"), k.view(1, 1, 1, -1))
    x = F.conv2d(F.pad(x, (0, 0, r, r), mode="reflect"), k.view(1, 1, -1, 1))
    return x[0, 0]


def make_clip(seed: int, H: int = 1080, W: int = 1920, frames: int = 32, device="cpu", fg_frac=(0.25, 0.25),
              velocity=(3.0, -1.5), out=None):
    """-> uint8 tensor (frames, H, W).  Background: uniform noise -> Gaussian sigma 3 -> min-max to 0..140;
    foreground: independently seeded texture (sigma 1.5) in 150..255, fh x fw = fg_frac * (H, W) (270 x 480 at
    1080p), top-left at (x0 + vx t, y0 + vy t)."""
    import torch
    g = torch.Generator(device="cpu").manual_seed(int(seed))
    bg = torch.rand((H, W), generator=g).to(device)
    bg = _blur(bg, 3.0, torch)
    bg = (bg - bg.min()) / (bg.max() - bg.min()) * 140.0
    fh, fw = max(8, int(H * fg_frac[0])), max(8, int(W * fg_frac[1]))
    tex = torch.rand((fh, fw), generator=g).to(device)
    tex = _blur(tex, 1.5, torch)
    tex = 150.0 + (tex - tex.min()) / (tex.max() - tex.min()) * 105.0
    # start so that the object stays inside the frame for the whole clip
    vx, vy = velocity
    span_x, span_y = abs(vx) * (frames - 1), abs(vy) * (frames - 1)
    x_lo, x_hi = 2 + (span_x if vx < 0 else 0), W - fw - 3 - (span_x if vx > 0 else 0)
    y_lo, y_hi = 2 + (span_y if vy < 0 else 0), H - fh - 3 - (span_y if vy > 0 else 0)
    u = torch.rand(2, generator=g)
    x0 = x_lo + float(u[0]) * max(0.0, x_hi - x_lo)
    y0 = y_lo + float(u[1]) * max(0.0, y_hi - y_lo)
    if out is None:
        out = torch.empty((frames, H, W), dtype=torch.uint8, device=device)
    texp = torch.zeros((fh + 2, fw + 2), dtype=torch.float32, device=device)
    alp = torch.zeros_like(texp)
    texp[1:-1, 1:-1] = tex
    alp[1:-1, 1:-1] = 1.0
    for t in range(frames):
        px, py = x0 + vx * t, y0 + vy * t
        ix, iy = int(np.floor(px)), int(np.floor(py))
        fx, fy = px - ix, py - iy
        # value at canvas pixel (iy + j, ix + i) samples the padded texture at (j + 1 - fy, i + 1 - fx)
        def shift(a):
            a00, a01, a10, a11 = a[:-1, :-1], a[:-1, 1:], a[1:, :-1], a[1:, 1:]
            return a00 * (fx * fy) + a01 * ((1 - fx) * fy) + a10 * (fx * (1 - fy)) + a11 * ((1 - fx) * (1 - fy))
        ta, aa = shift(texp * alp), shift(alp)
        canvas = bg.clone()
        sub = canvas[iy:iy + fh + 1, ix:ix + fw + 1]
        sub.copy_(ta + (1 - aa) * sub)
        out[t] = canvas.clamp_(0, 255).add_(0.5).to(torch.uint8)
    return out


def synthetic_events(H=240, W=320, box=(50, 50), speed_px_s=300.0, duration_s=1.5, events_per_edge_px=1.0, seed=0):
    """Box of `box` = (bh, bw) pixels sweeping left to right at `speed_px_s`; every pixel of the leading edge emits
    an ON event (p = 1) and every pixel of the trailing edge an OFF event (p = 0) each time the edge enters a new
    column.  Returns x, y (int16), p (int8), t (int64, microseconds, sorted)."""
    rng = np.random.default_rng(seed)
    bh, bw = box
    y_top = (H - bh) // 2
    n_cols = int(speed_px_s * duration_s)
    xs, ys, ps, ts = [], [], [], []
    for c in range(n_cols):
        t_us = int(c / speed_px_s * 1e6)
        lead = c + bw
        trail = c
        for col, pol in ((lead, 1), (trail, 0)):
            if 0 <= col < W:
                yy = np.arange(y_top, y_top + bh)
                keep = rng.random(bh) < events_per_edge_px
                yy = yy[keep]
                xs.append(np.full(len(yy), col, np.int16))
                ys.append(yy.astype(np.int16))
                ps.append(np.full(len(yy), pol, np.int8))
                ts.append(t_us + rng.integers(0, max(1, int(1e6 / speed_px_s)), len(yy)))
    x, y, p, t = np.concatenate(xs), np.concatenate(ys), np.concatenate(ps), np.concatenate(ts).astype(np.int64)
    order = np.argsort(t, kind="stable")
    return x[order], y[order], p[order], t[order]
