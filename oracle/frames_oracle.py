"""frames_oracle.py -- TEST INFRASTRUCTURE ONLY (parity checker for the frame-side neighbours of the path).

numpy restatement of two OpenCV 8-bit operations the reference scripts call around the hot path.  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this module.

PINNED: tests/golden/frames_reference.npz (tests/golden/make_golden.py `make_frames`) holds outputs of
`cv2.cvtColor(., COLOR_RGB2GRAY)` on real data/grasp frames and of the reference's own `task_results`
(optical_flow_prediction.py:257-361, source-sliced) on flow fields; tests/test_oracle_golden.py checks this file
against them bit for bit.  Third-party arithmetic (opencv-python 4.9.0 pinned by the reference, 4.13.0 here), restated
from OpenCV's documented fixed-point paths and validated black-box:
* RGB2GRAY, 8-bit: (c0 * 9798 + c1 * 19235 + c2 * 3735 + 2^14) >> 15 on the channels as stored;
* remap, 8-bit, INTER_LINEAR: sx = cvRound(map_x * 32), ix = sx >> 5, ax = sx & 31 (same in y); weights
  (32 - ax)(32 - ay) * 32, ax (32 - ay) * 32, (32 - ax) ay * 32, ax ay * 32; value = (sum + 2^14) >> 15;
  BORDER_REPLICATE clamps every tap to the image.
"""
import numpy as np


def rgb2gray(frame):
    """optical_flow_seg.py:442-443: cv2.cvtColor(frame, cv2.COLOR_RGB2GRAY), uint8 (H, W, 3) -> (H, W)."""
    c = np.asarray(frame).astype(np.int64)
    return ((c[..., 0] * 9798 + c[..., 1] * 19235 + c[..., 2] * 3735 + 16384) >> 15).astype(np.uint8)


def remap_linear_replicate(src, map_x, map_y):
    """cv2.remap(src, map_x, map_y, cv2.INTER_LINEAR, borderMode=cv2.BORDER_REPLICATE) for uint8 (H, W) `src` and
    float32 maps."""
    src = np.asarray(src)
    H, W = src.shape
    with np.errstate(invalid="ignore", over="ignore"):
        def cv_round(v):
            r = np.rint(v.astype(np.float32) * np.float32(32))
            bad = ~(np.abs(r) < 2147483648.0)
            return np.where(bad, -2147483648, np.where(bad, 0, r)).astype(np.int64)
        sx, sy = cv_round(np.asarray(map_x)), cv_round(np.asarray(map_y))
    ix, iy = np.clip(sx >> 5, -32768, 32767), np.clip(sy >> 5, -32768, 32767)
    ax, ay = sx & 31, sy & 31
    px = lambda yy, xx: src[np.clip(yy, 0, H - 1), np.clip(xx, 0, W - 1)].astype(np.int64)
    v = (px(iy, ix) * ((32 - ax) * (32 - ay) * 32) + px(iy, ix + 1) * (ax * (32 - ay) * 32)
         + px(iy + 1, ix) * ((32 - ax) * ay * 32) + px(iy + 1, ix + 1) * (ax * ay * 32) + 16384) >> 15
    return v.astype(np.uint8)


def remap_flow(next_frame, flow_region, x_min, y_min):
    """optical_flow_prediction.py:281-297 for one region."""
    fl = np.asarray(flow_region)
    h, w = fl.shape[:2]
    x = np.arange(x_min, x_min + w)[None, :] + fl[..., 0].astype(np.float64)
    y = np.arange(y_min, y_min + h)[:, None] + fl[..., 1].astype(np.float64)
    mx, my = x.astype(np.float32), y.astype(np.float32)
    img = np.asarray(next_frame)
    if img.ndim == 2:
        return remap_linear_replicate(img, mx, my)
    return np.stack([remap_linear_replicate(img[:, :, c], mx, my) for c in range(img.shape[2])], axis=-1)


def task_results(prev_frame, next_frame, flow, rect):
    """optical_flow_prediction.py:257-361, EST_FLAG == 2."""
    out = np.array(next_frame, copy=True)
    x0, y0, x1, y1 = rect
    if y1 > y0 and x1 > x0:
        out[y0:y1, x0:x1] = remap_flow(next_frame, flow[y0:y1, x0:x1], x0, y0)
    return out


def ssim_u8(im1, im2, data_range=255.0):
    """skimage.metrics.structural_similarity(im1, im2, data_range=data_range) with its defaults for 2-D images
    (optical_flow_prediction.py:113-115 passes channel 2 of two uint8 frames): win_size 7, uniform filter
    (scipy.ndimage.uniform_filter, the call skimage itself makes), use_sample_covariance=True, K1 = 0.01, K2 = 0.03,
    mean over the result cropped by (win_size - 1) // 2.  scikit-image is not installed in the build container or on
    the GPU box: PARITY UNPINNED against skimage itself; the restatement follows skimage/metrics/_structural_similarity.py
    (0.2x) line by line and uses the same scipy primitive."""
    from scipy.ndimage import uniform_filter
    x, y = np.asarray(im1).astype(np.float64), np.asarray(im2).astype(np.float64)
    win, NP = 7, 49
    cov_norm = NP / (NP - 1)
    ux, uy = uniform_filter(x, size=win), uniform_filter(y, size=win)
    uxx, uyy, uxy = uniform_filter(x * x, size=win), uniform_filter(y * y, size=win), uniform_filter(x * y, size=win)
    vx, vy, vxy = cov_norm * (uxx - ux * ux), cov_norm * (uyy - uy * uy), cov_norm * (uxy - ux * uy)
    C1, C2 = (0.01 * data_range) ** 2, (0.03 * data_range) ** 2
    S = ((2 * ux * uy + C1) * (2 * vxy + C2)) / ((ux ** 2 + uy ** 2 + C1) * (vx + vy + C2))
    pad = (win - 1) // 2
    return float(S[pad:-pad, pad:-pad].mean(dtype=np.float64))
