"""seg_oracle.py -- TEST INFRASTRUCTURE ONLY (the parity checker for the flow -> motion-mask post-process).

numpy restatement of what `task_results` + `process_flow_region` return (optical_flow_seg.py:255-320, 322-357;
identical code in optical_flow_yolo.py).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this module; the product path never does.

PINNED: tests/golden/segment_reference.npz holds masks produced by the reference's own `process_flow_region`
(loaded by source slicing, tests/golden/make_golden.py `make_segment`) on `cv2.cartToPolar` of float64 flow regions,
and tests/test_oracle_golden.py checks this file against them bit for bit.

The third-party arithmetic behind the reference's calls (opencv-python 4.9.0 pinned by requirements.txt:1; 4.13.0 in
the build container) is restated from OpenCV's documented behaviour:
* `cv2.cartToPolar` converts float64 inputs to float32 and returns magnitude = sqrtf(fmaf(x, x, y*y)) widened to
  double (found by probing cv2 4.13 and pinned by the golden case `edge_values`: (0.7071068, 0.7071068) has
  magnitude exactly 1.0 there, 1.00000007 in double arithmetic);
* `cv2.getStructuringElement(MORPH_ELLIPSE, (w, h))`: row i of the element is the run [c - dx, c + dx] with
  r = h // 2, c = w // 2, dy = i - r, dx = cvRound(c * sqrt((r*r - dy*dy) / r^2)) (empty when |dy| > r);
* `cv2.dilate` / `cv2.erode` with the default anchor (element centre, ksize // 2) and default border: pixels outside
  the image do not take part in the max / min.
"""
import math

import numpy as np

SEG_TH = 1  # optical_flow_seg.py:49
MORPH_RECT, MORPH_CROSS, MORPH_ELLIPSE = 0, 1, 2


def structuring_element(shape, ksize):
    """cv2.getStructuringElement(shape, (kw, kh)) -> uint8 (kh, kw)."""
    kw, kh = int(ksize[0]), int(ksize[1])
    out = np.zeros((kh, kw), np.uint8)
    r, c = kh // 2, kw // 2
    inv_r2 = 1.0 / (r * r) if r else 0.0
    for i in range(kh):
        j1 = j2 = 0
        if shape == MORPH_RECT or (shape == MORPH_CROSS and i == r):
            j1, j2 = 0, kw
        elif shape == MORPH_CROSS:
            j1, j2 = c, c + 1
        else:
            dy = i - r
            if abs(dy) <= r:
                dx = int(np.rint(c * math.sqrt((r * r - dy * dy) * inv_r2)))  # cvRound: half to even
                j1, j2 = max(c - dx, 0), min(c + dx + 1, kw)
        out[i, j1:j2] = 1
    return out


def _morph(mask, element, erode):
    """mask: bool (h, w).  dst(y, x) = max (dilate) / min (erode) of src(y + i - ay, x + j - ax) over the set
    element pixels (i, j) that fall inside the image."""
    h, w = mask.shape
    kh, kw = element.shape
    ay, ax = kh // 2, kw // 2
    acc = np.ones((h, w), bool) if erode else np.zeros((h, w), bool)
    for i in range(kh):
        for j in range(kw):
            if not element[i, j]:
                continue
            dy, dx = i - ay, j - ax
            # destination rows / columns whose source (y + dy, x + dx) is inside the image
            y0, y1 = max(0, -dy), min(h, h - dy)
            x0, x1 = max(0, -dx), min(w, w - dx)
            if y1 <= y0 or x1 <= x0:
                continue
            src = mask[y0 + dy:y1 + dy, x0 + dx:x1 + dx]
            if erode:
                acc[y0:y1, x0:x1] &= src
            else:
                acc[y0:y1, x0:x1] |= src
    return acc


def process_flow_region(flow_region, seg_th=SEG_TH, ksize=(10, 10), iterations=5, shape=MORPH_ELLIPSE):
    """optical_flow_seg.py:308-311 + 322-357 on a flow region (h, w, 2): uint8 mask, 0 / 255.
    motion_mask = mag > SEG_TH (:345-346) with cv2.cartToPolar's float32 magnitude (the reference's flow is float64,
    :215, holding float32 values);
    5 x dilate, erode with the 10 x 10 ellipse (:349-352); threshold(., 1, 255) (:355).  The HSV / BGR / grey images
    the function also builds (:327-342) do not reach its return value."""
    f = np.asarray(flow_region).astype(np.float32)
    x, y = f[..., 0].astype(np.float64), f[..., 1]
    with np.errstate(invalid="ignore", over="ignore"):
        s = (x * x + (y * y).astype(np.float64)).astype(np.float32)  # fmaf(x, x, y*y): exact product + one rounding
        mag = np.sqrt(s).astype(np.float64)
        m = mag > seg_th
    el = structuring_element(shape, ksize)
    for _ in range(iterations):
        m = _morph(m, el, erode=False)
        m = _morph(m, el, erode=True)
    return np.where(m, 255, 0).astype(np.uint8)


def task_results(frame_shape, flow, rect, seg_th=SEG_TH):
    """optical_flow_seg.py:255-320, EST_FLAG == 2: motion_binary zeros, the merged region pasted."""
    h, w = frame_shape
    out = np.zeros((h, w), np.uint8)
    x0, y0, x1, y1 = rect
    if y1 > y0 and x1 > x0:
        out[y0:y1, x0:x1] = process_flow_region(flow[y0:y1, x0:x1], seg_th)
    return out
