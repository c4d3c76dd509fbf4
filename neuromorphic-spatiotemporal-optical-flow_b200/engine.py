"""Host side of the hot path: one `Engine` per (GPU, stream) wrapping a libnsof handle.

Mirrors the reference's operator interface for the path (same names, return shapes and error behaviour; same
argument meaning except where noted: `process_merged_region` / `process_separate_regions` take the rectangle(s) that
`opticalFlow3D` has already computed on the device instead of the reference's `(stats, ..., pixel_width, pixel_height)`
-- call `opticalFlow3D`, which has the reference's exact signature, unless you hold rectangles already):

* ``calcOpticalFlowFarneback(prev, next, flow, pyr_scale, levels, winsize, iterations, poly_n,
  poly_sigma, flags)`` -- cv2's signature as the scripts call it (optical_flow_seg.py:203).
* ``update_transition_pic``, ``opticalFlow3D``, ``process_merged_region``,
  ``process_separate_regions`` (optical_flow_seg.py:115-252).
* ``update_state``, ``bincount_2d``, ``resistance_exp``, ``slice_indices``
  (eventsim/event_mem_sim.py:40-104) and the frame-mode simulator (simulation/*.m).

numpy arrays go through the ``*_host`` C entry points (H2D, kernels, D2H, synchronous); torch CUDA
tensors go through the device entry points on the current torch stream.
"""
from __future__ import annotations

import ctypes as C
import time
from dataclasses import dataclass, field

import numpy as np

from . import _lib
from ._lib import (CropDesc, FbParams, FrameSimParams, Library, Rect, RoiParams, SegDesc, SegParams, SynapseParams,
                   WarpDesc, error)

# PARAMS of eventsim/event_mem_sim.py:20-27 == simulation/simulationcode_v4_transistor_vehicle.m:26-33
PARAMS = dict(alphaoff=1, alphaon=1, voff=-0.2, von=0.1, koff=51.03, kon=-2.91, son=0.2, soff=0.8, bon=-5.12,
              boff=3.10, Ron=163_305, Roff=2_104_377, won=1, woff=0, wini=0.5)
DT = 5e-4            # eventsim/event_mem_sim.py:30
THETA_EVENTS = 1     # eventsim/event_mem_sim.py:33
REFRACTORY_US = 800  # eventsim/event_mem_sim.py:34
SEG_TH = 1           # optical_flow_seg.py:49, optical_flow_yolo.py:80

# farneback_params of optical_flow_seg.py:73-81 (also optical_flow_yolo.py:114-122)
FARNEBACK_PARAMS = dict(pyr_scale=0.5, levels=3, winsize=15, iterations=3, poly_n=5, poly_sigma=1.2, flags=0)


@dataclass
class RoiConfig:
    """The module-level constants the scripts edit by hand (optical_flow_seg.py:24-49)."""
    MEMSIZE: int = 80
    OFFSET: int = 0
    THRES: int = 250
    EXTEND_WIDTH_LEFT: int = 20
    EXTEND_WIDTH_RIGHT: int = 20
    EXTEND_HEIGHT_UPPER: int = 20
    EXTEND_HEIGHT_LOWER: int = 20
    FLAG: int = 2          # 1 separate regions, 2 merged region
    CONNECT: int = 4       # connectedComponents connectivity (only 4 is implemented, as the scripts use)
    PADDING: int = 20      # task_results, FLAG 1 with MERGE_FLAG: margin of the box around all regions (optical_flow_seg.py:48)
    MAP_A: float = -3366.0  # u8 = uint8(clip(MAP_A / log10(I) + MAP_B, 0, 255))  (optical_flow_seg.py:426-435)
    MAP_B: float = -306.0
    farneback_params: dict = field(default_factory=lambda: dict(FARNEBACK_PARAMS))

    @classmethod
    def grasp(cls):  # optical_flow_seg.py:24-49
        return cls()

    @classmethod
    def autodriving(cls):  # optical_flow_yolo.py:68-80
        return cls(MEMSIZE=200, OFFSET=15, THRES=114, EXTEND_WIDTH_LEFT=60, EXTEND_WIDTH_RIGHT=60,
                   EXTEND_HEIGHT_UPPER=60, EXTEND_HEIGHT_LOWER=60, PADDING=60)

    # The three datasets below have no checked-in script of their own; their constants are data/<ds>/Parameters.txt.
    @classmethod
    def uav(cls):  # data/uav/Parameters.txt
        return cls(MEMSIZE=40, THRES=114, FLAG=1, EXTEND_WIDTH_LEFT=30, EXTEND_WIDTH_RIGHT=30, EXTEND_HEIGHT_UPPER=30,
                   EXTEND_HEIGHT_LOWER=30, PADDING=30, farneback_params=dict(pyr_scale=0.6, levels=3, winsize=3, iterations=3,
                                                                 poly_n=10, poly_sigma=1.05, flags=0))

    @classmethod
    def uavnew2(cls):  # data/uavnew2/Parameters.txt
        return cls(MEMSIZE=40, THRES=245, FLAG=1, EXTEND_WIDTH_LEFT=60, EXTEND_WIDTH_RIGHT=60, EXTEND_HEIGHT_UPPER=60,
                   EXTEND_HEIGHT_LOWER=60, PADDING=60)

    @classmethod
    def tabletennis(cls):  # data/tabletennis/Parameters.txt
        return cls(MEMSIZE=10, THRES=245, FLAG=2, EXTEND_WIDTH_LEFT=20, EXTEND_WIDTH_RIGHT=20, EXTEND_HEIGHT_UPPER=20,
                   EXTEND_HEIGHT_LOWER=20, farneback_params=dict(pyr_scale=0.6, levels=3, winsize=4, iterations=2,
                                                                 poly_n=1, poly_sigma=1.05, flags=0))


def _fb_params(pyr_scale, levels, winsize, iterations, poly_n, poly_sigma, flags) -> FbParams:
    return FbParams(float(pyr_scale), int(levels), int(winsize), int(iterations), int(poly_n), float(poly_sigma),
                    int(flags))


def _synapse_params(p) -> SynapseParams:
    return SynapseParams(*[float(p[k]) for k, _ in SynapseParams._fields_])


def current_to_u8_reference_formula(I, a=-3366.0, b=-306.0):
    """The reference's host formula, verbatim semantics (optical_flow_seg.py:426-435). Used only to build
    the monotone threshold table; the map itself is evaluated on the device."""
    with np.errstate(divide="ignore", invalid="ignore"):
        v = a / np.log10(np.asarray(I, dtype=np.float64)) + b
    return np.clip(v, 0, 255).astype(np.uint8)


_THR_CACHE: dict = {}


def u8_threshold_table(a=-3366.0, b=-306.0) -> np.ndarray:
    """thr[v] = smallest float64 current in (0,1) whose uint8 map value is >= v (v = 1..255), found by
    bisection over the float64 bit pattern with numpy's own log10, so the device predicate `I >= thr[v]`
    reproduces numpy's map bit for bit (the map is monotone increasing on (0,1) for a < 0)."""
    key = (float(a), float(b))
    if key in _THR_CACHE:
        return _THR_CACHE[key]
    if not a < 0:
        raise NotImplementedError("current->uint8 map must be increasing on (0,1): MAP_A < 0")
    thr = np.full(256, np.inf, dtype=np.float64)
    thr[0] = 0.0
    lo_bits = np.float64(np.finfo(np.float64).tiny).view(np.uint64)
    hi_bits = np.nextafter(np.float64(1.0), np.float64(0.0)).view(np.uint64)
    f = lambda bits: int(current_to_u8_reference_formula(np.array([bits], dtype=np.uint64).view(np.float64), a, b)[0])
    for v in range(1, 256):
        lo, hi = int(lo_bits), int(hi_bits)
        if f(hi) < v:
            continue  # never reached: stays +inf
        if f(lo) >= v:
            thr[v] = np.array([lo], dtype=np.uint64).view(np.float64)[0]
            continue
        while hi - lo > 1:  # invariant: f(lo) < v <= f(hi); positive doubles order like their bit patterns
            mid = (lo + hi) // 2
            if f(mid) >= v:
                hi = mid
            else:
                lo = mid
        thr[v] = np.array([hi], dtype=np.uint64).view(np.float64)[0]
    _THR_CACHE[key] = thr
    return thr


class Engine:
    """One libnsof handle: owns the device workspace; calls are serialised by the caller."""

    def __init__(self, device: int = 0, lib: Library | None = None, _emulated: bool = False):
        self.lib = lib or _lib.default_library()
        self.device = int(device)
        # tests/emu only: the block-emulator build treats host memory as device memory
        self._emulated = bool(_emulated)
        self.handle = self.lib.create_handle(self.device)
        self.config = RoiConfig()
        # timing lists, like the module globals of optical_flow_seg.py:51-59
        self.mem_cal_times: list = []
        self.mem_velocity_times: list = []
        self.mem_opticalflow_times: list = []

    def close(self):
        if getattr(self, "handle", None):
            self.lib.nsof_destroy(self.handle)
            self.handle = None

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ stage 2
    def calcOpticalFlowFarneback(self, prev, next, flow, pyr_scale, levels, winsize, iterations, poly_n, poly_sigma,
                                 flags):
        """Drop-in for cv2.calcOpticalFlowFarneback on numpy arrays (uint8 single channel, any row stride)."""
        prev, next = self._check_pair(prev, next)
        h, w = prev.shape
        if flow is None:
            if int(flags) & 4:  # cv2: assertion on the size / type of the initial flow
                raise error("(-215:Assertion failed) OPTFLOW_USE_INITIAL_FLOW needs a float32 (h, w, 2) flow")
            out = np.empty((h, w, 2), dtype=np.float32)
        else:
            out = flow
            if not (isinstance(out, np.ndarray) and out.dtype == np.float32 and out.shape == (h, w, 2)
                    and out.strides[1:] == (8, 4)):
                raise error("flow must be None or a float32 (h, w, 2) array")
        if int(flags) & ~(4 | 256):
            raise NotImplementedError(f"flags={flags}: cv2 defines OPTFLOW_USE_INITIAL_FLOW (4) and OPTFLOW_FARNEBACK_GAUSSIAN (256)")
        p = _fb_params(pyr_scale, levels, winsize, iterations, poly_n, poly_sigma,
                       int(flags) | (_lib.NSOF_FB_SRC_F32 if prev.dtype == np.float32 else 0))
        rc = self.lib.nsof_farneback_u8_host(self.handle, prev.ctypes.data, next.ctypes.data, h, w, prev.strides[0],
                                             next.strides[0], C.byref(p), out.ctypes.data, out.strides[0])
        self.lib.check(rc, "calcOpticalFlowFarneback")
        return out

    @staticmethod
    def _check_pair(prev, next):
        prev, next = np.asarray(prev), np.asarray(next)
        if prev.size == 0 or next.size == 0:
            raise error("(-215:Assertion failed) !_src.empty()")
        if prev.ndim == 3 and prev.shape[2] == 1:
            prev = prev[:, :, 0]
        if next.ndim == 3 and next.shape[2] == 1:
            next = next[:, :, 0]
        if prev.ndim != 2 or next.ndim != 2 or prev.shape != next.shape:
            raise error("(-215:Assertion failed) prev0.size() == next0.size() && prev0.channels() == "
                        "next0.channels() && prev0.channels() == 1")
        if prev.dtype == np.float32 or next.dtype == np.float32:  # cv2 takes CV_32F frames too (no script passes them)
            prev, next = np.asarray(prev, dtype=np.float32), np.asarray(next, dtype=np.float32)
        elif prev.dtype != np.uint8 or next.dtype != np.uint8:
            raise error("(-210:Unsupported format or combination of formats) frames must be uint8 or float32")
        esz = prev.dtype.itemsize
        if prev.strides[1] != esz or prev.strides[0] < prev.shape[1] * esz or prev.strides[0] % esz:
            prev = np.ascontiguousarray(prev)
        if next.strides[1] != esz or next.strides[0] < next.shape[1] * esz or next.strides[0] % esz:
            next = np.ascontiguousarray(next)
        return prev, next

    def farneback_batch(self, pairs, **farneback_params):
        """`pairs`: list of (prev, next) numpy crops of any sizes -> list of float32 (h, w, 2) flows, one launch
        sequence for the whole batch."""
        kw = dict(FARNEBACK_PARAMS)
        kw.update(farneback_params)
        p = _fb_params(**kw)
        n = len(pairs)
        descs = (CropDesc * max(n, 1))()
        keep, outs = [], []
        for i, (a, b) in enumerate(pairs):
            a, b = self._check_pair(a, b)
            if a.dtype != np.uint8:
                raise NotImplementedError("farneback_batch: uint8 crops (float32 frames go through calcOpticalFlowFarneback)")
            out = np.empty(a.shape + (2,), dtype=np.float32)
            keep.append((a, b))
            outs.append(out)
            descs[i] = CropDesc(a.ctypes.data, b.ctypes.data, a.shape[0], a.shape[1], a.strides[0], b.strides[0],
                                out.ctypes.data, out.strides[0])
        self.lib.check(self.lib.nsof_farneback_batch_host(self.handle, descs, n, C.byref(p)), "farneback_batch")
        return outs

    def farneback_batch_device(self, crops, stream_ptr: int = 0, **farneback_params):
        """Device path.  `crops`: list of dicts/tuples (prev_ptr, next_ptr, h, w, prev_pitch, next_pitch, flow_ptr,
        flow_pitch) with raw device addresses (e.g. torch `data_ptr()` plus offsets).  Asynchronous."""
        kw = dict(FARNEBACK_PARAMS)
        kw.update(farneback_params)
        p = _fb_params(**kw)
        n = len(crops)
        descs = (CropDesc * max(n, 1))()
        for i, c in enumerate(crops):
            descs[i] = CropDesc(*c)
        self.lib.check(self.lib.nsof_farneback_batch(self.handle, descs, n, C.byref(p), C.c_void_p(stream_ptr)),
                       "farneback_batch_device")

    def farneback_descs(self, crops):
        """Pre-built ctypes descriptor array for repeated device batches (bench inner loop)."""
        descs = (CropDesc * max(len(crops), 1))()
        for i, c in enumerate(crops):
            descs[i] = CropDesc(*c)
        return descs

    def farneback_batch_descs(self, descs, n, fb: FbParams, stream_ptr: int = 0):
        self.lib.check(self.lib.nsof_farneback_batch(self.handle, descs, n, C.byref(fb), C.c_void_p(stream_ptr)),
                       "farneback_batch")

    def reserve(self, total_crop_pixels: int, max_items: int):
        self.lib.check(self.lib.nsof_reserve(self.handle, int(total_crop_pixels), int(max_items)), "nsof_reserve")

    def level_geometry(self, h, w, pyr_scale, levels):
        hk = (C.c_int * 64)()
        wk = (C.c_int * 64)()
        ks = (C.c_int * 64)()
        sg = (C.c_double * 64)()
        n = self.lib.nsof_fb_level_geometry(h, w, pyr_scale, levels, hk, wk, ks, sg)
        return [(hk[i], wk[i], ks[i], sg[i]) for i in range(n)]

    def algorithmic_bytes(self, h, w, **farneback_params) -> float:
        kw = dict(FARNEBACK_PARAMS)
        kw.update(farneback_params)
        p = _fb_params(**kw)
        return float(self.lib.nsof_fb_algorithmic_bytes(int(h), int(w), C.byref(p)))

    # launch accounting (bench.py)
    def launch_count(self) -> int:
        return int(self.lib.nsof_launch_count(self.handle))

    def profile(self, enable: bool):
        self.lib.check(self.lib.nsof_profile(self.handle, 1 if enable else 0), "nsof_profile")

    def profile_read(self):
        """-> {kernel class: dict(launches, ms, algorithmic_bytes)} for the launches since profile(True)."""
        rows = (_lib.KernelStat * 128)()
        n = C.c_int()
        self.lib.check(self.lib.nsof_profile_read(self.handle, rows, 128, C.byref(n)), "nsof_profile_read")
        return {rows[i].name.decode(): dict(launches=rows[i].launches, ms=rows[i].ms,
                                            algorithmic_bytes=rows[i].algorithmic_bytes) for i in range(n.value)}

    # test hooks
    def debug_truncate(self, stop_level=-1, stop_iters=-1):
        self.lib.check(self.lib.nsof_debug_truncate(self.handle, stop_level, stop_iters))

    def debug_fuse(self, enable: bool):
        self.lib.check(self.lib.nsof_debug_fuse(self.handle, 1 if enable else 0))

    def debug_fetch(self, item, what, h, w):
        ch = {0: 1, 1: 1, 2: 5, 3: 5, 4: 5, 5: 2}[what]
        out = np.empty((ch, h, w) if what != 5 else (h, w, 2), dtype=np.float32)
        oh, ow = C.c_int(), C.c_int()
        self.lib.check(self.lib.nsof_debug_fetch(self.handle, item, what, out.ctypes.data, C.byref(oh), C.byref(ow)))
        assert (oh.value, ow.value) == (h, w), ((oh.value, ow.value), (h, w))
        return out

    # ------------------------------------------------------------------ stage 1: ROI tail
    def _roi_params(self, cells_shape, frame_shape, pixel_width, pixel_height, cfg: RoiConfig) -> RoiParams:
        if cfg.CONNECT != 4:
            raise NotImplementedError("only connectivity=4 is implemented (CONNECT in every reference script)")
        return RoiParams(int(cells_shape[0]), int(cells_shape[1]), int(frame_shape[0]), int(frame_shape[1]),
                         int(pixel_height), int(pixel_width), cfg.EXTEND_WIDTH_LEFT, cfg.EXTEND_WIDTH_RIGHT,
                         cfg.EXTEND_HEIGHT_UPPER, cfg.EXTEND_HEIGHT_LOWER, cfg.FLAG, cfg.THRES, cfg.MAP_A, cfg.MAP_B)

    def roi_from_currents(self, currents, frame_shape, pixel_width=None, pixel_height=None, cfg: RoiConfig | None = None):
        """Stage-1 tail for `n` array states at once.  currents: float64 (n, Hc, Wc) or (Hc, Wc) [A].
        Returns dict(memimg uint8, labels int32, num_labels, rects list per state, stats)."""
        cfg = cfg or self.config
        cur = np.ascontiguousarray(np.asarray(currents, dtype=np.float64))
        single = cur.ndim == 2
        if single:
            cur = cur[None]
        thr = u8_threshold_table(cfg.MAP_A, cfg.MAP_B)
        return self._roi(cur, thr, frame_shape, pixel_width or cfg.MEMSIZE, pixel_height or cfg.MEMSIZE, cfg, single)

    def roi_from_memimg(self, memimg, frame_shape, pixel_width=None, pixel_height=None, cfg: RoiConfig | None = None):
        """Same, starting from the uint8 map (what opticalFlow3D receives).  The map is re-encoded as the
        pseudo-current (u8 + 0.5)/256 with the identity table thr[v] = v/256, so the same kernel runs."""
        cfg = cfg or self.config
        m = np.asarray(memimg)
        single = m.ndim == 2
        if single:
            m = m[None]
        cur = (m.astype(np.float64) + 0.5) / 256.0
        thr = np.arange(256, dtype=np.float64) / 256.0
        return self._roi(np.ascontiguousarray(cur), thr, frame_shape, pixel_width or cfg.MEMSIZE,
                         pixel_height or cfg.MEMSIZE, cfg, single)

    def _roi(self, cur, thr, frame_shape, pixel_width, pixel_height, cfg, single):
        n, Hc, Wc = cur.shape
        H, W = int(frame_shape[0]), int(frame_shape[1])
        th, tw = int(H / pixel_height), int(W / pixel_width)  # transition_pic shape, optical_flow_seg.py:218
        if Hc > th or Wc > tw:
            raise error(f"array state {Hc}x{Wc} exceeds the transition picture {th}x{tw} "
                        f"(the reference's numba loop would write out of bounds)")
        if (Hc, Wc) != (th, tw):  # smaller state: the rest of transition_pic stays 0
            pad = np.zeros((n, th, tw), dtype=np.float64)
            pad[:, :Hc, :Wc] = cur
            cur, Hc, Wc = pad, th, tw
        rp = self._roi_params((Hc, Wc), (H, W), pixel_width, pixel_height, cfg)
        memimg = np.empty((n, Hc, Wc), np.uint8)
        labels = np.empty((n, Hc, Wc), np.int32)
        ncomp = np.empty(n, np.int32)
        nrect = np.empty(n, np.int32)
        rects = np.empty((n, _lib.NSOF_MAX_RECTS, 4), np.int32)
        stats = np.empty((n, _lib.NSOF_MAX_RECTS, 5), np.int32)
        thr = np.ascontiguousarray(thr, dtype=np.float64)
        rc = self.lib.nsof_roi_from_state_f64_host(self.handle, cur.ctypes.data, n, C.byref(rp), thr.ctypes.data,
                                                   memimg.ctypes.data, labels.ctypes.data, ncomp.ctypes.data,
                                                   nrect.ctypes.data, rects.ctypes.data, stats.ctypes.data)
        self.lib.check(rc, "roi_from_state")
        if (nrect > _lib.NSOF_MAX_RECTS).any():
            raise NotImplementedError(f"more than {_lib.NSOF_MAX_RECTS} separate regions in one array state")
        out = dict(memimg=memimg, labels=labels, num_labels=ncomp + 1,
                   rects=[[tuple(int(v) for v in rects[i, j]) for j in range(nrect[i])] for i in range(n)],
                   stats=[stats[i, :min(int(ncomp[i]), _lib.NSOF_MAX_RECTS)].copy() for i in range(n)])
        if single:
            out = {k: v[0] for k, v in out.items()}
        return out

    def current_to_u8(self, currents, cfg: RoiConfig | None = None):
        """current -> uint8 map on the device (optical_flow_seg.py:426-435 / process_mem_state)."""
        cfg = cfg or self.config
        cur = np.asarray(currents, dtype=np.float64)
        shape = cur.shape
        Hc, Wc = shape[-2], shape[-1]
        r = self.roi_from_currents(cur.reshape(-1, Hc, Wc), (Hc * cfg.MEMSIZE, Wc * cfg.MEMSIZE), cfg=cfg)
        return r["memimg"].reshape(shape)

    def update_transition_pic(self, prev_memristor, transition_pic, thres):
        """optical_flow_seg.py:115-121: transition_pic[i, j] = 255 where prev_memristor[i, j] >= thres (in place)."""
        m = np.asarray(prev_memristor)
        cfg = RoiConfig(THRES=int(thres), MEMSIZE=1, EXTEND_WIDTH_LEFT=0, EXTEND_WIDTH_RIGHT=0, EXTEND_HEIGHT_UPPER=0,
                        EXTEND_HEIGHT_LOWER=0)
        r = self.roi_from_memimg(m, m.shape, 1, 1, cfg)
        transition_pic[:m.shape[0], :m.shape[1]][r["labels"] > 0] = 255
        return transition_pic

    def process_merged_region(self, rect, rgbimg1, rgbimg2, flow, farneback_params=None):
        """FLAG=2 (optical_flow_seg.py:168-209) for an already computed merged rectangle."""
        fp = farneback_params or self.config.farneback_params
        x0, y0, x1, y1 = rect
        region_list = []
        t0 = time.time()
        prev_region, next_region = rgbimg1[y0:y1, x0:x1], rgbimg2[y0:y1, x0:x1]
        self.mem_cal_times.append(time.time() - t0)
        region_list.append(prev_region.shape[0] * prev_region.shape[1] / (rgbimg1.shape[0] * rgbimg1.shape[1]) * 100)
        if prev_region.size > 0 and next_region.size > 0:
            t0 = time.time()
            flow[y0:y1, x0:x1] = self.calcOpticalFlowFarneback(prev_region, next_region, None, **fp)
            self.mem_velocity_times.append(time.time() - t0)
        return flow, self.mem_cal_times, self.mem_velocity_times, region_list, (x0, y0, x1, y1)

    # ------------------------------------------------------------------ flow -> motion mask
    def segment_flow(self, flow, seg_th=SEG_TH, ksize=(10, 10), iterations=5, shape=2):
        """cv2.cartToPolar + process_flow_region (optical_flow_seg.py:308-311, 322-357) on a numpy flow region
        (h, w, 2), float32 or the reference's float64 holding float32 values -> uint8 (h, w) mask, 0 / 255."""
        f = np.asarray(flow)
        if f.ndim != 3 or f.shape[2] != 2 or f.shape[0] < 1 or f.shape[1] < 1:
            raise error("flow region must be a non-empty (h, w, 2) array")
        f32 = np.ascontiguousarray(f, dtype=np.float32)
        if f.dtype != np.float32 and not np.array_equal(f32.astype(f.dtype), f, equal_nan=True):
            raise error("flow region holds values that are not float32-representable")
        h, w = f32.shape[:2]
        out = np.empty((h, w), dtype=np.uint8)
        p = SegParams(float(seg_th), int(shape), int(ksize[0]), int(ksize[1]), int(iterations))
        self.lib.check(self.lib.nsof_flow_segment_host(self.handle, f32.ctypes.data, f32.strides[0], h, w, C.byref(p),
                                                       out.ctypes.data, out.strides[0]), "segment_flow")
        return out

    def segment_flow_device(self, regions, stream_ptr: int = 0, seg_th=SEG_TH, ksize=(10, 10), iterations=5, shape=2):
        """`regions`: list of (flow_ptr, flow_pitch_bytes, h, w, mask_ptr, mask_pitch_bytes) with raw device
        addresses; all regions in 2 + 2 * iterations launches, asynchronous on the stream."""
        n = len(regions)
        descs = (SegDesc * max(n, 1))()
        for i, r in enumerate(regions):
            descs[i] = SegDesc(*r)
        p = SegParams(float(seg_th), int(shape), int(ksize[0]), int(ksize[1]), int(iterations))
        self.lib.check(self.lib.nsof_flow_segment_batch(self.handle, descs, n, C.byref(p), C.c_void_p(stream_ptr)),
                       "segment_flow_device")

    def task_results(self, prev_frame, next_frame, flow, num_labels, regions_info, EST_FLAG=2, MERGE_FLAG=True,
                     padding=0, seg_th=SEG_TH):
        """optical_flow_seg.py:255-320: motion_binary of one frame pair from the pasted flow and the region(s)."""
        h, w = prev_frame.shape[:2]
        motion_binary = np.zeros((h, w), dtype=np.uint8)
        if num_labels > 1:
            if EST_FLAG == 1 and MERGE_FLAG:
                boxes = [(max(0, min(r[0] for r in regions_info) - padding), max(0, min(r[1] for r in regions_info) - padding),
                          min(w, max(r[2] for r in regions_info) + padding), min(h, max(r[3] for r in regions_info) + padding))]
            elif EST_FLAG == 1:
                boxes = [tuple(r) for r in regions_info]
            else:
                boxes = [tuple(regions_info)]
            for x0, y0, x1, y1 in boxes:
                if y1 > y0 and x1 > x0:
                    motion_binary[y0:y1, x0:x1] = self.segment_flow(flow[y0:y1, x0:x1], seg_th)
        return motion_binary

    # ------------------------------------------------------------------ frames either side of the path (8(f) 2, 3)
    def cvtColor_rgb2gray(self, frame):
        """cv2.cvtColor(frame, cv2.COLOR_RGB2GRAY) for a uint8 (H, W, 3) array, channels as stored
        (optical_flow_seg.py:442-443 passes the BGR frame of cv2.imread)."""
        f = np.asarray(frame)
        if f.dtype != np.uint8 or f.ndim != 3 or f.shape[2] != 3 or f.shape[0] < 1 or f.shape[1] < 1:
            raise error("frame must be a non-empty uint8 (H, W, 3) array")
        if f.strides[1:] != (3, 1):
            f = np.ascontiguousarray(f)
        H, W = f.shape[:2]
        out = np.empty((H, W), dtype=np.uint8)
        self.lib.check(self.lib.nsof_rgb2gray_u8_host(self.handle, f.ctypes.data, f.strides[0], H, W, out.ctypes.data,
                                                      out.strides[0]), "cvtColor_rgb2gray")
        return out

    def imread_gray_device(self, jpeg_blobs, out=None, stream_ptr: int = 0):
        """`cv2.cvtColor(cv2.imread(path), cv2.COLOR_RGB2GRAY)` (optical_flow_seg.py:439-443) for a list of JPEG byte
        strings of one size -> uint8 torch tensor (n, H, W) on the device (`out`: write into this tensor instead).
        Decoder = nvJPEG, not cv2's libjpeg-turbo: grey levels agree to a few units, not bit for bit."""
        import torch
        blobs = [bytes(b) for b in jpeg_blobs]
        n = len(blobs)
        if n == 0:
            raise error("(-215:Assertion failed) !_src.empty()")
        bufs = [np.frombuffer(b, dtype=np.uint8) for b in blobs]
        hh, ww = C.c_int(), C.c_int()
        self.lib.check(self.lib.nsof_jpeg_info(self.handle, bufs[0].ctypes.data, bufs[0].size, C.byref(hh), C.byref(ww)), "jpeg_info")
        H, W = hh.value, ww.value
        if out is None:
            out = torch.empty((n, H, W), dtype=torch.uint8, device=f"cuda:{self.device}")
        assert out.dtype == torch.uint8 and tuple(out.shape) == (n, H, W) and out.stride(2) == 1
        ptrs = (C.c_void_p * n)(*[b.ctypes.data for b in bufs])
        lens = (C.c_size_t * n)(*[b.size for b in bufs])
        self.lib.check(self.lib.nsof_jpeg_decode_gray(self.handle, ptrs, lens, n, H, W, out.data_ptr(), out.stride(1), out.stride(0),
                                                      C.c_void_p(stream_ptr or torch.cuda.current_stream(self.device).cuda_stream)),
                       "jpeg_decode_gray")
        return out

    def remap_flow(self, next_frame, flow_region, x_min, y_min, flow_sign=1):
        """optical_flow_prediction.py:281-297 for one region: flow_map = (x, y) + flow_region as float32, every channel
        of `next_frame` (uint8 (H, W) or (H, W, C)) remapped with INTER_LINEAR / BORDER_REPLICATE -> uint8 region."""
        img = np.asarray(next_frame)
        if img.dtype != np.uint8 or img.ndim not in (2, 3) or img.size == 0:
            raise error("next_frame must be a non-empty uint8 (H, W[, C]) array")
        img3 = np.ascontiguousarray(img.reshape(img.shape[0], img.shape[1], -1))
        fl = np.asarray(flow_region)
        if fl.ndim != 3 or fl.shape[2] != 2 or fl.shape[0] < 1 or fl.shape[1] < 1:
            raise error("flow region must be a non-empty (h, w, 2) array")
        f32 = np.ascontiguousarray(fl, dtype=np.float32)
        if fl.dtype != np.float32 and not np.array_equal(f32.astype(fl.dtype), fl, equal_nan=True):
            raise error("flow region holds values that are not float32-representable")
        H, W, ch = img3.shape
        h, w = f32.shape[:2]
        out = np.empty((h, w, ch), dtype=np.uint8)
        self.lib.check(self.lib.nsof_remap_flow_host(self.handle, img3.ctypes.data, img3.strides[0], H, W, ch, f32.ctypes.data,
                                                     f32.strides[0], int(x_min), int(y_min), h, w, int(flow_sign),
                                                     out.ctypes.data, out.strides[0]), "remap_flow")
        return out if img.ndim == 3 else out[..., 0]

    def remap_flow_device(self, regions, stream_ptr: int = 0, flow_sign=1):
        """`regions`: list of WarpDesc field tuples (frame_ptr, frame_pitch, H, W, channels, flow_ptr, flow_pitch, x0, y0,
        h, w, out_ptr, out_pitch) with raw device addresses; one launch, asynchronous on the stream."""
        n = len(regions)
        descs = (WarpDesc * max(n, 1))()
        for i, r in enumerate(regions):
            descs[i] = WarpDesc(*r)
        self.lib.check(self.lib.nsof_remap_flow_batch(self.handle, descs, n, int(flow_sign), C.c_void_p(stream_ptr)),
                       "remap_flow_device")

    def predict_frame(self, prev_frame, next_frame, flow, num_labels, regions_info, EST_FLAG=2, MERGE_FLAG=True, padding=0):
        """task_results of optical_flow_prediction.py:257-361: next_frame with every region replaced by its
        flow-extrapolated version."""
        h, w = prev_frame.shape[:2]
        new_frame = np.array(next_frame, copy=True)
        if num_labels > 1:
            if EST_FLAG == 1 and MERGE_FLAG:
                boxes = [(max(0, min(r[0] for r in regions_info) - padding), max(0, min(r[1] for r in regions_info) - padding),
                          min(w, max(r[2] for r in regions_info) + padding), min(h, max(r[3] for r in regions_info) + padding))]
            elif EST_FLAG == 1:
                boxes = [tuple(r) for r in regions_info]
            else:
                boxes = [tuple(regions_info)]
            for x0, y0, x1, y1 in boxes:
                if y1 > y0 and x1 > x0:
                    new_frame[y0:y1, x0:x1] = self.remap_flow(next_frame, flow[y0:y1, x0:x1], x0, y0)
        return new_frame

    def structural_similarity(self, im1, im2, data_range=255.0):
        """skimage.metrics.structural_similarity(im1, im2, data_range=...) for 2-D uint8 arrays with skimage's defaults,
        as calculateIntegralError calls it on channel 2 of the true and the predicted frame
        (optical_flow_prediction.py:113-115).  Channel views of interleaved frames (`frame[:, :, 2]`) are read in place."""
        a, b = np.asarray(im1), np.asarray(im2)
        if a.shape != b.shape:
            raise ValueError("Input images must have the same dimensions.")      # skimage's check_shape_equality
        if a.ndim != 2 or a.dtype != np.uint8 or b.dtype != np.uint8:
            raise NotImplementedError("structural_similarity: 2-D uint8 images (one channel), as the prediction script passes")
        if min(a.shape) < 7:
            raise ValueError("win_size exceeds image extent.")
        fix = lambda v: v if (v.strides[1] >= 1 and v.strides[0] >= v.shape[1] * v.strides[1]) else np.ascontiguousarray(v)
        a, b = fix(a), fix(b)
        out = C.c_double()
        self.lib.check(self.lib.nsof_ssim_u8_host(self.handle, a.ctypes.data, a.strides[0], a.strides[1], b.ctypes.data,
                                                  b.strides[0], b.strides[1], a.shape[0], a.shape[1], float(data_range),
                                                  C.byref(out)), "structural_similarity")
        return float(out.value)

    def calculateIntegralError(self, prediction, true):
        """optical_flow_prediction.py:113-115."""
        return self.structural_similarity(true[:, :, 2], prediction[:, :, 2], data_range=255.0)

    def structuring_element(self, shape, ksize):
        """cv2.getStructuringElement(shape, ksize) as the segmentation passes build it."""
        out = np.zeros((int(ksize[1]), int(ksize[0])), dtype=np.uint8)
        self.lib.check(self.lib.nsof_structuring_element(int(shape), int(ksize[0]), int(ksize[1]), out.ctypes.data),
                       "structuring_element")
        return out

    def process_separate_regions(self, rects, rgbimg1, rgbimg2, flow, farneback_params=None):
        """FLAG=1 (optical_flow_seg.py:123-166): every region's crop pair goes into ONE batched launch sequence."""
        fp = farneback_params or self.config.farneback_params
        region_list, regions_info, pairs, where = [], [], [], []
        total = rgbimg1.shape[0] * rgbimg1.shape[1]
        for (x0, y0, x1, y1) in rects:
            t0 = time.time()
            regions_info.append((x0, y0, x1, y1))
            a, b = rgbimg1[y0:y1, x0:x1], rgbimg2[y0:y1, x0:x1]
            self.mem_cal_times.append(time.time() - t0)
            region_list.append(a.shape[0] * a.shape[1] / total * 100)
            if a.size > 0 and b.size > 0:
                pairs.append((a, b))
                where.append((x0, y0, x1, y1))
            else:
                self.mem_velocity_times.append(0)
        if pairs:
            t0 = time.time()
            flows = self.farneback_batch(pairs, **fp)
            for (x0, y0, x1, y1), f in zip(where, flows):  # later regions overwrite earlier ones, as in the loop
                flow[y0:y1, x0:x1] = f
            dt = time.time() - t0
            self.mem_velocity_times.extend([dt / len(pairs)] * len(pairs))
        return flow, self.mem_cal_times, self.mem_velocity_times, region_list, len(rects) + 1, regions_info

    def opticalFlow3D(self, memimg1, memimg2, rgbimg1, rgbimg2, pixel_width, pixel_height, cfg: RoiConfig | None = None):
        """optical_flow_seg.py:211-252.  Returns what the reference returns: FLAG==2 ->
        (flow float64 HxWx2, cal_times, vel_times, region_list, regions_info); FLAG==1 adds num_labels
        before regions_info.  memimg1 is unused by the reference too."""
        cfg = cfg or self.config
        t_start = time.time()
        h, w = rgbimg1.shape[:2]
        flow = np.zeros((h, w, 2))
        roi = self.roi_from_memimg(memimg2, (h, w), pixel_width, pixel_height, cfg)
        num_labels = int(roi["num_labels"])
        if num_labels == 1:
            dt = time.time() - t_start
            self.mem_cal_times.append(dt)
            self.mem_opticalflow_times.append(dt)
            if cfg.FLAG == 1:
                return flow, self.mem_cal_times, self.mem_opticalflow_times, [], num_labels, []
            return flow, self.mem_cal_times, self.mem_opticalflow_times, [], (0, 0, 0, 0)
        if cfg.FLAG == 1:
            res = self.process_separate_regions(roi["rects"], rgbimg1, rgbimg2, flow, cfg.farneback_params)
        else:
            res = self.process_merged_region(roi["rects"][0], rgbimg1, rgbimg2, flow, cfg.farneback_params)
        self.mem_opticalflow_times.append(time.time() - t_start)
        return res

    # ------------------------------------------------------------------ stage 1: simulators (host arrays)
    def _dev_roundtrip(self, arrays):
        import torch
        dev = "cpu" if self._emulated else f"cuda:{self.device}"
        return [torch.from_numpy(np.ascontiguousarray(a)).clone().to(dev) for a in arrays]

    def _stream(self):
        if self._emulated:
            return C.c_void_p(0)
        import torch
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def update_state(self, w, V, p=PARAMS, dt=DT):
        """eventsim/event_mem_sim.py:40-57 on float32 arrays (returns a new array)."""
        import torch
        w32 = np.ascontiguousarray(w, dtype=np.float32)
        V32 = np.ascontiguousarray(np.broadcast_to(np.asarray(V, dtype=np.float32), w32.shape))
        dw, dV = self._dev_roundtrip([w32, V32])
        sp = _synapse_params(p)
        self.lib.check(self.lib.nsof_synapse_step_f32(self.handle, dw.data_ptr(), dV.data_ptr(), dw.numel(), C.byref(sp),
                                                      float(dt), self._stream()))
        return dw.cpu().numpy()

    def bincount_2d(self, x, y, H, W):
        """eventsim/event_mem_sim.py:100-104 -> int32 (H, W)."""
        import torch
        x16 = np.ascontiguousarray(x, dtype=np.int16)
        y16 = np.ascontiguousarray(y, dtype=np.int16)
        if x16.size and (x16.min() < 0 or y16.min() < 0 or x16.max() >= W or y16.max() >= H):
            raise ValueError("event coordinates outside the H x W array (np.bincount would mis-bin or raise)")
        dx, dy = self._dev_roundtrip([x16, y16])
        out = torch.empty((H, W), dtype=torch.int32, device=dx.device)
        self.lib.check(self.lib.nsof_event_hist(self.handle, dx.data_ptr(), dy.data_ptr(), x16.size, H, W, out.data_ptr(),
                                                self._stream()))
        return out.cpu().numpy()

    def resistance_exp(self, w, p=PARAMS):
        """eventsim/event_mem_sim.py:60-63 for float32 w -> float64."""
        import torch
        w32 = np.ascontiguousarray(w, dtype=np.float32)
        (dw,) = self._dev_roundtrip([w32])
        out = torch.empty(w32.shape, dtype=torch.float64, device=dw.device)
        sp = _synapse_params(p)
        self.lib.check(self.lib.nsof_resistance_exp_f32(self.handle, dw.data_ptr(), dw.numel(), C.byref(sp), out.data_ptr(),
                                                        self._stream()))
        return out.cpu().numpy()

    def simulate_scheme1(self, x, y, t, H, W, slice_us, active_v=-6.0, silent_v=0.0, p=PARAMS, dt=DT, w0=None,
                         theta=THETA_EVENTS):
        """Scheme 1 of simulate() (eventsim/event_mem_sim.py:204-228) over the whole stream; returns w_final."""
        import torch
        t = np.asarray(t)
        bounds = slice_bounds(t, slice_us)
        x16 = np.ascontiguousarray(x, dtype=np.int16)
        y16 = np.ascontiguousarray(y, dtype=np.int16)
        dx, dy, db = self._dev_roundtrip([x16, y16, bounds.astype(np.int64)])
        w = np.full((H, W), p["wini"], dtype=np.float32) if w0 is None else np.ascontiguousarray(w0, dtype=np.float32)
        (dw,) = self._dev_roundtrip([w])
        sp = _synapse_params(p)
        rc = self.lib.nsof_event_scheme1_f32(self.handle, dx.data_ptr(), dy.data_ptr(), db.data_ptr(), len(bounds) - 1, H, W,
                                             int(theta), float(active_v), float(silent_v), C.byref(sp), float(dt),
                                             dw.data_ptr(), self._stream())
        self.lib.check(rc, "event_scheme1")
        return dw.cpu().numpy()

    def simulate_scheme2(self, x, y, pol, t, H, W, slice_us, active_v=-6.0, silent_v=0.0, polarity="split", p=PARAMS, dt=DT):
        """Scheme 2 of simulate() (eventsim/event_mem_sim.py:231-286) over the whole stream: DC bias + event overlay
        with the per-pixel refractory period.  Returns (w_a, w_b) -- w_b is None for polarity='magnitude'."""
        assert polarity in ("split", "magnitude"), "polarity must be 'split' or 'magnitude'"
        t = np.asarray(t)
        bounds = slice_bounds(t, slice_us)
        arrs = [np.ascontiguousarray(x, dtype=np.int16), np.ascontiguousarray(y, dtype=np.int16),
                np.ascontiguousarray(pol, dtype=np.int8), np.ascontiguousarray(t, dtype=np.int64), bounds.astype(np.int64)]
        dx, dy, dp, dt_, db = self._dev_roundtrip(arrs)
        two = polarity == "split"
        w0 = np.full((H, W), p["wini"], dtype=np.float32)
        nk0 = np.zeros((H, W), dtype=np.int64)
        dwa, dna = self._dev_roundtrip([w0, nk0])
        dwb, dnb = self._dev_roundtrip([w0, nk0]) if two else (None, None)
        sp = _synapse_params(p)
        rc = self.lib.nsof_event_scheme2_f32(self.handle, dx.data_ptr(), dy.data_ptr(), dp.data_ptr(), dt_.data_ptr(),
                                             db.data_ptr(), len(bounds) - 1, H, W, 0 if two else 1, float(active_v),
                                             float(silent_v), int(REFRACTORY_US), C.byref(sp), float(dt), dwa.data_ptr(),
                                             dwb.data_ptr() if two else None, dna.data_ptr(),
                                             dnb.data_ptr() if two else None, self._stream())
        self.lib.check(rc, "event_scheme2")
        return dwa.cpu().numpy(), (dwb.cpu().numpy() if two else None)

    def event_cell_currents(self, w, memsize, p=PARAMS):
        """Event-mode ROI glue: float32 state (H, W) -> float64 cell currents (H // memsize, W // memsize), the
        largest pixel current 1 / resistance_exp(w) of every cell (feeds roi_from_currents)."""
        import torch
        w32 = np.ascontiguousarray(w, dtype=np.float32)
        H, W = w32.shape
        (dw,) = self._dev_roundtrip([w32])
        out = torch.empty((H // memsize, W // memsize), dtype=torch.float64, device=dw.device)
        sp = _synapse_params(p)
        self.lib.check(self.lib.nsof_event_cell_currents_f32(self.handle, dw.data_ptr(), H, W, memsize, memsize,
                                                             C.byref(sp), out.data_ptr(), self._stream()))
        return out.cpu().numpy()

    def compress_images(self, frames_u8, m, n):
        """compress_image (vehicle.m:111-121) for a stack of grey uint8 frames (T, H, W) -> float64 (T, H//n, W//m)."""
        import torch
        fr = np.ascontiguousarray(frames_u8, dtype=np.uint8)
        if fr.ndim == 2:
            fr = fr[None]
        T, H, W = fr.shape
        Hc, Wc = H // n, W // m
        (d,) = self._dev_roundtrip([fr])
        out = torch.empty((T, Hc, Wc), dtype=torch.float64, device=d.device)
        rc = self.lib.nsof_lanczos3_compress_u8(self.handle, d.data_ptr(), T, H, W, W, H * W, Hc, Wc, out.data_ptr(),
                                                self._stream())
        self.lib.check(rc, "lanczos3_compress")
        return out.cpu().numpy()

    def _w_init(self, w0, Hc, Wc, p, device):
        import torch
        if w0 is None:
            return torch.full((Hc, Wc), float(p["wini"]), dtype=torch.float64, device=device)
        w0 = np.ascontiguousarray(np.broadcast_to(np.asarray(w0, dtype=np.float64), (Hc, Wc)))
        return self._dev_roundtrip([w0])[0]

    def simulate_memristor_array_stepwise(self, compressed_images, dt=5e-4, nSubSteps=1000, th1=2.0, th2=1.5, p=PARAMS,
                                          vds=1.0, w0=None):
        """Same as simulate_memristor_array through the per-pair entry point nsof_frame_sim_step_f64 (streaming use:
        one call per arriving frame).  `w0`: initial state per cell instead of wini (resuming a sequence)."""
        import torch
        cells = np.ascontiguousarray(compressed_images, dtype=np.float64)
        T, Hc, Wc = cells.shape
        (d,) = self._dev_roundtrip([cells])
        w = self._w_init(w0, Hc, Wc, p, d.device)
        cur = torch.empty((T, Hc, Wc), dtype=torch.float64, device=d.device)
        sp = _synapse_params(p)
        lam = np.log(p["Roff"] / p["Ron"])
        cur[0] = vds / (p["Ron"] / torch.exp(-lam * (1.0 - w)))
        fp = FrameSimParams(float(th1), float(th2), float(dt), int(nSubSteps), float(vds))
        st = self._stream()
        for i in range(T - 1):
            rc = self.lib.nsof_frame_sim_step_f64(self.handle, d[i].data_ptr(), d[i + 1].data_ptr(), w.data_ptr(),
                                                  cur[i + 1].data_ptr(), 1, Hc, Wc, C.byref(sp), C.byref(fp), st)
            self.lib.check(rc, "frame_sim_step")
        return w.cpu().numpy(), cur.cpu().numpy()

    def simulate_memristor_array(self, compressed_images, dt=5e-4, nSubSteps=1000, th1=2.0, th2=1.5, p=PARAMS, vds=1.0,
                                 w0=None):
        """simulate_memristor_array (vehicle.m:187-227): returns (w_array, currents (T, Hc, Wc) with slice 0 the
        initial state -- the layout of constructed3DMatrix = vds / resistances_over_time).  `w0`: initial state per
        cell instead of wini."""
        import torch
        cells = np.ascontiguousarray(compressed_images, dtype=np.float64)
        T, Hc, Wc = cells.shape
        (d,) = self._dev_roundtrip([cells])
        w = self._w_init(w0, Hc, Wc, p, d.device)
        cur = torch.empty((T, Hc, Wc), dtype=torch.float64, device=d.device)
        sp = _synapse_params(p)
        lam = np.log(p["Roff"] / p["Ron"])
        cur[0] = vds / (p["Ron"] / torch.exp(-lam * (1.0 - w)))
        fp = FrameSimParams(float(th1), float(th2), float(dt), int(nSubSteps), float(vds))
        rc = self.lib.nsof_frame_sim_run_f64(self.handle, d.data_ptr(), T, 1, Hc, Wc, w.data_ptr(), cur[1:].data_ptr(),
                                             C.byref(sp), C.byref(fp), self._stream())
        self.lib.check(rc, "frame_sim_run")
        return w.cpu().numpy(), cur.cpu().numpy()


def slice_bounds(t, slice_us):
    """Event index bounds of the fixed-duration windows of slice_indices (eventsim/event_mem_sim.py:78-83)."""
    t = np.asarray(t)
    bounds = np.arange(t[0], t[-1] + slice_us, slice_us, dtype=t.dtype)
    return np.searchsorted(t, bounds)


def slice_indices(t, slice_us):
    """eventsim/event_mem_sim.py:78-83 (host index arithmetic, identical results)."""
    idx = slice_bounds(t, slice_us)
    for i in range(len(idx) - 1):
        yield slice(idx[i], idx[i + 1])
