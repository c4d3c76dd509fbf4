"""nsof-b200: B200-native (sm_100a) ROI-gated Farneback hot path of
RTCartist/Neuromorphic-Spatiotemporal-Optical-Flow.

Module-level functions mirror the reference's call sites and run on a lazily created default
`Engine` on cuda:0.  Importing the package does not need a GPU; the first call does (no CPU fallback).

    import importlib; nsof = importlib.import_module("neuromorphic-spatiotemporal-optical-flow_b200")
    flow = nsof.calcOpticalFlowFarneback(prev_region, next_region, None, **nsof.FARNEBACK_PARAMS)
"""
from ._lib import DEFAULT_LIB_PATH, EXPORTED_SYMBOLS, Library, error  # noqa: F401
from .engine import (DT, FARNEBACK_PARAMS, PARAMS, REFRACTORY_US, SEG_TH, THETA_EVENTS, Engine, RoiConfig,  # noqa: F401
                     current_to_u8_reference_formula, slice_bounds, slice_indices, u8_threshold_table)

_engines: dict = {}


def default_engine(device: int = 0) -> Engine:
    if device not in _engines:
        _engines[device] = Engine(device)
    return _engines[device]


def calcOpticalFlowFarneback(prev, next, flow, pyr_scale, levels, winsize, iterations, poly_n, poly_sigma, flags):
    """cv2.calcOpticalFlowFarneback drop-in (optical_flow_seg.py:203 call site)."""
    return default_engine().calcOpticalFlowFarneback(prev, next, flow, pyr_scale, levels, winsize, iterations, poly_n,
                                                     poly_sigma, flags)


def opticalFlow3D(memimg1, memimg2, rgbimg1, rgbimg2, pixel_width, pixel_height):
    """optical_flow_seg.py:211."""
    return default_engine().opticalFlow3D(memimg1, memimg2, rgbimg1, rgbimg2, pixel_width, pixel_height)


def update_transition_pic(prev_memristor, transition_pic, thres):
    """optical_flow_seg.py:116."""
    return default_engine().update_transition_pic(prev_memristor, transition_pic, thres)


def process_flow_region(mag, ang=None, flow_region=None):
    """optical_flow_seg.py:322 drop-in.  The reference's result depends on `mag > SEG_TH` only, and `mag` is
    cv2.cartToPolar of the flow region: pass `flow_region=` (h, w, 2) to have the magnitude test evaluated on the
    device exactly as cv2.cartToPolar does (float32 arithmetic).  With `mag` alone the region (mag, 0) is
    segmented; a float64 magnitude is rounded to float32 without moving it across SEG_TH."""
    import numpy as np
    if flow_region is None:
        m = np.asarray(mag)
        m32 = m.astype(np.float32)
        th32 = np.float32(SEG_TH)
        m32 = np.where((m > SEG_TH) & (m32 <= th32), np.nextafter(th32, np.float32(np.inf)), m32)
        flow_region = np.stack([m32, np.zeros_like(m32)], axis=-1)
    return default_engine().segment_flow(flow_region)


def update_state(w, V, p=PARAMS, dt=DT):
    """eventsim/event_mem_sim.py:40."""
    return default_engine().update_state(w, V, p, dt)


def bincount_2d(x, y, H, W):
    """eventsim/event_mem_sim.py:100."""
    return default_engine().bincount_2d(x, y, H, W)


def resistance_exp(w, p=PARAMS):
    """eventsim/event_mem_sim.py:60."""
    return default_engine().resistance_exp(w, p)
