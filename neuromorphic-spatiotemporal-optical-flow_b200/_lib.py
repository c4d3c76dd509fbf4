"""ctypes binding of libnsof (include/nsof.h).

The product library is ``csrc/libnsof.so`` (hand-written sm_100a CUDA, built in-tree by
``__graft_entry__.build()`` / ``csrc/Makefile``).  There is NO CPU fallback: if the shared library is
missing, or no CUDA device is usable, loading / ``Library.create_handle`` raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# NSOF_LIB_PATH: kernel-tuning experiments load an alternative build of the same sources
DEFAULT_LIB_PATH = os.environ.get("NSOF_LIB_PATH") or os.path.join(_HERE, "csrc", "libnsof.so")

NSOF_OK = 0
NSOF_ERR_BAD_ARG = -1
NSOF_ERR_UNSUPPORTED = -2
NSOF_ERR_CUDA = -3
NSOF_ERR_NOMEM = -4
NSOF_MAX_RECTS = 256
NSOF_FB_SRC_F32 = 0x10000  # prev / next are float32 frames (include/nsof.h)


class error(ValueError):
    """Bad-argument error; plays the role of ``cv2.error`` (-215 assertion) for the drop-in calls."""


class FbParams(C.Structure):
    _fields_ = [("pyr_scale", C.c_double), ("levels", C.c_int), ("winsize", C.c_int), ("iterations", C.c_int),
                ("poly_n", C.c_int), ("poly_sigma", C.c_double), ("flags", C.c_int)]


class CropDesc(C.Structure):
    _fields_ = [("prev", C.c_void_p), ("next", C.c_void_p), ("h", C.c_int), ("w", C.c_int),
                ("prev_pitch", C.c_ssize_t), ("next_pitch", C.c_ssize_t), ("flow", C.c_void_p),
                ("flow_pitch", C.c_ssize_t)]


class Rect(C.Structure):
    _fields_ = [("x0", C.c_int), ("y0", C.c_int), ("x1", C.c_int), ("y1", C.c_int)]


class RoiParams(C.Structure):
    _fields_ = [("cells_h", C.c_int), ("cells_w", C.c_int), ("frame_h", C.c_int), ("frame_w", C.c_int),
                ("pixel_h", C.c_int), ("pixel_w", C.c_int), ("ext_left", C.c_int), ("ext_right", C.c_int),
                ("ext_upper", C.c_int), ("ext_lower", C.c_int), ("flag", C.c_int), ("thres", C.c_int),
                ("map_a", C.c_double), ("map_b", C.c_double)]


class SynapseParams(C.Structure):
    _fields_ = [(k, C.c_double) for k in ("alphaoff", "alphaon", "voff", "von", "koff", "kon", "son", "soff", "bon",
                                          "boff", "Ron", "Roff", "won", "woff", "wini")]


class FrameSimParams(C.Structure):
    _fields_ = [("th1", C.c_double), ("th2", C.c_double), ("dt", C.c_double), ("nsubsteps", C.c_int),
                ("vds", C.c_double)]


class TiledBuf(C.Structure):
    _fields_ = [("ptr", C.c_void_p), ("row_floats", C.c_size_t), ("plane_floats", C.c_size_t), ("planes", C.c_int),
                ("h", C.c_int), ("w", C.c_int), ("y_lo", C.c_int), ("y_hi", C.c_int)]


class CopyBlock(C.Structure):
    _fields_ = [("dst", C.c_void_p), ("src", C.c_void_p), ("bytes", C.c_size_t)]


class SegParams(C.Structure):  # nsof_seg_params
    _fields_ = [("mag_thresh", C.c_double), ("shape", C.c_int), ("ksize_w", C.c_int), ("ksize_h", C.c_int),
                ("iterations", C.c_int)]


class SegDesc(C.Structure):  # nsof_seg_desc
    _fields_ = [("flow", C.c_void_p), ("flow_pitch", C.c_ssize_t), ("h", C.c_int), ("w", C.c_int),
                ("mask", C.c_void_p), ("mask_pitch", C.c_ssize_t)]


class WarpDesc(C.Structure):  # nsof_warp_desc
    _fields_ = [("frame", C.c_void_p), ("frame_pitch", C.c_ssize_t), ("H", C.c_int), ("W", C.c_int), ("channels", C.c_int),
                ("flow", C.c_void_p), ("flow_pitch", C.c_ssize_t), ("x0", C.c_int), ("y0", C.c_int), ("h", C.c_int),
                ("w", C.c_int), ("out", C.c_void_p), ("out_pitch", C.c_ssize_t)]


class KernelStat(C.Structure):
    _fields_ = [("name", C.c_char * 32), ("launches", C.c_int), ("ms", C.c_double), ("algorithmic_bytes", C.c_double)]


_P = C.POINTER
_SIGNATURES = {
    # name: (restype, argtypes)
    "nsof_last_error": (C.c_char_p, []),
    "nsof_version": (C.c_int, []),
    "nsof_create": (C.c_int, [C.c_int, _P(C.c_void_p)]),
    "nsof_destroy": (C.c_int, [C.c_void_p]),
    "nsof_reserve": (C.c_int, [C.c_void_p, C.c_size_t, C.c_int]),
    "nsof_workspace_bytes": (C.c_size_t, [C.c_void_p]),
    "nsof_launch_count": (C.c_ulonglong, [C.c_void_p]),
    "nsof_profile": (C.c_int, [C.c_void_p, C.c_int]),
    "nsof_profile_read": (C.c_int, [C.c_void_p, _P(KernelStat), C.c_int, _P(C.c_int)]),
    "nsof_farneback_batch": (C.c_int, [C.c_void_p, _P(CropDesc), C.c_int, _P(FbParams), C.c_void_p]),
    "nsof_farneback_u8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_ssize_t, C.c_ssize_t,
                                    _P(FbParams), C.c_void_p, C.c_ssize_t, C.c_void_p]),
    "nsof_farneback_batch_host": (C.c_int, [C.c_void_p, _P(CropDesc), C.c_int, _P(FbParams)]),
    "nsof_farneback_u8_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_ssize_t,
                                         C.c_ssize_t, _P(FbParams), C.c_void_p, C.c_ssize_t]),
    "nsof_flow_clear": (C.c_int, [C.c_void_p, _P(CropDesc), C.c_int, C.c_void_p]),
    "nsof_copy_2d_async": (C.c_int, [C.c_void_p, C.c_ssize_t, C.c_void_p, C.c_ssize_t, C.c_size_t, C.c_size_t, C.c_int,
                                     C.c_void_p]),
    "nsof_fb_tiled_begin": (C.c_int, [C.c_void_p, _P(CropDesc), C.c_int, _P(FbParams), C.c_int, C.c_int, C.c_int,
                                      C.c_void_p]),
    "nsof_fb_tiled_levels": (C.c_int, [C.c_void_p]),
    "nsof_fb_tiled_stage": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "nsof_fb_tiled_stage_rows": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "nsof_fb_tiled_query": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, _P(TiledBuf)]),
    "nsof_fb_tiled_guard": (C.c_int, [C.c_void_p, _P(C.c_int), C.c_void_p]),
    "nsof_fb_tiled_link_export": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, _P(C.c_void_p), _P(C.c_void_p),
                                            _P(C.c_size_t)]),
    "nsof_fb_tiled_link_attach": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]),
    "nsof_fb_tiled_run": (C.c_int, [C.c_void_p, C.c_int, C.c_ulonglong, _P(C.c_ulonglong), C.c_void_p]),
    "nsof_copy_blocks": (C.c_int, [C.c_void_p, _P(CopyBlock), C.c_int, C.c_void_p]),
    "nsof_copy_blocks_resident": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_size_t, C.c_double, C.c_void_p]),
    "nsof_flow_segment_batch": (C.c_int, [C.c_void_p, _P(SegDesc), C.c_int, _P(SegParams), C.c_void_p]),
    "nsof_mask_clear": (C.c_int, [C.c_void_p, _P(SegDesc), C.c_int, C.c_void_p]),
    "nsof_flow_segment_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_ssize_t, C.c_int, C.c_int, _P(SegParams),
                                         C.c_void_p, C.c_ssize_t]),
    "nsof_structuring_element": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "nsof_rgb2gray_u8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_ssize_t, C.c_ssize_t, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                   C.c_ssize_t, C.c_ssize_t, C.c_void_p]),
    "nsof_rgb2gray_u8_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_ssize_t, C.c_int, C.c_int, C.c_void_p, C.c_ssize_t]),
    "nsof_remap_flow_batch": (C.c_int, [C.c_void_p, _P(WarpDesc), C.c_int, C.c_int, C.c_void_p]),
    "nsof_remap_flow_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_ssize_t, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                       C.c_ssize_t, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_ssize_t]),
    "nsof_jpeg_info": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, _P(C.c_int), _P(C.c_int)]),
    "nsof_jpeg_decode_gray": (C.c_int, [C.c_void_p, _P(C.c_void_p), _P(C.c_size_t), C.c_int, C.c_int, C.c_int, C.c_void_p,
                                        C.c_ssize_t, C.c_ssize_t, C.c_void_p]),
    "nsof_ssim_u8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_ssize_t, C.c_int, C.c_void_p, C.c_ssize_t, C.c_int, C.c_int, C.c_int,
                               C.c_double, C.c_void_p, C.c_void_p]),
    "nsof_ssim_u8_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_ssize_t, C.c_int, C.c_void_p, C.c_ssize_t, C.c_int, C.c_int,
                                       C.c_int, C.c_double, C.c_void_p, C.c_void_p]),
    "nsof_ssim_u8_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_ssize_t, C.c_int, C.c_void_p, C.c_ssize_t, C.c_int, C.c_int,
                                    C.c_int, C.c_double, _P(C.c_double)]),
    "nsof_fb_level_geometry": (C.c_int, [C.c_int, C.c_int, C.c_double, C.c_int, _P(C.c_int), _P(C.c_int),
                                         _P(C.c_int), _P(C.c_double)]),
    "nsof_fb_algorithmic_bytes": (C.c_double, [C.c_int, C.c_int, _P(FbParams)]),
    "nsof_debug_truncate": (C.c_int, [C.c_void_p, C.c_int, C.c_int]),
    "nsof_debug_fuse": (C.c_int, [C.c_void_p, C.c_int]),
    "nsof_debug_fork": (C.c_int, [C.c_void_p, C.c_int]),
    "nsof_debug_share_frames": (C.c_int, [C.c_void_p, C.c_int]),
    "nsof_debug_lanczos_f64": (C.c_int, [C.c_void_p, C.c_int]),
    "nsof_debug_big_tiles": (C.c_int, [C.c_void_p, C.c_int]),
    "nsof_debug_generic_blur": (C.c_int, [C.c_void_p, C.c_int]),
    "nsof_debug_r1_window": (C.c_int, [C.c_void_p, C.c_int]),
    "nsof_debug_bs_prefetch": (C.c_int, [C.c_void_p, C.c_int]),
    "nsof_debug_r1_window_stats": (C.c_int, [C.c_void_p, _P(C.c_ulonglong), C.c_void_p]),
    "nsof_debug_fetch": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, _P(C.c_int), _P(C.c_int)]),
    "nsof_roi_from_state_f64": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, _P(RoiParams), C.c_void_p, C.c_void_p,
                                          C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "nsof_roi_from_state_f64_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, _P(RoiParams), C.c_void_p,
                                               C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                               C.c_void_p]),
    "nsof_event_hist": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p,
                                  C.c_void_p]),
    "nsof_synapse_step_f32": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, _P(SynapseParams),
                                        C.c_double, C.c_void_p]),
    "nsof_event_scheme1_f32": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                         C.c_int, C.c_float, C.c_float, _P(SynapseParams), C.c_double, C.c_void_p,
                                         C.c_void_p]),
    "nsof_event_scheme2_f32": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                         C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_longlong, _P(SynapseParams),
                                         C.c_double, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "nsof_resistance_exp_f32": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, _P(SynapseParams), C.c_void_p,
                                          C.c_void_p]),
    "nsof_event_cell_currents_f32": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                               _P(SynapseParams), C.c_void_p, C.c_void_p]),
    "nsof_frame_sim_step_f64": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                          C.c_int, C.c_int, _P(SynapseParams), _P(FrameSimParams), C.c_void_p]),
    "nsof_frame_sim_run_f64": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                         C.c_void_p, _P(SynapseParams), _P(FrameSimParams), C.c_void_p]),
    "nsof_lanczos3_compress_u8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_ssize_t,
                                            C.c_ssize_t, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
}
EXPORTED_SYMBOLS = tuple(_SIGNATURES)


class Library:
    """One loaded libnsof shared object with typed entry points."""

    def __init__(self, path: str | None = None):
        self.path = path or DEFAULT_LIB_PATH
        if not os.path.exists(self.path):
            raise ImportError(
                f"libnsof not found at {self.path}: build it with `python -c 'import __graft_entry__ as g; "
                f"g.build()'` (or `make -C {os.path.dirname(DEFAULT_LIB_PATH)}`). There is no CPU fallback.")
        self.cdll = C.CDLL(self.path)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(self.cdll, name)  # AttributeError here means the .so is stale
            fn.restype = res
            fn.argtypes = args
            setattr(self, name, fn)

    def last_error(self) -> str:
        msg = self.nsof_last_error()
        return msg.decode("utf-8", "replace") if msg else ""

    def check(self, rc: int, what: str = "") -> None:
        if rc == NSOF_OK:
            return
        msg = f"{what + ': ' if what else ''}{self.last_error()} (status {rc})"
        if rc == NSOF_ERR_BAD_ARG:
            raise error(msg)
        if rc == NSOF_ERR_UNSUPPORTED:
            raise NotImplementedError(msg)
        if rc == NSOF_ERR_NOMEM:
            raise MemoryError(msg)
        raise RuntimeError(msg)

    def create_handle(self, device: int = 0) -> C.c_void_p:
        h = C.c_void_p()
        self.check(self.nsof_create(int(device), C.byref(h)), "nsof_create")
        return h


_default: Library | None = None


def default_library() -> Library:
    global _default
    if _default is None:
        _default = Library()
    return _default
