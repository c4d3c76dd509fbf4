"""Sequence driver: the body of the reference scripts' main loop (optical_flow_seg.py:413-461) for whole clips,
with stage 1 taken from the frame-mode simulator (simulation/simulationcode_v4_transistor_vehicle.m:187-227)
instead of a pre-computed .mat, everything resident on one GPU.

For a rank's `C` clips of `F` grey frames (layout [F][C][H][W] uint8, so that time slice t of every clip is one
contiguous block) one `step` is:

  1. compress_image for all F*C frames                       nsof_lanczos3_compress_u8   (1 B / source pixel)
  2. simulate_memristor_array, all F-1 pairs in one launch   nsof_frame_sim_run_f64      (cells only)
  3. current -> u8 -> transition_pic -> CC -> rectangles     nsof_roi_from_state_f64     (cells only)
  4. rectangles D2H (the one host sync of the step), crop descriptors
  5. zero the flow frames, ROI-gated Farneback on every crop nsof_farneback_batch        (DESIGN.md byte model)

torch is used for device memory, streams and pinned host buffers only.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import FrameSimParams, RoiParams
from .engine import PARAMS, Engine, RoiConfig, _fb_params, _synapse_params, u8_threshold_table


class SequencePipeline:
    def __init__(self, engine: Engine, clips: int, frames: int, H: int, W: int, cfg: RoiConfig | None = None,
                 sim=dict(dt=5e-4, nSubSteps=1000, th1=2.0, th2=1.5, vds=1.0), full_frame: bool = False):
        import torch
        self.torch = torch
        self.eng, self.lib, self.h = engine, engine.lib, engine.handle
        self.C, self.F, self.H, self.W = int(clips), int(frames), int(H), int(W)
        self.cfg = cfg or RoiConfig()
        self.full_frame = bool(full_frame)  # ROI = 100 %: skip the gate, run Farneback on whole frames
        self.dev = torch.device(f"cuda:{engine.device}")
        M = self.cfg.MEMSIZE
        self.Hc, self.Wc = self.H // M, self.W // M
        self.n_pairs = self.C * (self.F - 1)
        f64, i32 = torch.float64, torch.int32
        self.cells = torch.empty((self.F, self.C, self.Hc, self.Wc), dtype=f64, device=self.dev)
        self.w = torch.empty((self.C, self.Hc, self.Wc), dtype=f64, device=self.dev)
        # currents[t] = array state after the pair (t, t+1): slice t of constructed3DMatrix
        self.currents = torch.empty((self.F - 1, self.C, self.Hc, self.Wc), dtype=f64, device=self.dev)
        self.thr = torch.from_numpy(u8_threshold_table(self.cfg.MAP_A, self.cfg.MAP_B)).to(self.dev)
        self.rects = torch.zeros((self.n_pairs, _lib.NSOF_MAX_RECTS, 4), dtype=i32, device=self.dev)
        self.n_rects = torch.zeros((self.n_pairs,), dtype=i32, device=self.dev)
        self.memimg = torch.zeros((self.n_pairs, self.Hc, self.Wc), dtype=torch.uint8, device=self.dev)
        self.rects_host = torch.zeros((self.n_pairs, 4), dtype=i32).pin_memory()
        self.n_rects_host = torch.zeros((self.n_pairs,), dtype=i32).pin_memory()
        self.flow = torch.zeros((self.F - 1, self.C, self.H, self.W, 2), dtype=torch.float32, device=self.dev)
        self.sp = _synapse_params(PARAMS)
        self.fsp = FrameSimParams(float(sim["th1"]), float(sim["th2"]), float(sim["dt"]), int(sim["nSubSteps"]),
                                  float(sim["vds"]))
        c = self.cfg
        self.rp = RoiParams(self.Hc, self.Wc, self.H, self.W, M, M, c.EXTEND_WIDTH_LEFT, c.EXTEND_WIDTH_RIGHT,
                            c.EXTEND_HEIGHT_UPPER, c.EXTEND_HEIGHT_LOWER, 2, c.THRES, c.MAP_A, c.MAP_B)
        self.fb = _fb_params(**c.farneback_params)
        self.descs = (_lib.CropDesc * self.n_pairs)()
        self.last_rects = None
        self.last_crop_pixels = 0
        engine.reserve(self.n_pairs * self.H * self.W if full_frame else self.n_pairs * self.H * self.W // 3,
                       self.n_pairs)

    def _stream(self):
        return C.c_void_p(self.torch.cuda.current_stream(self.dev).cuda_stream)

    # ---- stage 1 ---------------------------------------------------------------------------------------
    def stage1(self, frames):
        """frames: uint8 device tensor [F][C][H][W].  Leaves rectangles in self.rects / self.n_rects (device)."""
        lib, h, st = self.lib, self.h, self._stream()
        F, Cn, H, W = self.F, self.C, self.H, self.W
        lib.check(lib.nsof_lanczos3_compress_u8(h, frames.data_ptr(), F * Cn, H, W, W, H * W, self.Hc, self.Wc,
                                                self.cells.data_ptr(), st), "lanczos3_compress")
        self.w.fill_(float(PARAMS["wini"]))
        lib.check(lib.nsof_frame_sim_run_f64(h, self.cells.data_ptr(), F, Cn, self.Hc, self.Wc, self.w.data_ptr(),
                                             self.currents.data_ptr(), C.byref(self.sp), C.byref(self.fsp), st),
                  "frame_sim_run")
        lib.check(lib.nsof_roi_from_state_f64(h, self.currents.data_ptr(), self.n_pairs, C.byref(self.rp),
                                              self.thr.data_ptr(), self.memimg.data_ptr(), None, None,
                                              self.n_rects.data_ptr(), self.rects.data_ptr(), None, st), "roi")

    def fetch_rects(self):
        """The step's one host synchronisation: merged rectangle (FLAG 2) of every pair -> host."""
        self.rects_host.copy_(self.rects[:, 0, :], non_blocking=True)
        self.n_rects_host.copy_(self.n_rects, non_blocking=True)
        self.torch.cuda.current_stream(self.dev).synchronize()
        return self.rects_host.numpy(), self.n_rects_host.numpy()

    # ---- stage 2 ---------------------------------------------------------------------------------------
    def build_descs(self, frames, rects, n_rects):
        """Crop descriptors: views into the resident frames and into the full-frame flow buffers (no copies)."""
        H, W, Cn = self.H, self.W, self.C
        fbase, obase = frames.data_ptr(), self.flow.data_ptr()
        n, px = 0, 0
        for i in range(self.n_pairs):
            if self.full_frame:
                x0, y0, x1, y1 = 0, 0, W, H
            else:
                if n_rects[i] == 0:
                    continue
                x0, y0, x1, y1 = (int(v) for v in rects[i])
                if x1 <= x0 or y1 <= y0:  # prev_region.size == 0 guard of optical_flow_seg.py:201
                    continue
            t, c = divmod(i, Cn)
            off = y0 * W + x0
            d = self.descs[n]
            d.prev = fbase + ((t * Cn + c) * H * W + off)
            d.next = fbase + (((t + 1) * Cn + c) * H * W + off)
            d.h, d.w = y1 - y0, x1 - x0
            d.prev_pitch = d.next_pitch = W
            d.flow = obase + ((t * Cn + c) * H * W + off) * 8
            d.flow_pitch = W * 8
            n += 1
            px += (y1 - y0) * (x1 - x0)
        self.last_crop_pixels = px
        return n

    def stage2(self, n_items):
        self.flow.zero_()  # S2-g: the reference pastes every crop's flow into a zero frame
        self.lib.check(self.lib.nsof_farneback_batch(self.h, self.descs, n_items, C.byref(self.fb), self._stream()),
                       "farneback_batch")

    def step(self, frames):
        """One pass of the hot path over the rank's clips, inputs and outputs resident in HBM."""
        if self.full_frame:
            n = self.build_descs(frames, None, None)
        else:
            self.stage1(frames)
            rects, n_rects = self.fetch_rects()
            self.last_rects = (rects.copy(), n_rects.copy())
            n = self.build_descs(frames, rects, n_rects)
        self.stage2(n)
        return n

    # ---- host-buffer entry (the e2e leg) ---------------------------------------------------------------
    def alloc_host(self):
        """Pinned host buffers: frames in, packed ROI flow crops + rectangles out."""
        t = self.torch
        self.frames_dev = t.empty((self.F, self.C, self.H, self.W), dtype=t.uint8, device=self.dev)
        self.crop_host = t.empty((self.n_pairs * self.H * self.W * 2 // (1 if self.full_frame else 3),),
                                 dtype=t.float32).pin_memory()
        self.crop_dev = t.empty_like(self.crop_host, device=self.dev)

    def step_host(self, frames_host):
        """frames_host: pinned uint8 [F][C][H][W].  H2D of the frames, the device step, D2H of the rectangles and of
        every ROI's flow (packed float32 crops; outside its rectangle a pair's flow is zero by definition).
        Returns (rects, n_rects, offsets, packed flow host tensor)."""
        t = self.torch
        self.frames_dev.copy_(frames_host, non_blocking=True)
        n = self.step(self.frames_dev)
        # gather the ROI windows of the full-frame flow buffers into one packed buffer, then one D2H
        offs, o = [], 0
        flow = self.flow.view(self.n_pairs, self.H, self.W, 2)
        rects, n_rects = (None, None) if self.full_frame else self.last_rects
        for i in range(self.n_pairs):
            if self.full_frame:
                x0, y0, x1, y1 = 0, 0, self.W, self.H
            elif n_rects[i] == 0:
                offs.append(-1)
                continue
            else:
                x0, y0, x1, y1 = (int(v) for v in rects[i])
            sz = (y1 - y0) * (x1 - x0) * 2
            self.crop_dev[o:o + sz].view(y1 - y0, x1 - x0, 2).copy_(flow[i, y0:y1, x0:x1])
            offs.append(o)
            o += sz
        self.crop_host[:o].copy_(self.crop_dev[:o], non_blocking=True)
        t.cuda.current_stream(self.dev).synchronize()
        self.last_d2h_bytes = o * 4 + self.n_pairs * 20
        self.last_h2d_bytes = frames_host.numel()
        return rects, n_rects, offs, self.crop_host
