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
                 sim=dict(dt=5e-4, nSubSteps=1000, th1=2.0, th2=1.5, vds=1.0), full_frame: bool = False,
                 packed: bool = False):
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
        # output: full flow frames (zero outside each ROI, the reference's paste semantics) or, for the host path,
        # the ROI crops packed back to back (outside its rectangle a pair's flow is zero by definition)
        self.packed_mode = bool(packed)
        if packed:
            cap = self.n_pairs * self.H * self.W * 2 // (1 if full_frame else 3)
            self.packed = torch.empty((cap,), dtype=torch.float32, device=self.dev)
            self.flow = None
        else:
            self.flow = torch.zeros((self.F - 1, self.C, self.H, self.W, 2), dtype=torch.float32, device=self.dev)
        self.offsets = [-1] * self.n_pairs
        self.packed_floats = 0
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
        fbase = frames.data_ptr()
        obase = self.packed.data_ptr() if self.packed_mode else self.flow.data_ptr()
        n, px = 0, 0
        for i in range(self.n_pairs):
            self.offsets[i] = -1
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
            if self.packed_mode:
                if (px + (y1 - y0) * (x1 - x0)) * 2 > self.packed.numel():
                    raise MemoryError("packed flow buffer too small for this step's ROIs")
                self.offsets[i] = px * 2
                d.flow = obase + px * 8
                d.flow_pitch = (x1 - x0) * 8
            else:
                d.flow = obase + ((t * Cn + c) * H * W + off) * 8
                d.flow_pitch = W * 8
            n += 1
            px += (y1 - y0) * (x1 - x0)
        self.last_crop_pixels = px
        self.packed_floats = px * 2
        return n

    def stage2(self, n_items):
        if not self.packed_mode:
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



class HostPipeline:
    """Host-buffer entry of the sequence path (bench.py's e2e leg): pinned uint8 clips in, per pair its rectangle
    and its ROI flow crop (float32, packed) out, host<->device copies included.

    Clips are independent, so the work is software-pipelined over `slots` streams, one clip per stage:
    H2D of clip k+1 overlaps stage 1 / stage 2 of clip k and the D2H of clip k-1 (both copy engines and the SMs
    busy at once).  Every slot owns a libnsof handle (one handle per (GPU, stream)), its device frames and its
    packed output buffer; the only host synchronisation per clip is the event that its rectangles have arrived.
    """

    def __init__(self, device: int, clips: int, frames: int, H: int, W: int, cfg: RoiConfig | None = None,
                 slots: int = 4, full_frame: bool = False, lib=None):
        import torch
        self.torch = torch
        self.dev = torch.device(f"cuda:{device}")
        self.C, self.F, self.H, self.W = int(clips), int(frames), int(H), int(W)
        self.slots = []
        for _ in range(max(1, min(slots, self.C))):
            eng = Engine(device, lib=lib)
            pipe = SequencePipeline(eng, 1, frames, H, W, cfg, full_frame=full_frame, packed=True)
            self.slots.append(dict(eng=eng, pipe=pipe, stream=torch.cuda.Stream(self.dev),
                                   frames=torch.empty((frames, 1, H, W), dtype=torch.uint8, device=self.dev),
                                   ev_rects=torch.cuda.Event(), ev_done=torch.cuda.Event(), busy=False))
        per_clip = self.slots[0]["pipe"].packed.numel()
        # host outputs: one packed flow buffer and one rectangle table per clip
        self.flow_host = [torch.empty((per_clip,), dtype=torch.float32).pin_memory() for _ in range(self.C)]
        self.rects_out = [None] * self.C
        self.offsets_out = [None] * self.C
        self.h2d_bytes = self.d2h_bytes = 0

    def launch_count(self) -> int:
        return sum(s["eng"].launch_count() for s in self.slots)

    def _front(self, k, clip_host):
        """H2D + stage 1 + rectangles D2H of clip k on its slot's stream (asynchronous)."""
        t = self.torch
        s = self.slots[k % len(self.slots)]
        if s["busy"]:
            s["ev_done"].synchronize()  # the slot's previous clip has left the device
        with t.cuda.stream(s["stream"]):
            s["frames"][:, 0].copy_(clip_host, non_blocking=True)
            p = s["pipe"]
            if not p.full_frame:
                p.stage1(s["frames"])
                p.rects_host.copy_(p.rects[:, 0, :], non_blocking=True)
                p.n_rects_host.copy_(p.n_rects, non_blocking=True)
            s["ev_rects"].record()
        s["busy"] = True
        self.h2d_bytes += clip_host.numel()

    def _back(self, k):
        """Wait for clip k's rectangles, then enqueue its Farneback batch and the D2H of the packed crops."""
        t = self.torch
        s = self.slots[k % len(self.slots)]
        p = s["pipe"]
        s["ev_rects"].synchronize()
        if p.full_frame:
            rects, n_rects = None, None
        else:
            rects, n_rects = p.rects_host.numpy().copy(), p.n_rects_host.numpy().copy()
        n = p.build_descs(s["frames"], rects, n_rects)
        with t.cuda.stream(s["stream"]):
            p.stage2(n)
            nf = p.packed_floats
            self.flow_host[k][:nf].copy_(p.packed[:nf], non_blocking=True)
            s["ev_done"].record()
        self.rects_out[k] = (rects, n_rects)
        self.offsets_out[k] = list(p.offsets)
        self.d2h_bytes += nf * 4 + p.n_pairs * 20

    def run(self, clips_host):
        """clips_host: sequence of C pinned uint8 tensors (F, H, W).  Returns per clip (rects, n_rects, offsets,
        packed float32 host tensor); pair t's flow crop is packed[offsets[t] : offsets[t] + 2 h w].view(h, w, 2)."""
        assert len(clips_host) == self.C
        self.h2d_bytes = self.d2h_bytes = 0
        depth = len(self.slots) - 1  # clips in flight ahead of the one being finished
        for k in range(self.C):
            self._front(k, clips_host[k])
            if k >= depth:
                self._back(k - depth)
        for k in range(max(0, self.C - depth), self.C):
            self._back(k)
        for s in self.slots:
            s["ev_done"].synchronize()
            s["busy"] = False
        return [(self.rects_out[k][0], self.rects_out[k][1], self.offsets_out[k], self.flow_host[k]) for k in range(self.C)]
