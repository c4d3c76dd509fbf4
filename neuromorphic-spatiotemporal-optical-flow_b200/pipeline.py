"""Sequence driver: the body of the reference scripts' main loop (optical_flow_seg.py:413-461) for whole clips,
with stage 1 taken from the frame-mode simulator (simulation/simulationcode_v4_transistor_vehicle.m:187-227)
instead of a pre-computed .mat, everything resident on one GPU.

For a rank's `C` clips of `F` grey frames (layout [F][C][H][W] uint8, so that time slice t of every clip is one
contiguous block) one `step` is, per chunk of consecutive frame pairs:

  1. compress_image for the chunk's frames                   nsof_lanczos3_compress_u8   (1 B / source pixel)
  2. simulate_memristor_array over the chunk's pairs         nsof_frame_sim_run_f64      (cells only)
  3. current -> u8 -> transition_pic -> CC -> rectangles     nsof_roi_from_state_f64     (cells only)
  4. rectangles D2H (one host wait per chunk), crop descriptors
  5. ROI-gated Farneback on the chunk's crops                nsof_farneback_batch        (DESIGN.md byte model)

Stage 1 is a recurrence in time (1000 dependent Euler sub-steps per pair): latency-bound, a handful of warps.
Stage 2 is the bandwidth-bound part.  The pairs are therefore processed in a few time chunks: steps 1-4 of
ALL chunks are enqueued on a side stream (its own libnsof handle: one handle per stream), step 5 of chunk j is
enqueued on the caller's stream as soon as chunk j's rectangles have arrived, so the Euler chains of the later
chunks run underneath the Farneback kernels of the earlier ones.

torch is used for device memory, streams, events and pinned host buffers only.
"""
from __future__ import annotations

import ctypes as C
import time

import numpy as np

from . import _lib
from ._lib import CropDesc, FrameSimParams, RoiParams, SegDesc, SegParams
from .engine import PARAMS, SEG_TH, Engine, RoiConfig, _fb_params, _synapse_params, u8_threshold_table

_DESC_SIZE = C.sizeof(CropDesc)


def chunk_schedule(n_pairs: int, first: int = 1, grow: int = 2, cap: int = 8):
    """Pair ranges [p0, p1) of a step: a short first chunk (stage 2 starts early), then growing ones (fewer,
    larger Farneback batches).  (15 pairs -> 1, 2, 4, 8.)"""
    out, p, n = [], 0, max(1, int(first))
    while p < n_pairs:
        q = min(n_pairs, p + n)
        out.append((p, q))
        p, n = q, min(max(1, int(cap)), n * max(1, int(grow)))
    return out


class SequencePipeline:
    def __init__(self, engine: Engine, clips: int, frames: int, H: int, W: int, cfg: RoiConfig | None = None,
                 sim=dict(dt=5e-4, nSubSteps=1000, th1=2.0, th2=1.5, vds=1.0), full_frame: bool = False,
                 packed: bool = False, chunks=None, overlap: bool = True, s2_streams: int = 2, segment: bool = False,
                 roi_state: str = "scripts", stage1: str = "simulate"):
        """roi_state: which array state gates pair (t, t+1).  "scripts" (default): the state BEFORE the
        pair -- all four scripts overwrite memimg2 with memimg1 (optical_flow_seg.py:435), so pair i reads slice
        OFFSET + i of constructed3DMatrix (the initial state for pair 0: every cell above THRES, ROI = the whole cell
        grid).  "notebook": the state AFTER the pair, slice OFFSET + i + 1, the un-bugged form of demo.ipynb cell 11.
        stage1: "simulate" (frame-mode simulator on the clip's frames) or "given" (the caller fills `self.currents`
        [F][C][Hc][Wc] with pre-computed array states, slice t = constructed3DMatrix[:, :, OFFSET + t], as the
        scripts read them from the .mat: steps 1-2 are skipped)."""
        import torch
        self.torch = torch
        self.eng, self.lib, self.h = engine, engine.lib, engine.handle
        self.segment = bool(segment)  # also run task_results / process_flow_region on every ROI flow (8(f) rank 1)
        self.C, self.F, self.H, self.W = int(clips), int(frames), int(H), int(W)
        self.cfg = cfg or RoiConfig()
        if self.cfg.FLAG not in (1, 2):
            raise ValueError("RoiConfig.FLAG must be 1 (separate regions) or 2 (merged region)")
        if self.cfg.FLAG == 1 and packed and segment:
            raise NotImplementedError("FLAG 1 with packed output hands out one flow crop per region; the motion mask of "
                                      "task_results needs the pasted frame: use the device-resident pipelines "
                                      "(SequencePipeline / DeviceStream with flow frames) or segment=False")
        self.sep = self.cfg.FLAG == 1 and not full_frame   # several regions per pair
        if roi_state not in ("scripts", "notebook") or stage1 not in ("simulate", "given"):
            raise ValueError("roi_state must be 'scripts' or 'notebook', stage1 'simulate' or 'given'")
        self.roi_state, self.stage1_mode = roi_state, stage1
        self.roi_shift = 0 if roi_state == "scripts" else 1
        self.full_frame = bool(full_frame)  # ROI = 100 %: skip the gate, run Farneback on whole frames
        self.dev = torch.device(f"cuda:{engine.device}")
        M = self.cfg.MEMSIZE
        self.Hc, self.Wc = self.H // M, self.W // M
        self.n_pairs = self.C * (self.F - 1)
        self.chunks = list(chunks) if chunks is not None else chunk_schedule(self.F - 1)
        assert self.chunks[0][0] == 0 and self.chunks[-1][1] == self.F - 1
        # stage 1 runs on a side stream with its own handle (one handle per (GPU, stream)); overlap=False keeps
        # everything on the caller's stream and handle (per-kernel profiling, debugging)
        self.overlap = bool(overlap) and not self.full_frame
        self.s1_eng = Engine(engine.device, lib=engine.lib) if self.overlap else engine
        # high priority: stage 1 is the latency chain of the step and needs few SM resources; its blocks must not
        # queue behind the stage-2 CTAs that fill the machine
        import os
        self.side = (torch.cuda.Stream(self.dev, priority=int(os.environ.get("NSOF_S1_PRIORITY", "-1")))
                     if self.overlap else None)
        self.ev_chunk = [torch.cuda.Event(enable_timing=True) for _ in self.chunks]
        # stage 2 of successive chunks alternates between the caller's stream and `s2_streams - 1` extra streams (own
        # handles): the small coarse-level launches of one batch run beside the large fine-level ones of another
        n_extra = max(0, int(s2_streams) - 1) if self.overlap else 0
        self.s2_extra = [(Engine(engine.device, lib=engine.lib), torch.cuda.Stream(self.dev)) for _ in range(n_extra)]
        self.ev_s2 = [torch.cuda.Event(enable_timing=True) for _ in self.chunks]
        f64, i32 = torch.float64, torch.int32
        self.cells = torch.empty((self.F, self.C, self.Hc, self.Wc), dtype=f64, device=self.dev)
        self.w = torch.empty((self.C, self.Hc, self.Wc), dtype=f64, device=self.dev)
        # currents[t] = slice t of constructed3DMatrix: [0] the initial state (w = wini), [t + 1] the
        # array state after the pair (t, t+1).  Pair t is gated by slice t + roi_shift.
        self.currents = torch.empty((self.F, self.C, self.Hc, self.Wc), dtype=f64, device=self.dev)
        lam = np.log(PARAMS["Roff"] / PARAMS["Ron"])    # calculate_resistances_exp (vehicle.m:233-236), vds / R
        self.currents[0] = float(sim["vds"]) / (PARAMS["Ron"] / np.exp(-lam * (1.0 - PARAMS["wini"])))
        self.thr = torch.from_numpy(u8_threshold_table(self.cfg.MAP_A, self.cfg.MAP_B)).to(self.dev)
        self.memimg = torch.zeros((self.n_pairs, self.Hc, self.Wc), dtype=torch.uint8, device=self.dev)
        # the ROI kernel stores its rectangles straight into pinned host memory (device-addressable under unified
        # addressing): a D2H copy would queue on the copy engine behind whole flow crops of earlier chunks / steps
        self.rects_pinned = torch.zeros((self.n_pairs, _lib.NSOF_MAX_RECTS, 4), dtype=i32).pin_memory()
        self.rects_host = self.rects_pinned[:, 0, :]          # FLAG 2: one merged rectangle per pair
        # FLAG 1: up to MAX_REGIONS regions per pair get a descriptor each (the ROI kernel stores up to NSOF_MAX_RECTS)
        self.max_regions = 32 if self.sep else 1
        self.n_rects_host = torch.zeros((self.n_pairs,), dtype=i32).pin_memory()
        # output: full flow frames (zero outside each ROI, the reference's paste semantics) or, for the host path,
        # the ROI crops packed back to back (outside its rectangle a pair's flow is zero by definition)
        self.packed_mode = bool(packed)
        if packed:
            # ROI crops: about a third of the frames' pixels plus, in the scripts' form, one whole frame per clip (the
            # initial state opens every cell); a step that needs more raises MemoryError in build_descs
            cap = (self.n_pairs * self.H * self.W * 2 if full_frame
                   else (self.n_pairs * self.H * self.W // 3 + self.C * self.H * self.W) * 2)
            self.packed = torch.empty((cap,), dtype=torch.float32, device=self.dev)
            self.flow = None
        else:
            self.flow = torch.zeros((self.F - 1, self.C, self.H, self.W, 2), dtype=torch.float32, device=self.dev)
        self.offsets = [-1] * self.n_pairs
        self.region_offsets = [[] for _ in range(self.n_pairs)]   # FLAG 1, packed: one offset per region, in label order
        self.packed_floats = 0
        # motion_binary (optical_flow_seg.py:262): uint8 frames, zero outside the ROI; packed mode: the crops' masks
        # back to back, pair i's mask at packed_mask[offsets[i] // 2 :][: h * w]
        self.motion = self.packed_mask = None
        if self.segment:
            if packed:
                self.packed_mask = torch.empty((self.packed.numel() // 2,), dtype=torch.uint8, device=self.dev)
            else:
                self.motion = torch.zeros((self.F - 1, self.C, self.H, self.W), dtype=torch.uint8, device=self.dev)
            self.seg_descs = (SegDesc * self.n_pairs)()   # FLAG 1: one per pair too (task_results' merged padding box)
            self.seg_params = SegParams(float(SEG_TH), 2, 10, 10, 5)  # optical_flow_seg.py:49, 349-351
        self.sp = _synapse_params(PARAMS)
        self.fsp = FrameSimParams(float(sim["th1"]), float(sim["th2"]), float(sim["dt"]), int(sim["nSubSteps"]),
                                  float(sim["vds"]))
        c = self.cfg
        self.rp = RoiParams(self.Hc, self.Wc, self.H, self.W, M, M, c.EXTEND_WIDTH_LEFT, c.EXTEND_WIDTH_RIGHT,
                            c.EXTEND_HEIGHT_UPPER, c.EXTEND_HEIGHT_LOWER, 1 if self.sep else 2, c.THRES, c.MAP_A, c.MAP_B)
        self.fb = _fb_params(**c.farneback_params)
        self.descs = (CropDesc * (self.n_pairs * self.max_regions))()
        self.n_items = 0          # descriptors of the current step
        self.n_dirty = 0          # descriptors whose flow rectangles are non-zero in self.flow (previous step)
        self.n_seg = self.n_seg_dirty = 0   # the same for the mask descriptors (FLAG 1: one per pair, not per region)
        self._groups = []         # FLAG 1: descriptor ranges of the chunk just built, one per region rank (see build_descs)
        self._seg_range = (0, 0)
        self.last_rects = None
        self.last_crop_pixels = 0
        # workspace hint for the largest Farneback batch of a step: ROI crops of about a fifth of the frame plus, in the
        # scripts' form, one whole frame per clip; a batch that needs more grows the arena (nsof_farneback_batch)
        biggest = max(q - p for p, q in self.chunks) * self.C if not full_frame else self.n_pairs
        for e in [engine] + [e for e, _ in self.s2_extra]:
            e.reserve(biggest * self.H * self.W if full_frame
                      else biggest * self.H * self.W // 5 + self.C * self.H * self.W, biggest)

    def _stream(self):
        return C.c_void_p(self.torch.cuda.current_stream(self.dev).cuda_stream)

    def launch_count(self) -> int:
        n = self.eng.launch_count() + sum(e.launch_count() for e, _ in self.s2_extra)
        return n + (self.s1_eng.launch_count() if self.s1_eng is not self.eng else 0)

    # ---- stage 1 ---------------------------------------------------------------------------------------
    def stage1_chunk(self, frames, j, eng=None):
        """Steps 1-3 for chunk j on the current stream: leaves the chunk's rectangles in rects_host / n_rects_host
        (visible to the host once an event recorded behind this call has completed)."""
        eng = eng or self.s1_eng
        lib, h, st = self.lib, eng.handle, self._stream()
        Cn, H, W = self.C, self.H, self.W
        p0, p1 = self.chunks[j]
        f0 = 0 if j == 0 else p0 + 1      # frames not compressed yet
        ncell = self.Hc * self.Wc
        if self.stage1_mode == "simulate":
            lib.check(lib.nsof_lanczos3_compress_u8(h, frames.data_ptr() + f0 * Cn * H * W, (p1 + 1 - f0) * Cn, H, W, W,
                                                    H * W, self.Hc, self.Wc, self.cells.data_ptr() + f0 * Cn * ncell * 8,
                                                    st), "lanczos3_compress")
            if j == 0:
                self.w.fill_(float(PARAMS["wini"]))
            lib.check(lib.nsof_frame_sim_run_f64(h, self.cells.data_ptr() + p0 * Cn * ncell * 8, p1 - p0 + 1, Cn, self.Hc,
                                                 self.Wc, self.w.data_ptr(),
                                                 self.currents.data_ptr() + (p0 + 1) * Cn * ncell * 8,
                                                 C.byref(self.sp), C.byref(self.fsp), st), "frame_sim_run")
        i0, n = p0 * Cn, (p1 - p0) * Cn
        lib.check(lib.nsof_roi_from_state_f64(h, self.currents.data_ptr() + (i0 + self.roi_shift * Cn) * ncell * 8, n,
                                              C.byref(self.rp),
                                              self.thr.data_ptr(), self.memimg.data_ptr() + i0 * ncell, None, None,
                                              self.n_rects_host.data_ptr() + i0 * 4,
                                              self.rects_pinned.data_ptr() + i0 * _lib.NSOF_MAX_RECTS * 16, None, st),
                  "roi")

    def stage1(self, frames):
        """All chunks back to back on the current stream (tests, debugging)."""
        for j in range(len(self.chunks)):
            self.stage1_chunk(frames, j, self.s1_eng if not self.overlap else self.eng)

    # ---- stage 2 ---------------------------------------------------------------------------------------
    def build_descs(self, frames, rects, n_rects, pairs=None, n0=0, px0=0):
        """Crop descriptors of the pairs `pairs` = (p0, p1) appended at self.descs[n0:]: views into the resident
        frames and into the flow frames (or the packed buffer from pixel px0 on) -- no copies."""
        H, W, Cn = self.H, self.W, self.C
        fbase = frames.data_ptr()
        obase = self.packed.data_ptr() if self.packed_mode else self.flow.data_ptr()
        p0, p1 = pairs if pairs is not None else (0, self.F - 1)
        n, px = n0, px0
        if self.sep and self.packed_mode:
            return self._build_descs_separate_packed(frames, rects, n_rects, p0, p1, n0, px0)
        if self.sep:
            return self._build_descs_separate(frames, rects, n_rects, p0, p1, n0), px0
        for i in range(p0 * Cn, p1 * Cn):
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
            if self.segment:
                g = self.seg_descs[n]
                g.flow, g.flow_pitch, g.h, g.w = d.flow, d.flow_pitch, d.h, d.w
                if self.packed_mode:
                    g.mask, g.mask_pitch = self.packed_mask.data_ptr() + px, x1 - x0
                else:
                    g.mask, g.mask_pitch = self.motion.data_ptr() + ((t * Cn + c) * H * W + off), W
            n += 1
            px += (y1 - y0) * (x1 - x0)
        return n, px

    def _build_descs_separate(self, frames, rects_all, n_rects, p0, p1, n0):
        """FLAG 1 (process_separate_regions, optical_flow_seg.py:123-166): one Farneback item per connected region, each
        pasted into the pair's flow frame, LATER regions over earlier ones where their rectangles overlap (:159).  The
        descriptors of the chunk are ordered by region rank -- all first regions, then all second regions, ... -- and
        `stage2` runs one batch per rank on the same stream, so the paste order is the stream order.  Segmentation
        (task_results with EST_FLAG 1, MERGE_FLAG True, :265-283): ONE box per pair around all its regions plus PADDING,
        read from the pasted frame.  rects_all: (n_pairs, NSOF_MAX_RECTS, 4) as the ROI kernel stores them."""
        H, W, Cn = self.H, self.W, self.C
        fbase, obase = frames.data_ptr(), self.flow.data_ptr()
        n, self._groups = n0, []
        m0 = m = self.n_seg
        live = []
        for i in range(p0 * Cn, p1 * Cn):
            self.offsets[i] = -1
            k = int(n_rects[i])
            if k > self.max_regions:
                raise MemoryError(f"pair {i}: {k} regions, the pipeline holds {self.max_regions} per pair")
            if k:
                live.append((i, k))
        rank = 0
        while live:
            g0 = n
            for i, k in live:
                x0, y0, x1, y1 = (int(v) for v in rects_all[i, rank])
                if x1 <= x0 or y1 <= y0:  # prev_region.size == 0 guard of optical_flow_seg.py:155
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
            if n > g0:
                self._groups.append((g0, n))
            rank += 1
            live = [(i, k) for i, k in live if k > rank]
        if self.segment:
            pad = int(self.cfg.PADDING)
            for i in range(p0 * Cn, p1 * Cn):
                k = int(n_rects[i])
                if not k:
                    continue
                r = rects_all[i, :k]
                x0, y0 = max(0, int(r[:, 0].min()) - pad), max(0, int(r[:, 1].min()) - pad)
                x1, y1 = min(W, int(r[:, 2].max()) + pad), min(H, int(r[:, 3].max()) + pad)
                if x1 <= x0 or y1 <= y0:
                    continue
                t, c = divmod(i, Cn)
                off = (t * Cn + c) * H * W + y0 * W + x0
                g = self.seg_descs[m]
                g.flow, g.flow_pitch, g.h, g.w = obase + off * 8, W * 8, y1 - y0, x1 - x0
                g.mask, g.mask_pitch = self.motion.data_ptr() + off, W
                m += 1
        self._seg_range, self.n_seg = (m0, m), m
        return n

    def _build_descs_separate_packed(self, frames, rects_all, n_rects, p0, p1, n0, px0):
        """FLAG 1 with packed output: every region's flow crop is stored on its own (region_offsets[i][j], label order), so
        nothing overlaps in memory and the chunk is ONE batch; the consumer pastes in label order if it wants the frame."""
        H, W, Cn = self.H, self.W, self.C
        fbase, obase = frames.data_ptr(), self.packed.data_ptr()
        n, px = n0, px0
        for i in range(p0 * Cn, p1 * Cn):
            self.offsets[i], self.region_offsets[i] = -1, []
            k = int(n_rects[i])
            if k > self.max_regions:
                raise MemoryError(f"pair {i}: {k} regions, the pipeline holds {self.max_regions} per pair")
            t, c = divmod(i, Cn)
            for j in range(k):
                x0, y0, x1, y1 = (int(v) for v in rects_all[i, j])
                if x1 <= x0 or y1 <= y0:
                    self.region_offsets[i].append(-1)
                    continue
                if (px + (y1 - y0) * (x1 - x0)) * 2 > self.packed.numel():
                    raise MemoryError("packed flow buffer too small for this step's regions")
                off = y0 * W + x0
                d = self.descs[n]
                d.prev = fbase + ((t * Cn + c) * H * W + off)
                d.next = fbase + (((t + 1) * Cn + c) * H * W + off)
                d.h, d.w = y1 - y0, x1 - x0
                d.prev_pitch = d.next_pitch = W
                d.flow, d.flow_pitch = obase + px * 8, (x1 - x0) * 8
                self.region_offsets[i].append(px * 2)
                if self.offsets[i] < 0:
                    self.offsets[i] = px * 2
                n += 1
                px += (y1 - y0) * (x1 - x0)
        self._groups, self._seg_range = ([(n0, n)] if n > n0 else []), (0, 0)
        return n, px

    def _descs_at(self, n0):
        return C.cast(C.byref(self.descs, n0 * _DESC_SIZE), C.POINTER(CropDesc))

    def clear_previous(self):
        """S2-g: the reference pastes every crop's flow into a zero frame.  The flow frames belong to this object
        and only stage 2 writes them, so zeroing the previous step's rectangles restores all-zero frames."""
        if self.packed_mode or not self.n_dirty:
            return
        self.lib.check(self.lib.nsof_flow_clear(self.h, self.descs, self.n_dirty, self._stream()), "flow_clear")
        if self.segment and self.n_seg_dirty:  # the same for the mask frames (a whole-frame zero_() of 249 MB cost 0.5 ms per step)
            self.lib.check(self.lib.nsof_mask_clear(self.h, self.seg_descs, self.n_seg_dirty, self._stream()), "mask_clear")
        self.n_dirty = self.n_seg_dirty = 0

    def stage2(self, n0, n1, eng=None):
        if self.sep:   # FLAG 1: one batch per region rank, in stream order; then one mask box per pair
            for a, b in self._groups:
                self.lib.check(self.lib.nsof_farneback_batch((eng or self.eng).handle, self._descs_at(a), b - a,
                                                             C.byref(self.fb), self._stream()), "farneback_batch")
            m0, m1 = self._seg_range
            if self.segment and m1 > m0:
                segs = C.cast(C.byref(self.seg_descs, m0 * C.sizeof(SegDesc)), C.POINTER(SegDesc))
                self.lib.check(self.lib.nsof_flow_segment_batch((eng or self.eng).handle, segs, m1 - m0,
                                                                C.byref(self.seg_params), self._stream()), "flow_segment")
            return
        if n1 > n0:
            self.lib.check(self.lib.nsof_farneback_batch((eng or self.eng).handle, self._descs_at(n0), n1 - n0,
                                                         C.byref(self.fb), self._stream()), "farneback_batch")
            if self.segment:
                segs = C.cast(C.byref(self.seg_descs, n0 * C.sizeof(SegDesc)), C.POINTER(SegDesc))
                self.lib.check(self.lib.nsof_flow_segment_batch((eng or self.eng).handle, segs, n1 - n0,
                                                                C.byref(self.seg_params), self._stream()), "flow_segment")

    def stage2_chunk(self, j, n0, n1, cur):
        """Farneback batch of chunk j on its stream (chunk j -> stream j mod s2_streams, 0 = the caller's);
        records self.ev_s2[j] there."""
        k = j % (1 + len(self.s2_extra))
        if k == 0:
            self.stage2(n0, n1)
            self.ev_s2[j].record(cur)
        else:
            eng, st = self.s2_extra[k - 1]
            with self.torch.cuda.stream(st):
                self.stage2(n0, n1, eng)
                self.ev_s2[j].record(st)

    def step(self, frames):
        """One pass of the hot path over the rank's clips, inputs and outputs resident in HBM."""
        self.begin(frames)
        return self.finish(frames)

    def begin(self, frames):
        """First half of a step, no host wait: clear the previous step's flow rectangles and queue stage 1 of every
        chunk on the side stream (overlap=True).  `DeviceStream` calls this for step k before it finishes step k-1."""
        t = self.torch
        self.clear_previous()
        if self.full_frame or not self.overlap:
            return
        cur = t.cuda.current_stream(self.dev)
        self.side.wait_stream(cur)  # frames (and everything queued so far) are ordered before stage 1
        for _, st in self.s2_extra:
            st.wait_stream(cur)     # ... and before stage 2 on the extra streams (flow_clear included)
        with t.cuda.stream(self.side):
            for j in range(len(self.chunks)):
                self.stage1_chunk(frames, j)
                self.ev_chunk[j].record(self.side)

    def finish(self, frames):
        """Second half: per chunk, wait for its rectangles on the host and queue its Farneback batch."""
        t = self.torch
        if self.full_frame:
            n, px = self.build_descs(frames, None, None)
            self.stage2(0, n)
        else:
            cur = t.cuda.current_stream(self.dev)
            n = px = 0
            self.n_seg = 0
            rects_np = self.rects_pinned.numpy() if self.sep else self.rects_host.numpy()
            for j, pr in enumerate(self.chunks):
                if self.overlap:
                    self.ev_chunk[j].synchronize()  # the chunk's rectangles are on the host
                else:
                    self.stage1_chunk(frames, j, self.eng)
                    cur.synchronize()
                n1, px = self.build_descs(frames, rects_np, self.n_rects_host.numpy(), pr, n, px)
                self.stage2_chunk(j, n, n1, cur)
                n = n1
            for j in range(len(self.chunks)):  # the caller's stream ends the step after every batch
                if j % (1 + len(self.s2_extra)):
                    cur.wait_event(self.ev_s2[j])
            # FLAG 2: (n_pairs, 4); FLAG 1: (n_pairs, max_regions, 4), the first n_rects[i] rows of pair i valid
            self.last_rects = ((rects_np[:, :self.max_regions].copy() if self.sep else rects_np.copy()),
                               self.n_rects_host.numpy().copy())
        self.n_items = n
        self.n_dirty = 0 if self.packed_mode else n
        self.n_seg_dirty = 0 if (self.packed_mode or not self.segment) else (self.n_seg if self.sep else n)
        self.last_crop_pixels = px
        self.packed_floats = px * 2
        return n


class DeviceStream:
    """A stream of device-resident steps through `depth` SequencePipeline slots, ONE Farneback batch per step.

    `SequencePipeline.step` hides the latency-bound synapse chains of stage 1 under stage 2 of the SAME step, which
    forces the step into small time chunks (a Farneback batch needs its rectangles) and the coarse pyramid levels of a
    small batch do not fill the machine.  Here stage 1 of step k runs under the Farneback batch of step k-1 instead:
    every step is one batch of all its pairs (measured 4.0 ms against 4.9 ms chunked for 120 pairs), and the host
    waits for rectangles that have been ready for a whole step.

        submit(frames) -> the pipeline slot holding the results (.flow, .last_rects, .motion) of the PREVIOUS step,
                          valid until `depth - 1` more submits; None for the first
        drain()        -> finishes the step still in flight and returns its slot
    """

    def __init__(self, engine: Engine, clips: int, frames: int, H: int, W: int, cfg: RoiConfig | None = None,
                 depth: int = 2, segment: bool = False, full_frame: bool = False, roi_state: str = "scripts",
                 chunks=None, s2_streams: int = 1):
        """chunks / s2_streams: by default a step is ONE Farneback batch on the caller's stream; `chunks` = pair ranges
        [(p0, p1), ...] splits it into time chunks whose batches alternate over `s2_streams` streams (two half batches
        side by side fill the tails and the barrier-bound phases of each other)."""
        assert depth >= 2
        one = list(chunks) if chunks is not None else [(0, int(frames) - 1)]
        self.slots = [SequencePipeline(engine if i == 0 else Engine(engine.device, lib=engine.lib), clips, frames, H, W, cfg,
                                       chunks=one, overlap=True, s2_streams=s2_streams, segment=segment,
                                       full_frame=full_frame, roi_state=roi_state)
                      for i in range(depth)]
        self.k = 0
        self.pending = None  # (slot, frames) of the step whose stage 1 is queued

    @property
    def n_pairs(self):
        return self.slots[0].n_pairs

    def launch_count(self) -> int:
        return sum(s.launch_count() for s in self.slots)

    def submit(self, frames):
        slot = self.slots[self.k % len(self.slots)]
        self.k += 1
        slot.begin(frames)          # stage 1 of this step: side stream, behind everything queued so far
        done = self.drain()         # Farneback batch of the previous step: runs beside it
        self.pending = (slot, frames)
        return done

    def drain(self):
        if self.pending is None:
            return None
        slot, frames = self.pending
        self.pending = None
        slot.finish(frames)
        return slot


class HostPipeline:
    """Host-buffer entry of the sequence path (bench.py's e2e leg): pinned uint8 clips in, per pair its rectangle
    and its ROI flow crop (float32, packed) out, host<->device copies included.

    The same time-chunked schedule as `SequencePipeline.step`, with a copy stream on either side:

        in    : H2D of chunk j's frames of every clip (strided into the time-major device layout)
        side  : stage 1 of chunk j  (waits for in[j])  -> rectangles D2H
        main  : Farneback batch of chunk j (host waits for the rectangles, enqueues)
        out   : D2H of chunk j's packed flow crops (waits for main[j])

    so both copy engines and the SMs are busy at once and the latency-bound synapse chains hide under stage 2.
    """

    def __init__(self, device: int, clips: int, frames: int, H: int, W: int, cfg: RoiConfig | None = None,
                 slots: int | None = None, full_frame: bool = False, lib=None, chunks=None, copy_streams=None,
                 segment: bool = False, roi_state: str = "scripts", stage1: str = "simulate"):
        import torch
        self.torch = torch
        self.dev = torch.device(f"cuda:{device}")
        self.C, self.F, self.H, self.W = int(clips), int(frames), int(H), int(W)
        self.eng = Engine(device, lib=lib)
        with torch.cuda.device(self.dev):
            self.main = torch.cuda.Stream(self.dev)
            # one H2D and one D2H stream, shared by all slots of a HostStream: the driver binds copy engines to
            # streams, and an H2D stream of one slot that lands on the engine of another slot's D2H stream is
            # serialised with it (measured: 7.6 ms instead of 5.3 ms per step of copies)
            self.s_in, self.s_out = copy_streams or (torch.cuda.Stream(self.dev), torch.cuda.Stream(self.dev))
            if chunks is None:
                chunks = chunk_schedule(self.F - 1, first=1, grow=2, cap=4)
            self.pipe = SequencePipeline(self.eng, clips, frames, H, W, cfg, full_frame=full_frame, packed=True,
                                         chunks=chunks, overlap=True, segment=segment, roi_state=roi_state, stage1=stage1)
            if full_frame:  # no stage 1: the side stream/handle of the pipe is unused
                self.pipe.side = None
            self.frames = torch.empty((self.F, self.C, H, W), dtype=torch.uint8, device=self.dev)
            nchunk = len(self.pipe.chunks)
            self.ev_in = [torch.cuda.Event(enable_timing=True) for _ in range(nchunk)]
            self.ev_d2h = [torch.cuda.Event(enable_timing=True) for _ in range(nchunk)]
            self.ev_start = torch.cuda.Event(enable_timing=True)
            self.ev_out = torch.cuda.Event()
        self.flow_host = torch.empty((self.pipe.packed.numel(),), dtype=torch.float32).pin_memory()
        self.mask_host = (torch.empty((self.pipe.packed_mask.numel(),), dtype=torch.uint8).pin_memory()
                          if segment else None)
        self.slots = ([self.main, self.s_in, self.s_out] + ([self.pipe.side] if self.pipe.side is not None else [])
                      + [st for _, st in self.pipe.s2_extra])
        self.h2d_bytes = self.d2h_bytes = 0

    def launch_count(self) -> int:
        return self.pipe.launch_count()

    def timeline(self):
        """Device-side completion times (ms after the start of the last run) of every chunk's H2D, stage 1
        (rectangles on the host), stage 2 and D2H, plus the host clock when each chunk's stage 2 was enqueued."""
        rows = []
        for j, pr in enumerate(self.pipe.chunks):
            g = lambda e: round(self.ev_start.elapsed_time(e), 3)
            rows.append(dict(chunk=pr, h2d=g(self.ev_in[j]), stage1=None if self.pipe.full_frame else g(self.pipe.ev_chunk[j]),
                             stage2=g(self.pipe.ev_s2[j]), d2h=g(self.ev_d2h[j]),
                             host_enqueued=round((self.host_marks[j + 1] - self.host_marks[0]) * 1e3, 3)))
        return rows

    def _ptr(self, stream):
        return C.c_void_p(stream.cuda_stream)

    def run(self, clips_host):
        """clips_host: sequence of C pinned uint8 tensors (F, H, W).  Returns per clip (rects, n_rects, offsets,
        packed float32 host tensor); pair t's flow crop is packed[offsets[t] : offsets[t] + 2 h w].view(h, w, 2).
        With segment=True a fifth entry, the packed uint8 motion masks: pair t's is
        masks[offsets[t] // 2 : offsets[t] // 2 + h w].view(h, w).  The packed tensors are shared by all clips of
        the step.  FLAG 1: rects is (F-1, max_regions, 4) and offsets[t] a list with one offset per region of pair t (label
        order; segment=True is refused).  Synchronous: results are on the host on return."""
        self.enqueue_input(clips_host)
        self.enqueue_flow()
        return self.collect()

    def enqueue_input(self, clips_host):
        """Phase 1 (no host wait): the H2D copies of every chunk and stage 1 behind them."""
        assert len(clips_host) == self.C
        t, p, lib = self.torch, self.pipe, self.eng.lib
        Cn, F, H, W = self.C, self.F, self.H, self.W
        fsz = H * W
        self.h2d_bytes = self.d2h_bytes = 0
        with t.cuda.device(self.dev):
            self.ev_start.record(self.s_in)
            self.host_marks = [time.perf_counter()]
            # ---- in: all H2D copies, chunk by chunk
            for j, (p0, p1) in enumerate(p.chunks):
                f0 = 0 if j == 0 else p0 + 1
                nf = p1 + 1 - f0
                for c in range(Cn):
                    src = clips_host[c]
                    assert src.dtype == t.uint8 and tuple(src.shape) == (F, H, W) and src.is_contiguous()
                    lib.check(lib.nsof_copy_2d_async(self.frames.data_ptr() + (f0 * Cn + c) * fsz, Cn * fsz,
                                                     src.data_ptr() + f0 * fsz, fsz, fsz, nf, 1, self._ptr(self.s_in)),
                              "copy_2d H2D")
                self.ev_in[j].record(self.s_in)
                self.h2d_bytes += nf * Cn * fsz
            # ---- side: stage 1 of every chunk
            if not p.full_frame:
                with t.cuda.stream(p.side):
                    for j in range(len(p.chunks)):
                        p.side.wait_event(self.ev_in[j])
                        p.stage1_chunk(self.frames, j)
                        p.ev_chunk[j].record(p.side)

    def enqueue_flow(self):
        """Phase 2: per chunk, wait for its rectangles on the host, enqueue its Farneback batch and the D2H copy
        of its packed flow crops."""
        t, p = self.torch, self.pipe
        with t.cuda.device(self.dev):
            n = px = 0
            for j, pr in enumerate(p.chunks):
                if p.full_frame:
                    rects = n_rects = None
                    self.main.wait_event(self.ev_in[j])
                else:
                    p.ev_chunk[j].synchronize()
                    rects, n_rects = (p.rects_pinned if p.sep else p.rects_host).numpy(), p.n_rects_host.numpy()
                n1, px1 = p.build_descs(self.frames, rects, n_rects, pr, n, px)
                with t.cuda.stream(self.main):
                    p.stage2_chunk(j, n, n1, self.main)
                if px1 > px:
                    self.s_out.wait_event(p.ev_s2[j])
                    with t.cuda.stream(self.s_out):
                        self.flow_host[2 * px:2 * px1].copy_(p.packed[2 * px:2 * px1], non_blocking=True)
                        if p.segment:
                            self.mask_host[px:px1].copy_(p.packed_mask[px:px1], non_blocking=True)
                self.ev_d2h[j].record(self.s_out)
                self.host_marks.append(time.perf_counter())
                n, px = n1, px1
            self.ev_out.record(self.s_out)
        self._n, self._px = n, px

    def collect(self):
        """Phase 3: wait for the step's last D2H copy and hand out the results."""
        t, p = self.torch, self.pipe
        Cn, F = self.C, self.F
        n, px = self._n, self._px
        with t.cuda.device(self.dev):
            self.ev_out.synchronize()
            self.main.synchronize()
            for _, st in p.s2_extra:
                st.synchronize()
        p.n_items, p.last_crop_pixels, p.packed_floats = n, px, 2 * px
        self.d2h_bytes = px * (9 if p.segment else 8) + (0 if p.full_frame else p.n_pairs * 20)
        out = []
        if p.full_frame:
            for c in range(Cn):
                out.append((None, None, [p.offsets[tt * Cn + c] for tt in range(F - 1)], self.flow_host)
                           + ((self.mask_host,) if p.segment else ()))
        elif p.sep:  # FLAG 1: per pair up to max_regions rectangles and as many offsets (label order)
            R = p.rects_pinned.numpy()[:, :p.max_regions].reshape(F - 1, Cn, p.max_regions, 4)
            N = p.n_rects_host.numpy().reshape(F - 1, Cn)
            for c in range(Cn):
                out.append((R[:, c].copy(), N[:, c].copy(), [list(p.region_offsets[tt * Cn + c]) for tt in range(F - 1)],
                            self.flow_host))
        else:
            R, N = p.rects_host.numpy().reshape(F - 1, Cn, 4), p.n_rects_host.numpy().reshape(F - 1, Cn)
            for c in range(Cn):
                out.append((R[:, c].copy(), N[:, c].copy(), [p.offsets[tt * Cn + c] for tt in range(F - 1)],
                            self.flow_host) + ((self.mask_host,) if p.segment else ()))
        return out


class HostStream:
    """A stream of steps through `depth` HostPipeline slots (own device buffers, streams and library handles each),
    so the copy engines never wait for a step's tail: while step k's last Farneback batch and D2H copy run, step
    k+1's frames are already arriving and its synapse arrays are integrating.

        submit(clips_host)  -> enqueue the step; returns the results of the step that used this slot `depth`
                               submissions ago (None while the pipeline fills)
        drain()             -> results of the steps still in flight, oldest first

    Order of work per submit: H2D + stage 1 of the new step are queued first (no host wait), then the host turns to
    the PREVIOUS step's rectangles and queues its Farneback batches.  Results are identical to HostPipeline.run.
    The host buffers of a submitted step must stay untouched until its results have been handed out.
    """

    def __init__(self, device: int, clips: int, frames: int, H: int, W: int, cfg: RoiConfig | None = None,
                 depth: int = 3, full_frame: bool = False, lib=None, chunks=None, segment: bool = False,
                 roi_state: str = "scripts"):
        assert depth >= 2
        import torch
        dev = torch.device(f"cuda:{device}")
        if chunks is None:
            # two time chunks: with the next step's copies always queued, the tail a fine schedule shortens is hidden
            # anyway, and under bidirectional PCIe load a few large copies run ~15 % faster than many small ones
            # (scripts/copy_pattern.py: 5.4 ms vs 6.4 ms per step of frames).  Measured per step at 5 / 20 steps:
            # one chunk 7.85 / 6.01 ms, two 6.70 / 6.11, three 6.73 / 6.30, 1-2-4-4-4 9.4 / 6.7.
            n, parts = int(frames) - 1, min(2, int(frames) - 1)
            chunks = [((n * i + parts - 1) // parts, (n * (i + 1) + parts - 1) // parts) for i in range(parts)]
        self.copy_streams = (torch.cuda.Stream(dev), torch.cuda.Stream(dev))
        self.slots = [HostPipeline(device, clips, frames, H, W, cfg, full_frame=full_frame, lib=lib, chunks=chunks,
                                   copy_streams=self.copy_streams, segment=segment, roi_state=roi_state)
                      for _ in range(depth)]
        self.k = 0            # steps submitted
        self.flowed = 0       # steps whose Farneback batches are queued
        self.collected = 0    # steps handed out
        self.h2d_bytes = self.d2h_bytes = 0

    def launch_count(self) -> int:
        return self.slots[0].launch_count()

    def _collect_oldest(self):
        hp = self.slots[self.collected % len(self.slots)]
        out = hp.collect()
        self.h2d_bytes, self.d2h_bytes = hp.h2d_bytes, hp.d2h_bytes
        self.collected += 1
        return out

    def submit(self, clips_host):
        depth, done = len(self.slots), None
        if self.k - self.collected >= depth:     # the slot is still owned by step k - depth (its flow is queued)
            done = self._collect_oldest()
        self.slots[self.k % depth].enqueue_input(clips_host)
        self.k += 1
        while self.flowed < self.k - 1:          # everything but the step just submitted
            self.slots[self.flowed % depth].enqueue_flow()
            self.flowed += 1
        return done

    def _flow_pending(self):
        while self.flowed < self.k:
            self.slots[self.flowed % len(self.slots)].enqueue_flow()
            self.flowed += 1

    def drain(self):
        self._flow_pending()
        out = []
        while self.collected < self.k:
            out.append(self._collect_oldest())
        return out
