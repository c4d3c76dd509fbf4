"""Spatially tiled Farneback: one frame pair (or a batch of pairs) split by ROWS over the GPUs of one box
(BASELINE.json configs[4]: 8K frames, NCCL halo-row exchange per pyramid level).

The reference has no multi-GPU path (its Farneback is cv2's single-threaded CPU code, optical_flow_seg.py:494);
this is the B200 design for frames too large to be worth one GPU each.  One process per GPU.  Every rank plans the
WHOLE frame (full-size workspace, so addressing, image borders and resize geometry are those of the whole frame)
but computes only its band of rows of every pyramid level (`nsof_fb_tiled_*`, include/nsof.h).  What a stencil or
a gather reads beyond the band arrives from the row neighbours, in place, between the stages:

    per level, coarse to fine                                      rows per neighbour      bytes at 8K level 0
      pyramid + PolyExp of the band (local: the source frames are read with their own overlap)
      R1 halo                      <- UpdateMatrices gathers at y + dy    halo_r1 (32) x 5 planes    4.9 MB
      coarse flow halo             <- bilinear flow upsampling            4 rows x float2
      first UpdateMatrices
      M halo                       <- 15 x 15 box window                  winsize/2 (7) x 5 planes   1.1 MB
      iterations x [blur + solve + UpdateMatrices ; M halo]

The halo rows of all pairs and planes of one exchange travel as ONE message per neighbour (gathered and scattered
by a copy kernel on either side: per-plane messages made the path latency-bound on operation count).  They go through
torch.distributed point-to-point operations (NCCL send/recv over NVLink between row neighbours; gloo in the CPU
tests, where the same driver runs on the block-emulator build of the kernels).  The result is bit-identical to the
single-GPU result: per-pixel arithmetic does not depend on the tile decomposition.

A displacement larger than `halo_r1` rows at some level would gather rows that were never received; the kernels
flag that (nsof_fb_tiled_guard) and `run` raises instead of returning a wrong field.

`run_linked` is the same decomposition over PEER MEMORY (NVLink): the row neighbours' workspaces are mapped into
this process once (CUDA IPC handles exchanged through torch.distributed -- plumbing), and ONE library call
(nsof_fb_tiled_run) enqueues the whole level loop; after every stage one kernel stores the band's edge rows straight
into the neighbours' halo rows and raises their flags, and a one-thread kernel waits for this rank's own flags.  No
message buffers, no pack / unpack kernels, no NCCL point-to-point operations, no host round trips between stages.
`run` (NCCL / gloo messages) remains as the portable path and as the cross-check.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from ._lib import CopyBlock, CropDesc, TiledBuf
from .engine import FARNEBACK_PARAMS, Engine, _fb_params

WHAT_R1, WHAT_MA, WHAT_MB, WHAT_FLOW = 0, 1, 2, 3
FLOW_HALO_ROWS = 4


def band_rows(h: int, rank: int, world: int):
    """Rows [lo, hi) of a level of height h owned by `rank` (the same split the library uses)."""
    return rank * h // world, (rank + 1) * h // world


class _DevArray:
    """float32 view of foreign device memory for torch.as_tensor (CUDA array interface v2)."""

    def __init__(self, ptr: int, n: int):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": "<f4", "data": (int(ptr), False), "version": 2}


class TiledFarneback:
    def __init__(self, engine: Engine, rank: int, world: int, halo_r1: int = 32, group=None, overlap: bool = False):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.eng, self.lib, self.h = engine, engine.lib, engine.handle
        self.rank, self.world, self.halo_r1 = int(rank), int(world), int(halo_r1)
        self.group = group
        self.emulated = bool(getattr(engine, "_emulated", False))
        self.messages = 0       # point-to-point operations of the last run (sends + receives of this rank)
        self.bytes_sent = 0
        self._ops: dict = {}    # (exchange, workspace layout) -> (operation list, bytes sent)
        # overlap: on levels whose band is at least 4 edges high every stage runs as edges -> [exchange on a second
        # stream || interior]; the rows travel while the bulk of the band is computed (nsof_fb_tiled_stage_rows).
        # Bit-identical, but measured no faster on 2 and 8 B200 (8K, 8 pairs: 1 117 vs 1 182 pairs/s at 8 GPUs): the
        # step is bound by the HOST submitting ~190 launches and 20 exchange rounds, not by the rows in flight, and
        # the split triples the launches.  Off by default.
        self.overlap = bool(overlap)
        self.comm = None if self.emulated else torch.cuda.Stream(torch.device(f"cuda:{engine.device}"))
        self.split_levels = 0   # levels of the last run that took the overlapped schedule

    # ---- plumbing ----------------------------------------------------------------------------------------
    def _stream(self):
        if self.emulated:
            return C.c_void_p(0)
        return C.c_void_p(self.torch.cuda.current_stream(self.eng.device).cuda_stream)

    def _tensor(self, ptr: int, n: int):
        if self.emulated:
            return self.torch.from_numpy(np.ctypeslib.as_array((C.c_float * n).from_address(ptr)))
        return self.torch.as_tensor(_DevArray(ptr, n), device=f"cuda:{self.eng.device}")

    def _query(self, item, level, what) -> TiledBuf:
        b = TiledBuf()
        self.lib.check(self.lib.nsof_fb_tiled_query(self.h, item, level, what, C.byref(b)), "fb_tiled_query")
        return b

    def _exchange(self, n_items, level, what, rows):
        """Send the first / last `rows` rows of the band to the upper / lower neighbour, receive theirs into the
        rows just outside the band.  The row blocks of all items and planes are gathered into ONE message per
        neighbour by one copy kernel (nsof_copy_blocks), exchanged with a single grouped send/recv, and scattered
        into the halo rows by a second copy kernel.  The plan of an exchange (block tables, message buffers,
        operation list) is built once per workspace layout."""
        if self.world == 1:
            return
        dist, torch = self.dist, self.torch
        # the cached block tables hold absolute addresses of EVERY item: key the plan on all of them
        key = (n_items, level, what, rows) + tuple((q.ptr, q.h, q.w, q.row_floats) for q in
                                                   (self._query(i, level, what) for i in range(n_items)))
        plan = self._ops.get(key)
        if plan is None:
            up, dn = self.rank > 0, self.rank < self.world - 1
            per_dir = 0
            blocks = []  # (band rows to send up, halo rows above, band rows to send down, halo rows below, bytes)
            for item in range(n_items):
                b = self._query(item, level, what)
                r = min(rows, b.y_hi - b.y_lo)
                nbytes = 4 * r * b.row_floats
                for p in range(b.planes):
                    row = lambda y: b.ptr + 4 * (p * b.plane_floats + y * b.row_floats)
                    blocks.append((row(b.y_lo), row(b.y_lo - r), row(b.y_hi - r), row(b.y_hi), nbytes))
                    per_dir += nbytes
            mk = lambda: (np.zeros(per_dir // 4, dtype=np.float32) if self.emulated
                          else torch.empty(per_dir // 4, dtype=torch.float32, device=f"cuda:{self.eng.device}"))
            bufs = {k: mk() for k, use in (("su", up), ("ru", up), ("sd", dn), ("rd", dn)) if use}
            addr = lambda t: t.ctypes.data if self.emulated else t.data_ptr()
            as_t = lambda t: torch.from_numpy(t) if self.emulated else t
            pack, unpack, off = [], [], 0
            for s_up, h_up, s_dn, h_dn, nbytes in blocks:
                if up:
                    pack.append((addr(bufs["su"]) + off, s_up, nbytes))
                    unpack.append((h_up, addr(bufs["ru"]) + off, nbytes))
                if dn:
                    pack.append((addr(bufs["sd"]) + off, s_dn, nbytes))
                    unpack.append((h_dn, addr(bufs["rd"]) + off, nbytes))
                off += nbytes
            def tab(rows_):
                """Block table of one copy kernel: host array in the emulator, resident device tensor on CUDA
                (uploaded once: no staging copy and no upload kernel per exchange)."""
                arr = (CopyBlock * len(rows_))(*[CopyBlock(*r) for r in rows_])
                if self.emulated or not rows_:
                    return arr, None
                host = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8)
                return arr, host.to(f"cuda:{self.eng.device}")
            ops = []
            if up:
                ops += [dist.P2POp(dist.isend, as_t(bufs["su"]), self.rank - 1, self.group),
                        dist.P2POp(dist.irecv, as_t(bufs["ru"]), self.rank - 1, self.group)]
            if dn:
                ops += [dist.P2POp(dist.isend, as_t(bufs["sd"]), self.rank + 1, self.group),
                        dist.P2POp(dist.irecv, as_t(bufs["rd"]), self.rank + 1, self.group)]
            if len(self._ops) > 256:
                self._ops.clear()
            big = max((b[2] for b in pack), default=0)
            plan = self._ops[key] = (tab(pack), len(pack), tab(unpack), len(unpack), ops, per_dir * (int(up) + int(dn)), bufs,
                                     big, float(sum(b[2] for b in pack)))
        (pack, pack_dev), n_pack, (unpack, unpack_dev), n_unpack, ops, sent, _, big, total = plan
        st = self._stream()

        def copy(host_tab, dev_tab, cnt, what_):
            if dev_tab is None:
                self.lib.check(self.lib.nsof_copy_blocks(self.h, host_tab, cnt, st), what_)
            else:
                self.lib.check(self.lib.nsof_copy_blocks_resident(self.h, dev_tab.data_ptr(), cnt, big, total, st), what_)

        copy(pack, pack_dev, n_pack, "copy_blocks (pack)")
        for req in dist.batch_isend_irecv(ops):
            req.wait()
        copy(unpack, unpack_dev, n_unpack, "copy_blocks (unpack)")
        self.messages += len(ops)
        self.bytes_sent += sent

    # ---- the level loop of FarnebackOpticalFlowImpl::calc, stage by stage -----------------------------------
    def run(self, crops, **farneback_params):
        """`crops`: list of (prev_ptr, next_ptr, h, w, prev_pitch, next_pitch, flow_ptr, flow_pitch) with raw
        device addresses of FULL frames (all pairs the same size).  On return (asynchronous on the current
        stream, apart from the guard check) rows band_rows(h, rank, world) of every flow buffer hold this rank's
        share of the flow."""
        kw = dict(FARNEBACK_PARAMS)
        kw.update(farneback_params)
        fb = _fb_params(**kw)
        n = len(crops)
        descs = (CropDesc * max(n, 1))()
        for i, c in enumerate(crops):
            descs[i] = CropDesc(*c)
        lib, h, st = self.lib, self.h, self._stream()
        lib.check(lib.nsof_fb_tiled_begin(h, descs, n, C.byref(fb), self.rank, self.world, self.halo_r1, st),
                  "fb_tiled_begin")
        self._n, self._kw = n, kw
        levels = self._stages()
        self._check_guard()
        return levels

    def _stages(self):
        """Every stage and exchange of the batch planned by the last nsof_fb_tiled_begin, on the current stream."""
        lib, h = self.lib, self.h
        n, kw = self._n, self._kw
        self.messages = self.bytes_sent = 0
        levels = lib.nsof_fb_tiled_levels(h)
        m, T = kw["winsize"] // 2, kw["iterations"]
        torch = self.torch
        main = None if self.emulated else torch.cuda.current_stream(self.eng.device)
        pending = [None]  # event of the last exchange queued on the second stream

        def wait_halo():
            if pending[0] is not None:
                main.wait_event(pending[0])
                pending[0] = None

        def stage(k, s, it, exchanges, split):
            """One stage of level k followed by its halo exchanges [(level, what, rows), ...]."""
            run = lambda part: lib.check(lib.nsof_fb_tiled_stage_rows(h, k, s, it, part, self._stream()), "fb_tiled_stage")
            wait_halo()
            if not split:
                run(0)
                for lv, what, rows in exchanges:
                    self._exchange(n, lv, what, rows)
                return
            run(1)
            run(2)
            if self.emulated or not exchanges:
                for lv, what, rows in exchanges:
                    self._exchange(n, lv, what, rows)
            else:
                ev = torch.cuda.Event()
                ev.record(main)
                with torch.cuda.stream(self.comm):
                    self.comm.wait_event(ev)
                    for lv, what, rows in exchanges:
                        self._exchange(n, lv, what, rows)
                    done = torch.cuda.Event()
                    done.record(self.comm)
                pending[0] = done
            run(3)

        self.split_levels = 0
        for k in range(levels - 1, -1, -1):
            split = self.overlap and self.world > 1 and all(
                (lambda b: b.y_hi - b.y_lo >= 4 * self.halo_r1)(self._query(i, k, WHAT_R1)) for i in range(n))
            self.split_levels += int(split)
            stage(k, 0, 0, [(k, WHAT_R1, self.halo_r1)], split)
            stage(k, 1, 0, [(k, WHAT_MA, m)], split)
            for it in range(T):
                if it < T - 1:
                    ex = [(k, WHAT_MB if it % 2 == 0 else WHAT_MA, m)]
                else:  # the level's flow: the next finer level upsamples it (bilinear: rows beyond the band)
                    ex = [(k, WHAT_FLOW, FLOW_HALO_ROWS)] if k > 0 else []
                stage(k, 2, it, ex, split)
        wait_halo()
        return levels

    # ---- the same over peer memory ---------------------------------------------------------------------------
    def _link(self, peers=None):
        """(Re-)attach the row neighbours when this rank's workspace is new.  `peers`: for ranks that share a process
        (tests on the emulator, several GPUs driven by one process) a dict rank -> TiledFarneback; otherwise the IPC
        handles travel through torch.distributed.all_gather_object."""
        lib, h = self.lib, self.h
        mine = (self._export[2], self._export[4])
        if peers is not None:   # one process: raw addresses; every rank re-attaches when ANY workspace is new
            key = tuple((r, peers[r]._export[2], peers[r]._export[3]) for r in sorted(peers))
            if getattr(self, "_linked", None) == (mine, key):
                return
            up, dn = peers.get(self.rank - 1), peers.get(self.rank + 1)
            lib.check(lib.nsof_fb_tiled_link_attach(h, up._export[2] if up else None, up._export[3] if up else None,
                                                    dn._export[2] if dn else None, dn._export[3] if dn else None, 0),
                      "fb_tiled_link_attach")
            self._linked = (mine, key)
            return
        # several processes: has any rank a new workspace?  (cheap: one small object collective per run)
        changed = [None] * self.world
        self.dist.all_gather_object(changed, getattr(self, "_linked", None) != mine, group=self.group)
        if not any(changed):
            return
        table = [None] * self.world
        self.dist.all_gather_object(table, self._export[:2], group=self.group)
        up = table[self.rank - 1] if self.rank > 0 else (None, None)
        dn = table[self.rank + 1] if self.rank < self.world - 1 else (None, None)
        buf = lambda b: (C.c_ubyte * 64).from_buffer_copy(b) if b is not None else None
        lib.check(lib.nsof_fb_tiled_link_attach(h, buf(up[0]), buf(up[1]), buf(dn[0]), buf(dn[1]), 1), "fb_tiled_link_attach")
        self._linked = mine

    def begin_linked(self, crops, **farneback_params):
        """Plan the batch (nsof_fb_tiled_begin) and export this rank's workspace; `run_linked` = begin + link + go."""
        kw = dict(FARNEBACK_PARAMS)
        kw.update(farneback_params)
        fb = _fb_params(**kw)
        n = len(crops)
        descs = (CropDesc * max(n, 1))()
        for i, c in enumerate(crops):
            descs[i] = CropDesc(*c)
        self.lib.check(self.lib.nsof_fb_tiled_begin(self.h, descs, n, C.byref(fb), self.rank, self.world, self.halo_r1,
                                                    self._stream()), "fb_tiled_begin")
        self._n, self._kw = n, kw
        a_ipc, f_ipc = (C.c_ubyte * 64)(), (C.c_ubyte * 64)()
        a_ptr, f_ptr, nbytes = C.c_void_p(), C.c_void_p(), C.c_size_t()
        self.lib.check(self.lib.nsof_fb_tiled_link_export(self.h, a_ipc, f_ipc, C.byref(a_ptr), C.byref(f_ptr),
                                                          C.byref(nbytes)), "fb_tiled_link_export")
        self._export = (bytes(a_ipc), bytes(f_ipc), a_ptr.value, f_ptr.value, int(nbytes.value))

    def go_linked(self):
        nxt = C.c_ulonglong(0)
        self.lib.check(self.lib.nsof_fb_tiled_run(self.h, FLOW_HALO_ROWS, getattr(self, "_seq", 0), C.byref(nxt), self._stream()),
                       "fb_tiled_run")
        self._seq = int(nxt.value)

    def run_linked(self, crops, check_guard: bool = True, **farneback_params):
        """`run` over peer memory: same arguments, same (bit-identical) result, one library call for the level loop.
        Every rank must call it with the same batch geometry, in the same order."""
        self.begin_linked(crops, **farneback_params)
        if self.world > 1:
            self._link()
        self.go_linked()
        if check_guard:
            self._check_guard()
        return self.lib.nsof_fb_tiled_levels(self.h)

    def _check_guard(self):
        lib, h, st = self.lib, self.h, self._stream()
        tripped = C.c_int(0)
        lib.check(lib.nsof_fb_tiled_guard(h, C.byref(tripped), st), "fb_tiled_guard")
        flag = int(tripped.value)
        if self.world > 1:  # every rank must reach the same verdict
            t = self.torch.tensor([flag], dtype=self.torch.int32,
                                  device="cpu" if self.emulated else f"cuda:{self.eng.device}")
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX, group=self.group)
            flag = int(t.item())
        if flag:
            raise RuntimeError(f"tiled Farneback: a displacement exceeded the R1 halo of {self.halo_r1} rows on some "
                               f"rank; rerun with a larger halo_r1")
