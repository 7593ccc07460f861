"""Asynchronous parameter-server mode (reference: 1-ps-cpu trainer — TF_CONFIG cluster with one PS + one worker per
instance, every variable on `/job:ps`, workers pull variables / push gradients over gRPC with NO barrier and the
optimizer applied on the PS without locks, README-EN:14, PDF p35; SURVEY X3/N8, BASELINE.json configs 1 and 4).

B200-box re-design: the "parameter server" is a HOST-RESIDENT parameter store in shared, page-locked memory that every
worker process of the box maps (`/dev/shm` file + cudaHostRegister) — no server process, no gRPC:

    pull    (pull stream) the batch's unique rows are found ON THE GPU (touched_bitmap + union_rowlist: a device-resident
            row list with a device-resident count, so the host never waits for a size), the ids are remapped to list
            positions, and the GPU gathers those rows straight out of the mapped host store (zero-copy) together with
            the MLP weights
    compute (main stream) forward + backward with the fused sm_100a kernels over a COMPACT table (ids = list
            positions), replayed from a CUDA graph; no local optimizer step
    push    (push stream) the listed gradient rows, their ids and the count are stored by a kernel into a ring of
            page-locked host buffers, the MLP gradients follow with one D2H copy; a background applier thread waits for
            that step's event and runs the TF-1 update rule on the store in place, unlocked (Hogwild — the reference's
            async semantics)
    The three stages of consecutive steps overlap: two buffer sets alternate, so pull(k+1) runs under compute(k) and
    push(k-1) under both; the training thread only enqueues work and never synchronises with the GPU. Workers run ahead
    of their applied pushes by at most `staleness` steps (bounded staleness; 0 = apply before the next pull, no overlap).

Only rows present in the batch are updated and regularised (`l2_mode='touched'`): the whole-table L2 sweep of the
reference would mean moving the full table both ways every step (it does: ~15 MB each way).
On CPU-only boxes (BASELINE config 1: gloo, world_size=2, no GPU) the same store is used with the torch backend;
`torch.distributed` (gloo) is used only for start-up ordering — "chief initialises, others start after" (README-EN:62).
"""
from __future__ import annotations

import os
import queue
import threading
from collections import OrderedDict
from typing import Dict, Iterable, List, Optional

import numpy as np
import torch
import torch.distributed as dist

from ..models import DeepFMConfig
from ..models.config import DeepFMConfig as _Cfg
from ..models.oracle import oracle_forward, oracle_grads
from ..models.params import init_params, native_names
from ..optim.tf1 import SLOT_INIT, SLOT_SUFFIXES, TF1Optimizer
from ..utils.auc import StreamingAUC


class SharedParamStore:
    """All model variables + optimizer slots in one shared-memory file mapped by every worker of the box."""

    def __init__(self, cfg: DeepFMConfig, opt: TF1Optimizer, path: str, create: bool, seed: int = 0, pin: bool = False):
        self.cfg, self.opt, self.path = cfg, opt, path
        specs = [(n, s) for n, s, tr in native_names(cfg)]
        nslots = len(SLOT_INIT[opt.name])
        self.offsets: "OrderedDict[str, tuple]" = OrderedDict()
        off = 0
        for n, shape in specs:
            numel = int(np.prod(shape))
            self.offsets[n] = (off, shape)
            off += (numel + 63) // 64 * 64
            for i in range(nslots):
                self.offsets["%s/%s" % (n, SLOT_SUFFIXES[opt.name][i])] = (off, shape)
                off += (numel + 63) // 64 * 64
        self.offsets["__step__"] = (off, (16,))
        off += 64
        self.total = off
        if create:
            with open(path, "wb") as f:
                f.truncate(self.total * 4)
        self.flat = torch.from_file(path, shared=True, size=self.total, dtype=torch.float32)
        self.pinned = False
        self.dev_base = 0  # device-visible alias of the store (0 = not mapped)
        if pin and torch.cuda.is_available():
            try:
                import ctypes
                from ..ops import lib as _nlib
                dp = ctypes.c_void_p(0)
                self.pinned = _nlib.load().dfm_host_register(ctypes.c_void_p(self.flat.data_ptr()),
                                                             ctypes.c_ulonglong(self.total * 4), ctypes.byref(dp)) == 0
                self.dev_base = int(dp.value or 0) if self.pinned else 0
            except Exception:
                self.pinned = False
        if create:
            src = init_params(cfg, seed)
            for n, _ in specs:
                self.view(n).copy_(src[n])
                for i in range(nslots):
                    self.view("%s/%s" % (n, SLOT_SUFFIXES[opt.name][i])).fill_(SLOT_INIT[opt.name][i])
            self.view("__step__").zero_()

    def view(self, name: str) -> torch.Tensor:
        off, shape = self.offsets[name]
        n = int(np.prod(shape))
        return self.flat[off:off + n].view(shape)

    def dev_ptr(self, name: str) -> int:
        """Device-visible address of a variable inside the mapped store (for the zero-copy gather kernel)."""
        return self.dev_base + self.offsets[name][0] * 4 if self.dev_base else 0

    def slots(self, name: str) -> List[torch.Tensor]:
        return [self.view("%s/%s" % (name, s)) for s in SLOT_SUFFIXES[self.opt.name]]

    @property
    def step_counter(self) -> torch.Tensor:
        """int64 view of the first 8 bytes of the step region (a float32 word is exact only up to 2^24 steps)."""
        return self.view("__step__")[:2].view(torch.int64)

    def bump_step(self) -> int:
        c = self.step_counter
        c[0] += 1  # unlocked on purpose; an approximate global step is all async PS ever had
        return int(c[0])

    def close(self, unlink: bool = False):
        if self.pinned:
            try:
                import ctypes
                from ..ops import lib as _nlib
                _nlib.load().dfm_host_unregister(ctypes.c_void_p(self.flat.data_ptr()))
            except Exception:
                pass
        if unlink:
            try:
                os.remove(self.path)
            except OSError:
                pass


class AsyncPSTrainer:
    """Worker side of the async PS. Same surface as Trainer (train_batch / drain / evaluate / predict / state_dict)."""

    backend = "ps"

    def __init__(self, cfg: DeepFMConfig, batch_size: int, optimizer: TF1Optimizer, seed: int, env, backend: str = "auto",
                 staleness: int = 4, store_path: Optional[str] = None):
        self.cfg, self.B, self.opt, self.env = cfg, batch_size, optimizer, env
        self.world, self.rank = env.world_size, env.rank
        self.dev = env.device if env.device is not None else torch.device("cpu")
        self.cuda = self.dev.type == "cuda"
        self.use_fused = self.cuda and backend in ("auto", "fused") and not cfg.batch_norm
        self.staleness = max(0, int(staleness))
        self.comm_name = "host-resident shared store; zero-copy pull / pinned push pipelined under compute; hogwild apply"
        path = store_path or os.path.join("/dev/shm" if os.path.isdir("/dev/shm") else "/tmp",
                                          "dfm_ps_%s.bin" % os.environ.get("MASTER_PORT", str(os.getppid())))
        self.store_path = path
        chief = self.rank == 0
        if chief:
            self.store = SharedParamStore(cfg, optimizer, path, True, seed, pin=self.cuda)
        if self.world > 1 and dist.is_initialized():
            dist.barrier()  # chief initialises, the others map afterwards (README-EN:62)
        if not chief:
            self.store = SharedParamStore(cfg, optimizer, path, False, seed, pin=self.cuda)
        self.names = [n for n, _, _ in native_names(cfg)]
        self.dense_names = [n for n in self.names if n not in ("fm_w", "fm_v")]
        # pushes are applied by the C++ host optimizer unless DFM_PS_NATIVE_APPLY=0 (then: the torch-op reference path)
        self._native_apply = os.environ.get("DFM_PS_NATIVE_APPLY", "1") != "0"
        self._apply_threads = max(1, min(8, (os.cpu_count() or 4) // max(1, self.world)))
        if self._native_apply:
            try:
                from ..data import native as _nio
                _nio.load()
            except Exception:
                self._native_apply = False
        self._q: "queue.Queue" = queue.Queue()
        self._applier = threading.Thread(target=self._apply_loop, daemon=True)
        self._applier.start()
        self._inflight = 0
        self._lock = threading.Condition()
        self._losses: List[float] = []
        self._last_loss = float("nan")
        # where a step's host time goes (seconds, cumulative): issuing the three stages, the applier waiting for a push to
        # land, the applier running the optimizer on the store
        self.prof = {"issue_s": 0.0, "applier_wait_s": 0.0, "apply_s": 0.0, "steps": 0}
        self._local_steps = 0
        self.kernels_per_step = 0
        self.max_observed_staleness = 0
        if self.use_fused:
            self._init_fused()

    # ------------------------------------------------------------------------------------------ fused GPU worker
    def _init_fused(self):
        """Two buffer sets (compact-table engine + bitmap / row list / remapped ids) alternate between consecutive steps;
        a ring of `staleness + 2` page-locked push records feeds the applier thread."""
        from ..models.fused import FusedDeepFM
        from ..ops import kernels as FK
        c = self.cfg
        self.umax = self.B * c.field_size
        N, K, dev = c.feature_size, c.embedding_size, self.dev
        ccfg = _Cfg(field_size=c.field_size, feature_size=self.umax, embedding_size=K, deep_layers=c.deep_layers,
                    dropout=c.dropout, l2_reg=0.0)
        words = FK.bitmap_words(N)
        i32 = dict(device=dev, dtype=torch.int32)
        self._sets = []
        for _ in range(2):
            m = FusedDeepFM(ccfg, self.B, TF1Optimizer("GD", 0.0), 0, dev, use_graph=False, input_slots=1)
            self._sets.append(dict(
                m=m, ids=torch.zeros(self.B, c.field_size, **i32), bm_own=torch.zeros(words, **i32),
                bm_union=torch.zeros(words, **i32), rows=torch.zeros(max(N, 1), **i32), count=torch.zeros(1, **i32),
                prefix=torch.zeros(words, **i32), graph=None, graphs={}, pull_done=torch.cuda.Event(),
                compute_done=torch.cuda.Event(), push_done=None))
        self.model = self._sets[0]["m"]
        L = self.model.layout
        self._dense_lo, self._dense_hi = L.shadow_range[0], L.trainable_numel  # MLP block of the flat buffer (contiguous)
        self.pull_s = torch.cuda.Stream(device=dev)
        self.push_s = torch.cuda.Stream(device=dev)
        pin = lambda shape, dt=torch.float32: torch.zeros(shape, dtype=dt).pin_memory()
        R = (self.staleness + 3) // 2 * 2  # >= staleness + 2 and even: record r always pairs with buffer set r % 2
        self._ring = [dict(gv=pin((self.umax, K)), gw=pin((self.umax,)), rows=pin((self.umax,), torch.int32),
                           count=pin((1,), torch.int32), gdense=pin((self._dense_hi - self._dense_lo,)), gbias=pin((1,)),
                           loss=pin((1,)), free=True) for _ in range(R)]
        self._ring_cv = threading.Condition()
        self._k = 0

    def _fused_issue(self, ids, vals, labels) -> None:
        """Enqueue pull / compute / push of one step; returns without waiting for the GPU."""
        import ctypes
        import time as _time
        _t0 = _time.perf_counter()
        from ..ops import kernels as FK
        from ..ops import lib as _nlib
        lib = _nlib.load()
        c, st = self.cfg, self.store
        k = self._k
        S = self._sets[k % 2]
        m = S["m"]
        r = k % len(self._ring)
        R = self._ring[r]
        with self._ring_cv:  # the applier has finished with this record (always true within the staleness bound)
            while not R["free"]:
                self._ring_cv.wait()
            R["free"] = False
        main = torch.cuda.current_stream(self.dev)
        slot = m.in_slots[0]
        vp = lambda t: ctypes.c_void_p(t.data_ptr())

        def staged(cache: dict, key, stream, body):
            """Run `body` on `stream`: eagerly the first time (then captured into a CUDA graph), as one graph launch after
            that. Every pointer a stage touches is fixed per (buffer set, ring record), so the three stages of a step cost
            three graph launches on the host instead of ~25 Python-issued copies and kernels (measured: 0.8 ms -> below)."""
            g = cache.get(key)
            with torch.cuda.stream(stream):
                if g is not None:
                    g.replay()
                    return
                body()
                if os.environ.get("DFM_PS_GRAPHS", "1") == "0":
                    return
                stream.synchronize()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=stream, capture_error_mode="thread_local"):
                    body()
                cache[key] = g

        def pull_body():
            sp = ctypes.c_void_p(torch.cuda.current_stream(self.dev).cuda_stream)
            FK.touched_bitmap(S["ids"], c.feature_size, S["bm_own"])
            tab = (ctypes.c_void_p * 1)(S["bm_own"].data_ptr())
            _nlib.check(lib.dfm_union_rowlist_prefix(tab, ctypes.c_int(1), ctypes.c_int(1), ctypes.c_int(c.feature_size),
                                                     vp(S["bm_union"]), vp(S["rows"]), vp(S["count"]), vp(S["prefix"]), sp))
            _nlib.check(lib.dfm_ps_remap_ids(vp(S["ids"]), ctypes.c_longlong(self.umax), ctypes.c_int(c.feature_size),
                                             vp(S["bm_union"]), vp(S["prefix"]), ctypes.c_int(self.umax), vp(slot["ids"]),
                                             sp))
            # the GPU reads the touched rows straight out of the mapped host store
            _nlib.check(lib.dfm_ps_pull_list(ctypes.c_void_p(st.dev_ptr("fm_v")), ctypes.c_void_p(st.dev_ptr("fm_w")),
                                             vp(S["rows"]), vp(S["count"]), ctypes.c_int(self.umax),
                                             ctypes.c_int(c.embedding_size), vp(m._w("fm_v")), vp(m._w("fm_w")), sp))
            for n in self.dense_names:
                m._w(n).copy_(st.view(n), non_blocking=True)
            m.refresh_shadow()

        def push_body():
            sp = ctypes.c_void_p(torch.cuda.current_stream(self.dev).cuda_stream)
            _nlib.check(lib.dfm_ps_push_list(vp(m._g("fm_v")), vp(m._g("fm_w")), vp(S["rows"]), vp(S["count"]),
                                             ctypes.c_int(self.umax), ctypes.c_int(c.embedding_size), vp(R["gv"]),
                                             vp(R["gw"]), vp(R["rows"]), vp(R["count"]), sp))
            R["gdense"].copy_(m.G[self._dense_lo:self._dense_hi], non_blocking=True)
            R["gbias"].copy_(m._g("fm_bias"), non_blocking=True)
            R["loss"].copy_(m.loss_sum, non_blocking=True)
            m.G[self._dense_lo:self._dense_hi].zero_()
            m._g("fm_bias").zero_()

        # ---- pull (pull stream): after the previous user of this buffer set has read its inputs and parameters
        self.pull_s.wait_event(S["compute_done"])
        with torch.cuda.stream(self.pull_s):
            S["ids"].copy_(ids.reshape(self.B, c.field_size), non_blocking=True)
            slot["vals"].copy_(vals.reshape(self.B, c.field_size), non_blocking=True)
            slot["labels"].copy_(labels.reshape(-1), non_blocking=True)
        if st.dev_base:
            staged(S["graphs"], "pull", self.pull_s, pull_body)
        else:  # store not device-mapped (registration failed): host gather, needs the list on the host: one synchronisation
            with torch.cuda.stream(self.pull_s):
                FK.touched_bitmap(S["ids"], c.feature_size, S["bm_own"])
                tab = (ctypes.c_void_p * 1)(S["bm_own"].data_ptr())
                sp = ctypes.c_void_p(self.pull_s.cuda_stream)
                _nlib.check(lib.dfm_union_rowlist_prefix(tab, ctypes.c_int(1), ctypes.c_int(1), ctypes.c_int(c.feature_size),
                                                         vp(S["bm_union"]), vp(S["rows"]), vp(S["count"]), vp(S["prefix"]),
                                                         sp))
                _nlib.check(lib.dfm_ps_remap_ids(vp(S["ids"]), ctypes.c_longlong(self.umax), ctypes.c_int(c.feature_size),
                                                 vp(S["bm_union"]), vp(S["prefix"]), ctypes.c_int(self.umax),
                                                 vp(slot["ids"]), sp))
                self.pull_s.synchronize()
                U = int(S["count"].item())
                rows_h = S["rows"][:U].cpu().long()
                m._w("fm_v")[:U].copy_(st.view("fm_v")[rows_h])
                m._w("fm_w")[:U].copy_(st.view("fm_w")[rows_h])
                for n in self.dense_names:
                    m._w(n).copy_(st.view(n))
                m.refresh_shadow()
        S["pull_done"].record(self.pull_s)
        # ---- compute (main stream): fused kernels on the compact table, dropout keyed by the local step
        main.wait_event(S["pull_done"])
        if S["push_done"] is not None:
            main.wait_event(S["push_done"])  # this set's previous gradients have left (and were zeroed)
        m.step_t.fill_(self._local_steps)
        if S["graph"] is None:
            m._train_kernels(slot, with_optimizer=False)  # warm-up (module load, attributes), then capture
            torch.cuda.synchronize(self.dev)
            m.G.zero_()
            torch.cuda.synchronize(self.dev)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, capture_error_mode="thread_local"):
                self.kernels_per_step = m._train_kernels(slot, with_optimizer=False)
            S["graph"] = g
        S["graph"].replay()
        S["compute_done"].record(main)
        # ---- push (push stream): rows + ids + count by kernel into the pinned record, dense gradients with one copy
        self.push_s.wait_event(S["compute_done"])
        staged(S["graphs"], ("push", r), self.push_s, push_body)
        ev = torch.cuda.Event()
        ev.record(self.push_s)
        S["push_done"] = ev
        self._k += 1
        with self._lock:
            self._inflight += 1
        self._q.put(("ring", r, ev))
        self.prof["issue_s"] += _time.perf_counter() - _t0
        self.prof["steps"] += 1

    def _apply_ring_native(self, r: int) -> float:
        """Applier fast path: the whole pushed step in ONE native call (csrc/io/host_optim.cc dfmio_ps_apply_push)."""
        import ctypes
        from ..data import native as _nio
        from ..optim.tf1 import OPT_IDS
        R, st, o = self._ring[r], self.store, self.opt
        args = R.get("native_args")
        if args is None:  # pointer tables are fixed per ring record: build them once
            L = self.model.layout
            ptr = lambda t: ctypes.c_void_p(t.data_ptr() if t is not None else 0)
            sl = lambda name, i: (st.slots(name)[i] if len(st.slots(name)) > i else None)
            names = self.dense_names
            arr = lambda xs: (ctypes.c_void_p * len(xs))(*[x.data_ptr() if x is not None else 0 for x in xs])
            gs = []
            for n in names:
                if n == "fm_bias":
                    gs.append(R["gbias"])
                else:
                    sg = L.segments[n]
                    gs.append(R["gdense"][sg.offset - self._dense_lo: sg.offset - self._dense_lo + sg.numel])
            keep = ([st.view(n) for n in names], [sl(n, 0) for n in names], [sl(n, 1) for n in names], gs)
            numel = (ctypes.c_longlong * len(names))(*[int(st.view(n).numel()) for n in names])
            tabs = (arr(keep[0]), arr(keep[1]), arr(keep[2]), arr(keep[3]))
            args = R["native_args"] = dict(
                head=(ptr(st.view("fm_v")), ptr(sl("fm_v", 0)), ptr(sl("fm_v", 1)), ptr(st.view("fm_w")), ptr(sl("fm_w", 0)),
                      ptr(sl("fm_w", 1)), ctypes.c_longlong(self.cfg.feature_size), ctypes.c_int(self.cfg.embedding_size),
                      ptr(R["rows"])),
                mid=(ptr(R["gv"]), ptr(R["gw"]), ctypes.c_int(len(names)), tabs[0], tabs[1], tabs[2], tabs[3], numel),
                keep=(keep, tabs))
        t = st.bump_step()
        U = int(R["count"][0])
        rc = _nio.load().dfmio_ps_apply_push(*args["head"], ctypes.c_longlong(U), *args["mid"], ctypes.c_int(OPT_IDS[o.name]),
                                             o.lr, o.beta1, o.beta2, o.eps, o.momentum, float(self.cfg.l2_reg), t,
                                             self._apply_threads)
        if rc != 0:
            raise RuntimeError("ps apply: bad arguments (row index out of range?)")
        return float(R["loss"][0]) / self.B

    def _ring_to_push(self, r: int) -> dict:
        """Applier side: the record's event has completed; view its contents as a push."""
        R, L = self._ring[r], self.model.layout
        U = int(R["count"][0])
        dense = {}
        for n in self.dense_names:
            if n == "fm_bias":
                dense[n] = R["gbias"]
            else:
                sg = L.segments[n]
                o = sg.offset - self._dense_lo
                dense[n] = R["gdense"][o:o + sg.numel].view(sg.shape)
        return {"rows": R["rows"][:U], "fm_v": R["gv"][:U], "fm_w": R["gw"][:U], "dense": dense,
                "loss": float(R["loss"][0]) / self.B}

    # ------------------------------------------------------------------------------------------ torch (CPU) worker
    def _torch_step(self, ids, vals, labels):
        st, cfg = self.store, self.cfg
        ids = ids.long()
        u, inv = torch.unique(ids, return_inverse=True)
        p = {"fm_w": st.view("fm_w")[u].clone(), "fm_v": st.view("fm_v")[u].clone()}  # pull touched rows
        for n in self.dense_names:
            p[n] = st.view(n).clone()                                                    # pull dense variables
        ccfg = _Cfg(field_size=cfg.field_size, feature_size=int(u.numel()), embedding_size=cfg.embedding_size,
                    deep_layers=cfg.deep_layers, dropout=cfg.dropout, l2_reg=0.0, batch_norm=cfg.batch_norm,
                    batch_norm_decay=cfg.batch_norm_decay)
        r = oracle_grads(p, ccfg, inv.reshape(ids.shape), vals, labels, True, 0, self._local_steps, 0)
        g = r["grads"]
        push = {"rows": u, "fm_v": g["fm_v"], "fm_w": g["fm_w"], "dense": {n: g[n] for n in self.dense_names if n in g},
                "bn": r["bn_updates"]}
        return float(r["data_loss"]), push

    # ------------------------------------------------------------------------------------------ push application
    def _apply_loop(self):
        while True:
            item = self._q.get()
            if item is None:
                return
            ring_slot = None
            try:
                import time as _time
                t0 = _time.perf_counter()
                if isinstance(item, tuple):  # ("ring", slot, event): a pipelined GPU step — wait for ITS push only
                    _, ring_slot, ev = item
                    ev.synchronize()
                    if self._native_apply:
                        t1 = _time.perf_counter()
                        loss = self._apply_ring_native(ring_slot)
                        self._losses.append(loss)
                        self._last_loss = loss
                        self.prof["applier_wait_s"] += t1 - t0
                        self.prof["apply_s"] += _time.perf_counter() - t1
                        continue
                    item = self._ring_to_push(ring_slot)
                    self._losses.append(item["loss"])
                    self._last_loss = item["loss"]
                t1 = _time.perf_counter()
                self._apply(item)
                self.prof["applier_wait_s"] += t1 - t0
                self.prof["apply_s"] += _time.perf_counter() - t1
            except Exception as e:  # keep the worker alive; the error surfaces at the next drain()
                self._apply_error = e
            finally:
                if ring_slot is not None:
                    with self._ring_cv:
                        self._ring[ring_slot]["free"] = True
                        self._ring_cv.notify_all()
                with self._lock:
                    self._inflight -= 1
                    self._lock.notify_all()

    def _apply(self, push):
        st, o, l2 = self.store, self.opt, self.cfg.l2_reg
        t = st.bump_step()
        rows = push["rows"]
        if self._native_apply:  # one fused pass per row / per variable in C++ (csrc/io/host_optim.cc)
            from ..data import native as _nio
            for name in ("fm_w", "fm_v"):
                _nio.sparse_opt_apply(st.view(name), st.slots(name), rows, push[name], o, l2, t, self._apply_threads)
            for name, g in push["dense"].items():
                _nio.dense_opt_apply(st.view(name), st.slots(name), g, o, 0.0, t, self._apply_threads)
            for name, v in push.get("bn", {}).items():
                st.view(name).copy_(v)
            return
        for name in ("fm_w", "fm_v"):
            tab, slots = st.view(name), st.slots(name)
            w = tab[rows]
            g = push[name] + l2 * w                     # touched-row L2
            sl = [s[rows] for s in slots]
            o.apply_dense(w, g, sl, t)
            tab[rows] = w                               # unlocked scatter: Hogwild
            for s, v in zip(slots, sl):
                s[rows] = v
        for name, g in push["dense"].items():
            o.apply_dense(st.view(name), g, st.slots(name), t)
        for name, v in push.get("bn", {}).items():
            st.view(name).copy_(v)

    # ------------------------------------------------------------------------------------------ Trainer surface
    def train_batch(self, ids, vals, labels):
        with self._lock:  # bounded staleness: never more than `staleness` un-applied pushes from this worker
            while self._inflight > self.staleness:
                self._lock.wait()
            self.max_observed_staleness = max(self.max_observed_staleness, self._inflight)
        if self.use_fused:
            self._fused_issue(ids, vals, labels)  # enqueues pull / compute / push; the applier collects the loss
            self._local_steps += 1
            return None
        loss, push = self._torch_step(ids, vals, labels)
        with self._lock:
            self._inflight += 1
        self._q.put(push)
        self._losses.append(loss)
        self._last_loss = loss
        self._local_steps += 1
        return None

    train_batch_device = train_batch

    def train_packed(self, packed):
        """One step on a batch in the loader's packed slot layout (labels | ids | vals, 256-byte aligned parts)."""
        B, F = self.B, self.cfg.field_size
        al = lambda x: (x + 255) & ~255
        o_ids = al(B * 4)
        o_vals = o_ids + al(B * F * 4)
        labels = packed[0:B * 4].view(torch.float32)
        ids = packed[o_ids:o_ids + B * F * 4].view(torch.int32).view(B, F)
        vals = packed[o_vals:o_vals + B * F * 4].view(torch.float32).view(B, F)
        return self.train_batch(ids, vals, labels)

    def drain(self, keep: bool = False):
        with self._lock:
            while self._inflight > 0:
                self._lock.wait()
        err = getattr(self, "_apply_error", None)
        if err is not None:
            self._apply_error = None
            raise err
        if keep:
            return []
        out, self._losses = self._losses, []
        return out

    def last_loss(self):
        return torch.tensor(self._losses[-1] if self._losses else self._last_loss)

    @property
    def global_step(self) -> int:
        return int(self.store.step_counter[0])

    def _params(self) -> Dict[str, torch.Tensor]:
        return OrderedDict((n, self.store.view(n)) for n in self.names)

    def parameter_names(self):
        return list(self.names)

    def _fused_scorer(self):
        """Full-table fused forward engine for evaluate / predict on a GPU worker (serving.Predictor: graph-captured,
        batch-size buckets), refreshed from the store on every call."""
        from ..serving.predictor import Predictor
        params = OrderedDict((n, self.store.view(n).clone()) for n in self.names)
        sc = getattr(self, "_scorer", None)
        if sc is None:
            sc = self._scorer = Predictor(self.cfg, params, max_batch=max(self.B, 1024), backend="fused")
        else:
            sc.set_params(params)
        return sc

    @torch.no_grad()
    def evaluate(self, batches: Iterable, distributed: bool = False):
        self.drain(keep=True)
        if self.use_fused:
            sc = self._fused_scorer()
            auc, tot, n = StreamingAUC(), 0.0, 0
            for ids, vals, labels in batches:
                p = sc.predict({"feat_ids": ids, "feat_vals": vals})["prob"].double().clamp(1e-7, 1 - 1e-7)
                lab = labels.double().reshape(-1).cpu()
                tot += float(-(lab * p.log() + (1 - lab) * (1 - p).log()).sum())
                auc.update(labels.cpu(), p.float())
                n += ids.shape[0]
            return {"loss": tot / max(n, 1), "auc": auc.result(), "global_step": self.global_step, "examples": n}
        p = self._params()
        auc, tot, n = StreamingAUC(), 0.0, 0
        for ids, vals, labels in batches:
            y = oracle_forward(p, self.cfg, ids, vals, False)
            tot += float(torch.nn.functional.binary_cross_entropy_with_logits(y, labels.float().reshape(-1), reduction="sum"))
            auc.update(labels, torch.sigmoid(y))
            n += ids.shape[0]
        return {"loss": tot / max(n, 1), "auc": auc.result(), "global_step": self.global_step, "examples": n}

    @torch.no_grad()
    def predict(self, batches: Iterable):
        if self.use_fused:
            sc = self._fused_scorer()
            for b in batches:
                yield sc.predict({"feat_ids": b[0], "feat_vals": b[1]})["prob"].float()
            return
        p = self._params()
        for b in batches:
            yield torch.sigmoid(oracle_forward(p, self.cfg, b[0], b[1], False)).float()

    def state_dict(self):
        self.drain(keep=True)
        sd = OrderedDict()
        for n in self.names:
            sd[n] = self.store.view(n).clone()
            for s, t in zip(SLOT_SUFFIXES[self.opt.name], self.store.slots(n)):
                sd["%s/%s" % (n, s)] = t.clone()
        sd["global_step"] = torch.tensor([self.global_step], dtype=torch.int64)
        return sd

    def load_state_dict(self, sd, strict: bool = True):
        for k, v in sd.items():
            if k == "global_step":
                self.store.step_counter[0] = int(v.reshape(-1)[0])
            elif k in self.store.offsets:
                self.store.view(k).copy_(v)
        self.opt.step = self.global_step

    def close(self):
        self.drain(keep=True)
        self._q.put(None)
        if self.world > 1 and dist.is_initialized():
            dist.barrier()
        self.store.close(unlink=self.rank == 0)
