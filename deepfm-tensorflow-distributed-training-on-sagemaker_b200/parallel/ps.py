"""Asynchronous parameter-server mode (reference: 1-ps-cpu trainer — TF_CONFIG cluster with one PS + one worker per
instance, every variable on `/job:ps`, workers pull variables / push gradients over gRPC with NO barrier and the
optimizer applied on the PS without locks, README-EN:14, PDF p35; SURVEY X3/N8, BASELINE.json configs 1 and 4).

B200-box re-design: the "parameter server" is a HOST-RESIDENT parameter store in shared, page-locked memory that every
worker process of the box maps (`/dev/shm` file + cudaHostRegister) — no server process, no gRPC:

    pull   worker gathers the rows its batch touches (unique ids) and the MLP weights from the store into a pinned
           staging buffer and copies them to its GPU on a side stream (H2D)
    compute forward + backward on the GPU with the fused sm_100a kernels over a COMPACT table (ids remapped to
           0..U-1), captured in a CUDA graph; no local optimizer step
    push   gradient rows + MLP gradients come back on the side stream (D2H, pinned) and a background applier thread
           runs the TF-1 update rule on the store rows in place, unlocked (Hogwild — the reference's async semantics).
           Workers run ahead of their pushes by at most `staleness` steps (bounded staleness; 0 = apply before next pull).

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
        return self.view("__step__")

    def bump_step(self) -> int:
        c = self.step_counter
        c[0] += 1.0  # unlocked on purpose; an approximate global step is all async PS ever had
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
        self.comm_name = "host-resident shared store; pinned pull/push on a side stream; hogwild apply"
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
        self._local_steps = 0
        self.kernels_per_step = 0
        self.max_observed_staleness = 0
        if self.use_fused:
            self._init_fused()

    # ------------------------------------------------------------------------------------------ fused GPU worker
    def _init_fused(self):
        from ..models.fused import FusedDeepFM
        c = self.cfg
        self.umax = self.B * c.field_size
        ccfg = _Cfg(field_size=c.field_size, feature_size=self.umax, embedding_size=c.embedding_size,
                    deep_layers=c.deep_layers, dropout=c.dropout, l2_reg=0.0)
        self.model = FusedDeepFM(ccfg, self.B, TF1Optimizer("GD", 0.0), 0, self.dev, use_graph=False, input_slots=1)
        self.side = torch.cuda.Stream(device=self.dev)
        K = c.embedding_size
        pin = lambda *s: torch.zeros(*s, dtype=torch.float32).pin_memory()
        self.h_rows, self.h_w = pin(self.umax, K), pin(self.umax)
        self.h_grows, self.h_gw = pin(self.umax, K), pin(self.umax)
        L = self.model.layout
        self.h_dense = {n: torch.zeros(L.segments[n].shape).pin_memory() for n in self.dense_names}
        self.h_gdense = {n: torch.zeros(L.segments[n].shape).pin_memory() for n in self.dense_names}
        self._graph = None

    def _fused_step(self, ids, vals, labels):
        m, st = self.model, self.store
        ids_d = ids.to(self.dev, non_blocking=True).long()
        u, inv = torch.unique(ids_d, return_inverse=True)
        U = int(u.numel())
        u_host = u.cpu()
        for n in self.dense_names:
            self.h_dense[n].copy_(st.view(n))
        self.side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(self.side):
            if st.dev_base:
                # ---- pull: the GPU gathers the touched rows straight out of the mapped host store (zero-copy)
                import ctypes
                from ..ops import lib as _nlib
                lib, sp = _nlib.load(), ctypes.c_void_p(self.side.cuda_stream)
                K = self.cfg.embedding_size
                _nlib.check(lib.dfm_ps_gather(ctypes.c_void_p(st.dev_ptr("fm_v")), ctypes.c_void_p(u.data_ptr()),
                                              ctypes.c_int(U), ctypes.c_int(K), ctypes.c_void_p(m._w("fm_v").data_ptr()), sp))
                _nlib.check(lib.dfm_ps_gather(ctypes.c_void_p(st.dev_ptr("fm_w")), ctypes.c_void_p(u.data_ptr()),
                                              ctypes.c_int(U), ctypes.c_int(1), ctypes.c_void_p(m._w("fm_w").data_ptr()), sp))
            else:
                # ---- pull: host gather into pinned staging, H2D on the side stream
                torch.index_select(st.view("fm_v"), 0, u_host, out=self.h_rows[:U])
                torch.index_select(st.view("fm_w"), 0, u_host, out=self.h_w[:U])
                m._w("fm_v")[:U].copy_(self.h_rows[:U], non_blocking=True)
                m._w("fm_w")[:U].copy_(self.h_w[:U], non_blocking=True)
            for n in self.dense_names:
                m._w(n).copy_(self.h_dense[n], non_blocking=True)
        torch.cuda.current_stream().wait_stream(self.side)
        m.refresh_shadow()
        # ---- compute: fused kernels on the compact table (ids remapped to 0..U-1), dropout keyed by the local step
        m.step_t.fill_(self._local_steps)
        m.load_inputs(inv.reshape(ids.shape).to(torch.int32), vals.to(self.dev, non_blocking=True),
                      labels.to(self.dev, non_blocking=True), 0)
        self.kernels_per_step = m._train_kernels(m.in_slots[0], with_optimizer=False)
        loss = m.loss_sum / self.B
        # ---- push: gradients D2H on the side stream, applied by the background thread
        self.side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(self.side):
            self.h_grows[:U].copy_(m._g("fm_v")[:U], non_blocking=True)
            self.h_gw[:U].copy_(m._g("fm_w")[:U], non_blocking=True)
            for n in self.dense_names:
                self.h_gdense[n].copy_(m._g(n), non_blocking=True)
            done = torch.cuda.Event()
            done.record(self.side)
        done.synchronize()
        m.G.zero_()
        push = {"rows": u_host, "fm_v": self.h_grows[:U].clone(), "fm_w": self.h_gw[:U].clone(),
                "dense": {n: self.h_gdense[n].clone() for n in self.dense_names}}
        return float(loss), push

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
            try:
                self._apply(item)
            finally:
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
        loss, push = self._fused_step(ids, vals, labels) if self.use_fused else self._torch_step(ids, vals, labels)
        with self._lock:
            self._inflight += 1
        self._q.put(push)
        self._losses.append(loss)
        self._local_steps += 1
        return None

    train_batch_device = train_batch

    def train_packed(self, packed):
        raise NotImplementedError("parameter-server workers take (ids, vals, labels) batches")

    def drain(self, keep: bool = False):
        with self._lock:
            while self._inflight > 0:
                self._lock.wait()
        if keep:
            return []
        out, self._losses = self._losses, []
        return out

    def last_loss(self):
        return torch.tensor(self._losses[-1] if self._losses else float("nan"))

    @property
    def global_step(self) -> int:
        return int(self.store.step_counter[0])

    def _params(self) -> Dict[str, torch.Tensor]:
        return OrderedDict((n, self.store.view(n)) for n in self.names)

    def parameter_names(self):
        return list(self.names)

    @torch.no_grad()
    def evaluate(self, batches: Iterable, distributed: bool = False):
        self.drain(keep=True)
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
                self.store.step_counter[0] = float(int(v.reshape(-1)[0]))
            elif k in self.store.offsets:
                self.store.view(k).copy_(v)
        self.opt.step = self.global_step

    def close(self):
        self.drain(keep=True)
        self._q.put(None)
        if self.world > 1 and dist.is_initialized():
            dist.barrier()
        self.store.close(unlink=self.rank == 0)
