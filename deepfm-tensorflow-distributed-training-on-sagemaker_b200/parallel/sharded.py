"""Row-sharded embedding tables with fused all-to-all kernels (BASELINE.json configs 3 and 5).

The reference keeps both FM tables whole on every GPU (HVD) or whole on one parameter server (PS) and only *advises*
partitioning (`fixed_size_partitioner`, PDF p32; "borrow another GPU's memory", PDF p27). Here fm_w / fm_v are split by
row across the W GPUs of the box (row id -> rank id % W, local row id // W) while the MLP stays data parallel through
the fused all-reduce + optimizer kernel (parallel/dp.py). The three kernels in csrc/sharded.cu do the embedding
all-to-all *inside* the gather / gradient kernels over symmetric memory — see that file's header.

Optimizer semantics for the tables: touched rows only ("lazy"; `l2_mode='touched'`), which is TF's exact sparse rule for
Adagrad / GD / FTRL and the LazyAdam variant for Adam / Momentum (TF-1's own sparse Adam still sweeps the whole table
every step — unaffordable at 33 M x 64).
"""
from __future__ import annotations

from collections import OrderedDict
from ctypes import c_float, c_int, c_longlong, c_void_p
from typing import Dict, Optional

import math

import torch
import torch.distributed as dist

from ..models.layout import FlatLayout
from ..ops import lib as _nlib
from ..optim.tf1 import OPT_IDS, SLOT_INIT
from .dp import FusedDataParallel
from .symm import SymmArena


class ShardedDeepFM(FusedDataParallel):
    def __init__(self, cfg, batch_size, optimizer, seed, device, world_size, rank, comm: str = "auto",
                 use_graph: bool = True, params: Optional[Dict[str, torch.Tensor]] = None, process_group=None,
                 a2a_mode: str = "auto"):
        """a2a_mode: 'pull' (peer loads from the owner's shard; lowest latency, small batches), 'push' (requests out,
        rows pushed back with P2P stores, gradients pushed into owner inboxes; highest bandwidth, large batches),
        'auto' = push from 64 Ki lookups per rank per step (env DFM_A2A overrides)."""
        import os as _os
        a2a_mode = _os.environ.get("DFM_A2A", a2a_mode)
        lookups = batch_size * cfg.field_size
        self.push = a2a_mode == "push" or (a2a_mode == "auto" and lookups >= 65536 and world_size > 1)
        self._table_src = params
        N, K = cfg.feature_size, cfg.embedding_size
        self.rows_local = (N + world_size - 1) // world_size
        self._tables_ready = False
        super().__init__(cfg, batch_size, optimizer, seed, device, world_size, rank, comm, use_graph, params,
                         process_group)
        lib = _nlib.load()
        R = self.rows_local
        self.words = (R + 31) // 32
        regions = [("V", R * K, torch.float32), ("W", R, torch.float32), ("GV", R * K, torch.float32),
                   ("GW", R, torch.float32), ("bitmap", self.words, torch.int32),
                   ("flags", lib.dfm_sharded_flag_words(), torch.int32)]
        self.tarena = SymmArena(regions, device, self.group)
        self.V = self.tarena.view("V").view(R, K)
        self.Wt = self.tarena.view("W")
        self.GV = self.tarena.view("GV").view(R, K)
        self.GW = self.tarena.view("GW")
        self.bitmap = self.tarena.view("bitmap")
        inits = SLOT_INIT[self.opt.name]
        f32 = dict(device=device, dtype=torch.float32)
        self.VS = [torch.full((R, K), v, **f32) for v in inits]
        self.WS = [torch.full((R,), v, **f32) for v in inits]
        self._init_tables(seed)
        self._t = {k: self.tarena.ptr_table(k) for k in ("V", "W", "GV", "GW", "bitmap", "flags")}
        self.sh_seq = torch.zeros(1, device=device, dtype=torch.int32)
        self.sh_done = torch.zeros(1, device=device, dtype=torch.int32)
        self.e32 = torch.zeros(self.B, cfg.field_size * K, **f32)
        self._fwd_blocks = max(1, min(1184, (self.B + 7) // 8))
        # sparse_row_opt: a warp takes up to 32 bitmap words per trip and ~4 rows per inner iteration. Enough warps for
        # ~8 touched rows each (at most B*F*W / W lookups land on this shard), one wave of 4 CTAs/SM (64 registers).
        want_warps = max(self.words // 32, min(R, self.B * cfg.field_size) // 8, 1)
        self._opt_blocks = int(_os.environ.get("DFM_SPARSE_OPT_BLOCKS", max(1, min(592, (want_warps + 7) // 8))))
        if self.push:
            cap = self.B * cfg.field_size
            self.cap = cap
            Wn = world_size
            pregions = [("req", Wn * cap * 4, torch.int32), ("req_cnt", 8, torch.int32), ("rows_in", cap * K, torch.float32),
                        ("wv_in", cap, torch.float32), ("grad_in", Wn * cap * K, torch.float32),
                        ("gw_in", Wn * cap, torch.float32), ("pflags", lib.dfm_a2a_push_flag_words(), torch.int32)]
            self.parena = SymmArena(pregions, device, self.group)
            self._pt = {k: self.parena.ptr_table(k) for k in ("req", "req_cnt", "rows_in", "wv_in", "grad_in", "gw_in", "pflags")}
            self.cnt_local = torch.zeros(8, device=device, dtype=torch.int32)
            self.pos_of_pair = torch.zeros(cap, device=device, dtype=torch.int32)
            self.push_done = torch.zeros(1, device=device, dtype=torch.int32)
            self._push_blocks = max(1, min(1184, (cap + 63) // 64))
        self.comm_name = "sharded-embeddings fused-a2a (%s) + %s" % ("P2P push" if self.push else "P2P pull", self.comm_name)
        self._tables_ready = True
        torch.cuda.synchronize(device)
        dist.barrier(group=self.group)

    # ---- layout without the tables
    def _make_layout(self, cfg):
        return FlatLayout(cfg, include_embeddings=False)

    def _default_params(self, cfg, seed):
        from ..models.params import native_names
        from ..models.init import xavier_uniform_
        gen = torch.Generator().manual_seed(seed)
        p = OrderedDict()
        for name, shape, _ in native_names(cfg):
            if name in ("fm_w", "fm_v"):
                continue
            t = torch.zeros(shape)
            if name.endswith(".weight"):
                xavier_uniform_(t, shape[0] if name.startswith("deep_out") else shape[1],
                                1 if name.startswith("deep_out") else shape[0], gen)
            p[name] = t
        return p

    def _init_tables(self, seed):
        cfg, W, r = self.cfg, self.world, self.rank
        N, K = cfg.feature_size, cfg.embedding_size
        src = self._table_src
        n_mine = len(range(r, N, W))
        if src is not None and "fm_v" in src:
            self.V[:n_mine].copy_(src["fm_v"][r::W].to(self.dev))
            self.Wt[:n_mine].copy_(src["fm_w"][r::W].to(self.dev))
        else:  # shard-local glorot-normal (2-sigma clamp instead of resampling: tables this large are never compared)
            g = torch.Generator(device=self.dev).manual_seed(seed * 1000 + r)
            std_v = math.sqrt(1.0 / ((N + K) / 2.0)) / 0.87962566103423978
            std_w = math.sqrt(1.0 / N) / 0.87962566103423978
            self.V.normal_(0, std_v, generator=g).clamp_(-2 * std_v, 2 * std_v)
            self.Wt.normal_(0, std_w, generator=g).clamp_(-2 * std_w, 2 * std_w)
        self.GV.zero_(); self.GW.zero_(); self.bitmap.zero_()

    # ---- embedding kernels with the all-to-all fused in
    _train_mode = False
    _emb_zeroes_scalars = False

    def _train_kernels(self, slot, with_optimizer: bool = True) -> int:
        self._train_mode = True  # the forward all-to-all synchronises with the peers only inside a training step
        try:
            return super()._train_kernels(slot, with_optimizer)
        finally:
            self._train_mode = False

    def _emb_forward(self, slot, zero_scalars: bool = False) -> None:
        cfg, t = self.cfg, self._t
        if self.push and self._train_mode:
            pt = self._pt
            p = lambda x: c_void_p(x.data_ptr())
            self.cnt_local.zero_()
            _nlib.check(_nlib.load().dfm_a2a_push_forward(
                pt["req"], pt["req_cnt"], pt["rows_in"], pt["wv_in"], pt["grad_in"], pt["gw_in"], pt["pflags"],
                c_int(self.rank), c_int(self.world), c_longlong(self.cap), p(slot["ids"]), p(slot["vals"]), p(self.V),
                p(self.Wt), p(self.cnt_local), p(self.pos_of_pair), p(self.push_done), p(self.x0), c_void_p(0),
                c_int(self.x0.stride(0)), p(self.Ssum), p(self.yfm), c_int(self.B), c_int(cfg.field_size),
                c_int(cfg.embedding_size), c_longlong(cfg.feature_size), p(self.sh_seq), c_int(self._push_blocks),
                c_void_p(torch.cuda.current_stream().cuda_stream)))
            return
        _nlib.check(_nlib.load().dfm_emb_fm_fwd_a2a(
            t["V"], t["W"], t["GV"], t["GW"], t["bitmap"], t["flags"], c_int(self.rank), c_int(self.world),
            c_void_p(slot["ids"].data_ptr()), c_void_p(slot["vals"].data_ptr()), c_void_p(self.x0.data_ptr()),
            c_void_p(self.e32.data_ptr()), c_int(self.x0.stride(0)), c_void_p(self.Ssum.data_ptr()),
            c_void_p(self.yfm.data_ptr()), c_int(self.B), c_int(cfg.field_size), c_int(cfg.embedding_size),
            c_longlong(cfg.feature_size), c_void_p(self.sh_seq.data_ptr()), c_int(1 if self._train_mode else 0),
            c_int(self._fwd_blocks),
            c_void_p(torch.cuda.current_stream().cuda_stream)))

    def _emb_backward(self, slot) -> int:
        cfg, t, o = self.cfg, self._t, self.opt
        lib = _nlib.load()
        st = c_void_p(torch.cuda.current_stream().cuda_stream)
        if self.push:
            pt = self._pt
            q = lambda x: c_void_p(x.data_ptr())
            _nlib.check(lib.dfm_a2a_push_backward(
                pt["req"], pt["req_cnt"], pt["rows_in"], pt["wv_in"], pt["grad_in"], pt["gw_in"], pt["pflags"],
                c_int(self.rank), c_int(self.world), c_longlong(self.cap), q(slot["ids"]), q(slot["vals"]),
                q(self.pos_of_pair), q(self.e32), q(self.Ssum), q(self.dy), q(self.dx0), c_int(self.dx0.stride(0)),
                q(self.GV), q(self.GW), q(self.bitmap), c_int(self.B), c_int(cfg.field_size), c_int(cfg.embedding_size),
                q(self.sh_seq), c_int(self._push_blocks), st))
        else:
            self._bwd_pull(slot, st)
        return self._sparse_opt(st) + (1 if self.push else 0)

    def _bwd_pull(self, slot, st) -> None:
        cfg, t, lib = self.cfg, self._t, _nlib.load()
        _nlib.check(lib.dfm_emb_bwd_a2a(
            t["V"], t["W"], t["GV"], t["GW"], t["bitmap"], t["flags"], c_int(self.rank), c_int(self.world),
            c_void_p(slot["ids"].data_ptr()), c_void_p(slot["vals"].data_ptr()), c_void_p(self.e32.data_ptr()),
            c_void_p(self.Ssum.data_ptr()), c_void_p(self.dy.data_ptr()), c_void_p(self.dx0.data_ptr()),
            c_int(self.dx0.stride(0)), c_int(self.B), c_int(cfg.field_size), c_int(cfg.embedding_size),
            c_longlong(cfg.feature_size), st))

    def _sparse_opt(self, st) -> int:
        cfg, t, o, lib = self.cfg, self._t, self.opt, _nlib.load()
        p = lambda x: c_void_p(x.data_ptr())
        z = c_void_p(0)
        _nlib.check(lib.dfm_sparse_row_opt(
            t["V"], t["W"], t["GV"], t["GW"], t["bitmap"], t["flags"], c_int(self.rank), c_int(self.world),
            p(self.VS[0]) if len(self.VS) > 0 else z, p(self.VS[1]) if len(self.VS) > 1 else z,
            p(self.WS[0]) if len(self.WS) > 0 else z, p(self.WS[1]) if len(self.WS) > 1 else z,
            c_longlong(self.rows_local), c_int(cfg.embedding_size), c_int(OPT_IDS[o.name]), c_float(o.lr),
            c_float(o.beta1), c_float(o.beta2), c_float(o.eps), c_float(o.momentum), c_float(cfg.l2_reg),
            p(self.step_t), c_int(1), p(self.sh_seq), p(self.sh_done), p(self.reg_loss), c_int(self._opt_blocks), st))
        return 2

    def time_a2a(self, iters: int = 50, slot_idx: int = 0) -> Dict[str, float]:
        """Device time of the embedding all-to-all kernels alone (BASELINE.json metric: "embedding a2a time/step"): the
        fused dispatch+gather+combine forward, and the gradient all-to-all + segment-sum + sparse row update backward, each
        timed with CUDA events over `iters` back-to-back launches on the inputs / upstream gradients of the last step.
        Collective: every rank must call it. Returns microseconds per launch (this rank; take the max over ranks)."""
        slot = self.in_slots[slot_idx]
        out = {}
        self._train_mode = True
        try:
            for name, fn in (("fwd_us", lambda: self._emb_forward(slot)), ("bwd_opt_us", lambda: self._emb_backward(slot))):
                for _ in range(5):
                    fn()
                torch.cuda.synchronize(self.dev)
                dist.barrier(group=self.group)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(iters):
                    fn()
                e1.record()
                torch.cuda.synchronize(self.dev)
                out[name] = e0.elapsed_time(e1) * 1e3 / iters
        finally:
            self._train_mode = False
        out["a2a_us_per_step"] = out["fwd_us"] + out["bwd_opt_us"]
        lookups = self.B * self.cfg.field_size
        # bytes that must cross NVLink per rank and step: (W-1)/W of the looked-up rows in (forward) and of the gradient rows
        # out (backward), fp32 rows of K floats + the first-order scalar
        frac = (self.world - 1) / max(self.world, 1)
        out["nvlink_bytes_per_step"] = 2 * lookups * (self.cfg.embedding_size + 1) * 4 * frac
        out["mode"] = "push" if self.push else "pull"
        return out

    def _snapshot(self):
        snap = super()._snapshot()
        self._tsnap = [self.V.clone(), self.Wt.clone()] + [t.clone() for t in self.VS + self.WS]
        return snap

    def _restore(self, snap):
        torch.cuda.synchronize()
        dist.barrier(group=self.group)
        for dst, src in zip([self.V, self.Wt] + self.VS + self.WS, self._tsnap):
            dst.copy_(src)
        self.GV.zero_(); self.GW.zero_(); self.bitmap.zero_()
        self._tsnap = None
        super()._restore(snap)

    def _reg_extra_ptr(self) -> int:
        # touched-row L2 term of THIS rank's shard (sparse_row_opt adds it to reg_loss before the MLP exchange runs): the
        # fused kernel adds it to this rank's partial, so the published L2 term covers every shard
        return self.reg_loss.data_ptr()

    # ---- state: the chief assembles full tables through peer-mapped views
    def _gather_table(self, name: str, cols: int) -> torch.Tensor:
        N, W = self.cfg.feature_size, self.world
        full = torch.empty((N, cols) if cols > 1 else (N,), device=self.dev, dtype=torch.float32)
        for r in range(W):
            n_r = len(range(r, N, W))
            pv = self.tarena.peer_view(name, r)
            full[r::W] = pv[: n_r * cols].view(n_r, cols) if cols > 1 else pv[:n_r]
        return full

    def named_parameters(self):
        out = OrderedDict()
        flat = super().named_parameters()
        out["fm_bias"] = flat["fm_bias"]
        out["fm_w"] = self._gather_table("W", 1)
        out["fm_v"] = self._gather_table("V", self.cfg.embedding_size)
        for k, v in flat.items():
            if k != "fm_bias":
                out[k] = v
        return out

    def state_dict(self):
        sd = super().state_dict()  # flat (MLP) parameters + slots + step; tables via named_parameters above
        # table slots stay rank-local in this version (they are re-created at init_value on load)
        return sd

    def load_state_dict(self, sd, strict: bool = True):
        W, r, N = self.world, self.rank, self.cfg.feature_size
        n_mine = len(range(r, N, W))
        if "fm_v" in sd:
            self.V[:n_mine].copy_(sd["fm_v"][r::W].to(self.dev))
        if "fm_w" in sd:
            self.Wt[:n_mine].copy_(sd["fm_w"][r::W].to(self.dev))
        sub = {k: v for k, v in sd.items() if not k.startswith(("fm_v", "fm_w"))}
        super().load_state_dict(sub, strict=False)

    def broadcast_state(self, src: int = 0):
        super().broadcast_state(src)
        # tables: every rank needs ITS rows of src's full table; src pushes them through the peer mappings
        N, W, K = self.cfg.feature_size, self.world, self.cfg.embedding_size
        if getattr(self, "_loaded_tables", None) is not None and self.rank == src:
            fv, fw = self._loaded_tables
            for r in range(W):
                n_r = len(range(r, N, W))
                self.tarena.peer_view("V", r)[: n_r * K].copy_(fv[r::W].reshape(-1).to(self.dev))
                self.tarena.peer_view("W", r)[:n_r].copy_(fw[r::W].to(self.dev))
        torch.cuda.synchronize(self.dev)
        dist.barrier(group=self.group)
