"""Synchronous data parallelism over replicated tables that exchanges only what differs between the replicas
(reference: Horovod, HVD:149,262,372 — one process per GPU, full replica, dense NCCL all-reduce of every gradient).

Why not a dense all-reduce: the reference's embedding gradient is dense only because of the whole-table L2 term
(SURVEY 2.2), and that term is identical on every replica. What differs is the data gradient, which is non-zero only in
the table rows some rank looked up this step. Per step and per rank (csrc/sparse_allreduce.cu, csrc/split_opt.cu):

    side branch   touched_bitmap -> bitmap_exchange (N/8 bytes to every peer, P2P stores / multimem.st, flag barrier)
                  -> union_rowlist: OR of the W bitmaps + ascending list of the touched rows (one CTA, deterministic)
                  -> sweep_untouched: every row NO rank touched gets its L2-only update, locally and identically on every
                  replica, while forward/backward run (optimizer slots are replicated)
    main chain    forward, backward (as on one GPU), then barrier A (also deposits this rank's log-loss sum in every
                  peer) -> sparse_allreduce: in-place all-reduce of [fm_bias | MLP block | union rows], each item reduced
                  by exactly one rank (multimem.ld_reduce in the switch, or peer loads in rank order) and stored into
                  every replica's gradient buffer (multimem.st / peer stores) -> barrier B -> tail_opt: every rank applies
                  the same update to the same values and publishes (global log-loss sum, L2 term) into the host ring.

Replicas stay bit-identical because every rank consumes the same reduced values. NCCL is not on this path; the dense
ZeRO-1 kernel (parallel/dp.py, csrc/allreduce_opt.cu) remains as `DFM_DP_EXCHANGE=dense` and for the MLP of the row-sharded
model. DP-N on N shards of a batch equals DP-1 on the concatenated batch (global-mean loss, dropout keyed by global row).
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_int, c_longlong, c_void_p

import torch
import torch.distributed as dist

from ..models.fused import FusedDeepFM
from ..ops import kernels as K
from ..ops import lib as _nlib
from ..optim.tf1 import SLOT_INIT


class SparseDataParallel(FusedDeepFM):
    loss_is_global = True  # the published / last_loss() log-loss sum covers the GLOBAL batch
    slots_replicated = True  # every rank holds (and updates identically) the full optimizer slots

    def __init__(self, cfg, batch_size, optimizer, seed, device, world_size, rank, comm: str = "auto",
                 use_graph: bool = True, params=None, process_group=None):
        from .symm import SymmArena
        self.world, self.rank = world_size, rank
        self.group = process_group if process_group is not None else dist.group.WORLD
        lib = _nlib.require_gpu()
        L = self._make_layout(cfg)
        if not L.include_embeddings:
            raise ValueError("SparseDataParallel needs the tables in the flat buffer")
        self._words = K.bitmap_words(cfg.feature_size)
        nslots = len(SLOT_INIT[optimizer.name])
        regions = [("P", L.total_numel, torch.float32), ("G", L.total_numel, torch.float32)]
        for i in range(nslots):  # replicated slots; in the arena so start/resume broadcasts them with the same kernel
            regions.append(("S%d" % i, L.trainable_numel, torch.float32))
        regions += [("flags", lib.dfm_sparse_allreduce_flag_words(), torch.int32), ("loss_parts", 8, torch.float32),
                    ("bm_in", 8 * self._words, torch.int32), ("step_bcast", 4, torch.float32)]
        self.arena = SymmArena(regions, device, self.group)
        use_nvls = (comm == "nvls") or (comm == "auto" and self.arena.mc_ptr != 0 and world_size >= 4)
        if comm == "nvls" and not self.arena.mc_ptr:
            raise RuntimeError("comm='nvls' requested but no multicast mapping is available on this fabric")
        self.use_nvls = bool(use_nvls)
        self.comm_name = "sparse-exchange " + ("nvls-multimem" if self.use_nvls else "p2p")
        super().__init__(cfg, batch_size, optimizer, seed, device, params, global_batch=batch_size * world_size,
                         row_offset=rank * batch_size, use_graph=use_graph)
        i32 = dict(device=device, dtype=torch.int32)
        self.seq_ab = torch.zeros(1, **i32)
        self.seq_c = torch.zeros(1, **i32)
        self.done_ar = torch.zeros(1, **i32)
        self.done_bm = torch.zeros(1, **i32)
        self._tables = {k: self.arena.ptr_table(k) for k in ("G", "P", "flags", "loss_parts", "bm_in", "step_bcast")}
        for i in range(nslots):
            self._tables["S%d" % i] = self.arena.ptr_table("S%d" % i)
        # grid of the all-reduce: this rank's share is (dense block)/W float4 + (touched rows)/W x K/4 row chunks, two
        # items in flight per thread
        seg_w = L.segments["fm_w"]
        dense4 = (seg_w.offset + (L.trainable_numel - L.shadow_range[0])) // 4
        row_items = cfg.feature_size * (cfg.embedding_size // 4) // world_size  # upper bound: every row touched
        blocks = max((dense4 // world_size + 511) // 512, (row_items + 511) // 512, 8)
        self._ar_blocks = int(os.environ.get("DFM_AR_BLOCKS", min(148 * 2, blocks)))
        if os.environ.get("DFM_L2_PERSIST", "1") != "0":
            lib.dfm_set_l2_persist.restype = ctypes.c_longlong
            end = self.arena.offsets["flags"][0]  # P | G | slots are contiguous at the start of the arena
            self.l2_persist_bytes = int(lib.dfm_set_l2_persist(c_void_p(self.arena.buf.data_ptr()),
                                                               ctypes.c_ulonglong(end)))
        torch.cuda.synchronize(device)
        dist.barrier(group=self.group)

    # ---- allocation: parameters, gradients and slots live in symmetric memory
    def _alloc(self, name, numel, dtype):
        if name in ("P", "G", "S0", "S1"):
            return self.arena.view(name)
        return super()._alloc(name, numel, dtype)

    def _want_split(self) -> bool:
        return True

    def _setup_split_optimizer(self) -> None:
        dev, W, words = self.dev, self.world, self._words
        self.bm_own = torch.zeros(words, device=dev, dtype=torch.int32)
        self.bm_union = torch.zeros(words, device=dev, dtype=torch.int32)
        self.reg_acc = torch.zeros(1, device=dev, dtype=torch.float32)
        inbox = self.arena.view("bm_in")
        self._bm_slices = [inbox[r * words:(r + 1) * words] for r in range(W)]
        self.loss_parts = self.arena.view("loss_parts")
        lt = self.layout.trainable_numel
        self.split = K.SplitOptimizer(self.layout, self.cfg, self.opt, self.P[:lt], self.G[:lt], self.S0, self.S1,
                                      self.shadow, self.step_t, self.reg_acc, self.reg_loss, self.done_counter,
                                      self.bm_union, self._bm_slices, False, self.loss_sum, self.loss_parts, W,
                                      sweep_blocks=self._sweep_blocks())

    def _stream(self):
        return c_void_p(torch.cuda.current_stream().cuda_stream)

    def _split_side_kernels(self, slot) -> int:
        K.touched_bitmap(slot["ids"], self.cfg.feature_size, self.bm_own)
        t = self._tables
        _nlib.check(_nlib.load().dfm_bitmap_exchange(
            t["bm_in"], c_void_p(self.arena.mc("bm_in")), t["flags"], c_void_p(self.bm_own.data_ptr()),
            c_int(self.cfg.feature_size), c_int(self.rank), c_int(self.world), c_void_p(self.done_bm.data_ptr()),
            c_void_p(self.seq_c.data_ptr()), c_int(1 if self.use_nvls else 0), self._stream()))
        self.split.union_rowlist()
        self.split.sweep()
        return 4

    def _optimizer_kernels(self) -> int:
        s, t = self.split, self._tables
        _nlib.check(_nlib.load().dfm_sparse_allreduce(
            t["G"], c_void_p(self.arena.mc("G")), t["flags"], t["loss_parts"], c_int(self.rank), c_int(self.world),
            c_longlong(s.fmw_off), c_longlong(s.fmv_off), c_int(s.N), c_int(s.K), c_longlong(s.d0[0]),
            c_longlong(s.d0[1]), c_longlong(s.d1[0]), c_longlong(s.d1[1]), c_void_p(s.rows.data_ptr()),
            c_void_p(s.row_count.data_ptr()),
            c_void_p(self.loss_sum.data_ptr()), c_void_p(self.done_ar.data_ptr()), c_void_p(self.seq_ab.data_ptr()),
            c_int(1 if self.use_nvls else 0), c_int(self._ar_blocks), self._stream()))
        ring = getattr(self, "_result_ring", None)
        s.tail(publish=(self.scalars, ring) if ring is not None else None)
        return 3

    def set_result_ring(self, ring) -> bool:
        if self._graphs:
            return False
        self._result_ring = ring
        return ring is not None

    def _restore(self, snap):
        super()._restore(snap)
        self.done_ar.zero_()
        self.done_bm.zero_()
        torch.cuda.synchronize()
        dist.barrier(group=self.group)

    # ---- start / resume: rank `src`'s parameters, slots and step into every replica, no NCCL collective on the path
    def broadcast_state(self, src: int = 0):
        """Reference: BroadcastGlobalVariablesHook(0), HVD:372 (an NCCL broadcast of ~36 tensors per train() call). Here:
        three passes of multicast (NVLS) or per-peer P2P stores from inside bcast_kernel over the flat buffers."""
        torch.cuda.synchronize(self.dev)
        dist.barrier(group=self.group)
        lib = _nlib.load()
        sb = self.arena.view("step_bcast").view(torch.int64)
        if self.rank == src:
            sb[0] = self.step_t[0]
            st = self._stream()
            nv = c_int(1 if self.use_nvls else 0)

            def bcast(name, n):
                n4 = n // 4 * 4
                _nlib.check(lib.dfm_bcast_f32(self._tables[name], c_void_p(self.arena.mc(name)), c_longlong(n4),
                                              c_int(self.rank), c_int(self.world), nv, st))
                if n % 4:  # tail elements the float4 kernel does not cover
                    for r in range(self.world):
                        if r != src:
                            self.arena.peer_view(name, r)[n4:n].copy_(self.arena.view(name)[n4:n])
            bcast("P", self.P.numel())
            for i in range(len(SLOT_INIT[self.opt.name])):
                bcast("S%d" % i, self.layout.trainable_numel)
            bcast("step_bcast", 4)
        torch.cuda.synchronize(self.dev)
        dist.barrier(group=self.group)
        self.step_t.copy_(sb[0:1])
        self.opt.step = int(self.step_t.item())
        self.refresh_shadow()
        self.G.zero_()
        torch.cuda.synchronize(self.dev)
        dist.barrier(group=self.group)
