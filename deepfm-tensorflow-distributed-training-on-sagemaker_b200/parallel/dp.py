"""Synchronous data parallelism (reference: Horovod, HVD:149,262,372 — one process per GPU, full replica, gradients
averaged by `hvd.DistributedOptimizer`, LR x world size, state broadcast from rank 0, chief-only checkpoints).

Two implementations behind `Trainer`:

  * FusedDataParallel — the product. The flat fp32 gradient buffer lives in symmetric memory
    (`torch.distributed._symmetric_memory`: one VMM allocation per rank mapped into every peer, plus an NVLS
    multicast mapping when the fabric offers one). ONE kernel per step (csrc/allreduce_opt.cu) then does what
    Horovod + TF need an NCCL all-reduce, a cast/scale pass and 11 ApplyAdam launches for:
        rank r reduces ITS 1/W slice of the gradients across all peers (P2P loads over NVSwitch, or
        `multimem.ld_reduce` in-switch), applies L2 decay + the TF-1 optimizer update on that slice only
        (optimizer state is therefore sharded W ways — ZeRO-1 — cutting the HBM sweep W-fold), and stores the
        updated fp32 weights and bf16 GEMM shadows straight into every peer's replica (P2P stores / `multimem.st`).
    NCCL is not on this path; `comm='nccl'` selects the baseline (NCCL all-reduce + local flat optimizer).
  * install_torch_allreduce — plain `torch.distributed.all_reduce` hook for the torch backend (gloo on CPU).

Because the loss is the mean over the GLOBAL batch and dropout masks are keyed by global row, DP-N on N shards of
a batch equals DP-1 on the concatenated batch (tests/test_dp_cpu.py, tests/test_multi_gpu.py).
"""
from __future__ import annotations

from typing import Dict

import torch
import torch.distributed as dist


def install_torch_allreduce(model, group=None) -> None:
    """Sum the (already 1/global-batch-scaled) data gradients over ranks with one flat all-reduce."""

    def hook(grads: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        keys = sorted(grads)
        flat = torch.cat([grads[k].reshape(-1) for k in keys])
        dist.all_reduce(flat, group=group)
        out, o = {}, 0
        for k in keys:
            n = grads[k].numel()
            out[k] = flat[o:o + n].view_as(grads[k])
            o += n
        return out

    model.grad_hook = hook


def broadcast_model(model, src: int = 0, group=None) -> None:
    """`BroadcastGlobalVariablesHook(0)` (HVD:372): parameters, optimizer slots and the step from rank `src`."""
    sd = model.state_dict()
    for k in sorted(sd):
        t = sd[k].contiguous()
        dist.broadcast(t, src, group=group)
        sd[k] = t
    model.load_state_dict(sd, strict=False)


# ------------------------------------------------------------------------------------------------------------
from ctypes import c_float, c_int, c_longlong, c_void_p  # noqa: E402

from ..models.fused import FusedDeepFM  # noqa: E402
from ..models.layout import FlatLayout  # noqa: E402
from ..ops import lib as _nlib  # noqa: E402
from ..optim.tf1 import OPT_IDS, SLOT_INIT, SLOT_SUFFIXES  # noqa: E402


class FusedDataParallel(FusedDeepFM):
    """FusedDeepFM whose optimizer step is the fused all-reduce + sharded optimizer + all-gather kernel."""

    def __init__(self, cfg, batch_size, optimizer, seed, device, world_size, rank, comm: str = "auto",
                 use_graph: bool = True, params=None, process_group=None):
        from .symm import SymmArena
        self.world, self.rank = world_size, rank
        self.group = process_group if process_group is not None else dist.group.WORLD
        self.comm = comm
        L = self._make_layout(cfg)
        n = L.trainable_numel
        self._nccl = comm == "nccl"
        # Buckets (SURVEY K12 "launched per bucket so it overlaps backward"; Horovod reduces each gradient as backprop
        # produces it, HVD:262-263): the flat buffer is exchanged as [fm_bias | fm_w | fm_v] and [MLP block]. Each bucket is
        # cut into `world` equal slices, so every rank owns one slice of EACH (and the optimizer slots of both). The MLP
        # bucket is launched on the wgrad branch as soon as the weight gradients exist and runs under emb_bwd; only the
        # table bucket — which needs emb_bwd — stays on the critical path, and it is 20 % shorter than the whole buffer.
        import os as _osb
        split = L.shadow_range[0]
        # OFF by default (DFM_DP_BUCKETS=1 turns it on): measured at 2 GPUs, batch 1024, it is slower (138 vs 126 us/step). The
        # weight gradients are complete only ~9 us before emb_bwd ends (scripts/step_trace.py: wgrad_0 ends at 57 us, emb_bwd
        # at 66 us), the early bucket needs ~20 us (two flag barriers + its sweep), shares NVLink and L2 with emb_bwd while it
        # runs, and the table bucket cannot be shorter than the link time of its 15 MB — at this step size there is no
        # backward left to hide a bucket under. Kept for larger batches / deeper MLPs, where the window grows.
        bucketed = (not self._nccl) and L.include_embeddings and world_size > 1 and _osb.environ.get("DFM_DP_BUCKETS", "0") == "1"
        rnd = lambda x: ((x + world_size - 1) // world_size + 1023) // 1024 * 1024
        self._buckets = [(0, split, rnd(split)), (split, n, rnd(n - split))] if bucketed else [(0, n, rnd(n))]
        self._bucketed = bucketed
        if bucketed:  # the wgrads must finish early enough for the MLP bucket to run beside emb_bwd
            self._wgrad_fork_default = "before_dgrad0"
        chunk = self._buckets[0][2]
        self._chunk = chunk  # slice length of the FIRST bucket (the one that holds the tables: owner routing uses it)
        lib = _nlib.require_gpu()
        nslots = len(SLOT_INIT[optimizer.name])
        slot_n = n if self._nccl else sum(c for _, _, c in self._buckets)
        regions = [("P", L.total_numel, torch.float32), ("G", L.total_numel, torch.float32),
                   ("shadow", L.shadow_range[1] - L.shadow_range[0], torch.bfloat16),
                   ("flags", lib.dfm_allreduce_opt_flag_words(), torch.int32), ("reg_parts", 8, torch.float32),
                   ("loss_parts", 8, torch.float32), ("flags2", lib.dfm_allreduce_opt_flag_words(), torch.int32),
                   ("reg_parts2", 8, torch.float32)]
        for i in range(nslots):
            regions.append(("S%d" % i, slot_n, torch.float32))
        self.arena = SymmArena(regions, device, self.group)
        self._slot_n = slot_n
        super().__init__(cfg, batch_size, optimizer, seed, device, params, global_batch=batch_size * world_size,
                         row_offset=rank * batch_size, use_graph=use_graph)
        self.reg_local = torch.zeros(1, device=device, dtype=torch.float32)
        self.seq = torch.zeros(1, device=device, dtype=torch.int32)
        self._tables = {k: self.arena.ptr_table(k) for k in ("G", "P", "shadow", "flags", "reg_parts", "loss_parts", "flags2",
                                                             "reg_parts2")}
        # scratch of the early (MLP) bucket: it runs beside the main chain, so it shares nothing with the final launch
        self.reg_local2 = torch.zeros(1, device=device, dtype=torch.float32)
        self.seq2 = torch.zeros(1, device=device, dtype=torch.int32)
        self.done2 = torch.zeros(1, device=device, dtype=torch.int32)
        # owner routing (csrc/embed.cu EmbRoute): emb_bwd adds every row gradient straight into the buffer of the rank that
        # reduces it, so the embedding share of the reduce-scatter happens inside emb_bwd and the fused kernel reads it locally
        import os as _osr
        self._routed = False
        self._local_range = (0, 0)
        # measured (bench.py, batch 1024 per GPU): 2 GPUs 124.5 vs 131.2 us/step with routing, 8 GPUs 155.6 vs 133.3 — 16-byte
        # remote reductions to seven peers cost more link time than the dense multimem.ld_reduce they replace, so routing
        # is the default only at 2 ranks (DFM_DP_ROUTED=1 / 0 forces it on / off)
        want = _osr.environ.get("DFM_DP_ROUTED", "1" if world_size == 2 else "0")
        if not self._nccl and L.include_embeddings and want != "0":
            sw, sv = L.segments["fm_w"], L.segments["fm_v"]
            if sv.offset % cfg.embedding_size == 0 and chunk % cfg.embedding_size == 0:
                self._routed = True
                self._local_range = (sw.offset, L.shadow_range[0])  # [fm_w | fm_v] incl. their alignment padding
        # the fused kernel exchanges the per-rank log-loss sums and L2 partials: what it reports is the GLOBAL batch's loss
        self.loss_is_global = not self._nccl
        # measured on 8xB200: P2P wins at 2 ranks (55 vs 75 us per 18.8 MB step), NVLS multimem from 4 ranks up (71 vs 82 us
        # at 8): in-switch reduction pays once several peers contribute
        use_nvls = (comm == "nvls") or (comm == "auto" and self.arena.mc_ptr != 0 and world_size >= 4)
        if comm == "nvls" and not self.arena.mc_ptr:
            raise RuntimeError("comm='nvls' requested but no multicast mapping is available on this fabric")
        self.use_nvls = bool(use_nvls)
        self.comm_name = "nccl-allreduce+flat_opt (baseline)" if self._nccl else \
            ("fused-nvls-multimem" if self.use_nvls else "fused-p2p")
        import os as _os

        def grid_for(ck):
            blocks = (ck // 4 + 511) // 512  # two float4 per thread per trip
            if self.use_nvls:
                # NVLS: ~4 trips per thread instead of 1. With a single wave every CTA reduces, then every CTA broadcasts,
                # and the link runs one direction at a time; a few trips per thread let the multimem.st of trip k overlap the
                # multimem.ld_reduce of trip k+1 (measured at 8 ranks: 287 CTAs 69.7 us, 74 CTAs 58.3 us, 37 CTAs 67.0 us)
                blocks = max(32, (ck // 4 + 2047) // 2048)
            if _os.environ.get("DFM_AR_BLOCKS"):  # tuning knob: must be set identically on every rank
                blocks = int(_os.environ["DFM_AR_BLOCKS"])
            return max(1, min(lib.dfm_allreduce_opt_max_blocks(), blocks))
        self._grids = [grid_for(c) for _, _, c in self._buckets]
        self._grid_blocks = self._grids[0]
        import ctypes as _ct, os as _os2
        if _os2.environ.get("DFM_L2_PERSIST", "1") != "0":  # the symmetric arena (P | G | shadow | slot slices) is contiguous
            lib.dfm_set_l2_persist.restype = _ct.c_longlong
            self.l2_persist_bytes = int(lib.dfm_set_l2_persist(_ct.c_void_p(self.arena.buf.data_ptr()),
                                                               _ct.c_ulonglong(self.arena.nbytes)))
        torch.cuda.synchronize(device)
        dist.barrier(group=self.group)

    # ---- allocation in symmetric memory
    def _alloc(self, name, numel, dtype):
        if name in ("P", "G", "shadow"):
            return self.arena.view(name)
        if name in ("S0", "S1"):
            return self.arena.view(name)
        return super()._alloc(name, numel, dtype)

    def _slot_numel(self) -> int:
        return self._slot_n

    def _pre_step(self, side) -> int:
        # gradients are zeroed on a forked branch that overlaps the forward pass (peers finished reading them at
        # barrier B of the previous step)
        cur = torch.cuda.current_stream()
        ev = torch.cuda.Event()
        ev.record(cur)
        side.wait_event(ev)
        with torch.cuda.stream(side):
            if self._routed:
                # the table gradients are zeroed by their OWNER inside the fused kernel (a fast peer may already be adding
                # next step's rows to them now); only the ranges this rank accumulates locally are cleared here
                a, b = self._local_range
                self.G[:a].zero_()
                self.G[b:].zero_()
                return 2
            self.G.zero_()
        return 1

    def _optimizer_kernels(self) -> int:
        o, L = self.opt, self.layout
        if self._nccl:
            lt = L.trainable_numel
            dist.all_reduce(self.G[:lt], group=self.group)
            from ..ops import kernels as K
            decay = L.decay_range if self.cfg.l2_reg > 0 else (0, 0)
            K.flat_opt_step(self.P[:lt], self.G[:lt], self.S0, self.S1, o.name, o.lr, o.beta1, o.beta2, o.eps, o.momentum,
                            1.0, decay, self.cfg.l2_reg, L.shadow_range, self.shadow, self.step_t, self.reg_loss,
                            self.done_counter)
            return 1
        return self._launch_bucket(len(self._buckets) - 1 if not self._bucketed else 0, final=True)

    def _pieces(self, r: int):
        """[(global_begin, global_end, offset in rank r's slot storage)] — one slice of every bucket."""
        out, off = [], 0
        for (b, e, c) in self._buckets:
            out.append((min(e, b + r * c), min(e, b + (r + 1) * c), off))
            off += c
        return out

    @property
    def slice(self):  # the first bucket's slice of this rank (kept for callers that predate buckets)
        p = self._pieces(self.rank)[0]
        return (p[0], p[1])

    def _launch_bucket(self, bi: int, final: bool) -> int:
        """The fused reduce-scatter + optimizer + all-gather kernel over this rank's slice of bucket `bi`. final=True: the
        launch that ends the step (advances the step counter, exchanges the loss / L2 sums, publishes the result)."""
        o, L, lib = self.opt, self.layout, _nlib.load()
        decay = L.decay_range if self.cfg.l2_reg > 0 else (0, 0)
        t = self._tables
        st = c_void_p(torch.cuda.current_stream().cuda_stream)
        ring = getattr(self, "_result_ring", None) if final else None
        gb, ge, off = self._pieces(self.rank)[bi]
        sp = lambda S: c_void_p(S.data_ptr() + 4 * off if S is not None else 0)
        fl, rp = (t["flags"], t["reg_parts"]) if final else (t["flags2"], t["reg_parts2"])
        regl, done, seq = (self.reg_local, self.done_counter, self.seq) if final else (self.reg_local2, self.done2, self.seq2)
        _nlib.check(lib.dfm_allreduce_opt_step_pub(
            t["G"], t["P"], t["shadow"], fl, rp, c_void_p(self.arena.mc("G")),
            c_void_p(self.arena.mc("P")), c_void_p(self.arena.mc("shadow")), sp(self.S0), sp(self.S1), c_int(self.rank),
            c_int(self.world), c_longlong(gb), c_longlong(ge), c_int(OPT_IDS[o.name]), c_float(o.lr), c_float(o.beta1),
            c_float(o.beta2), c_float(o.eps), c_float(o.momentum), c_float(1.0), c_longlong(decay[0]),
            c_longlong(decay[1]), c_float(self.cfg.l2_reg), c_longlong(L.shadow_range[0]), c_longlong(L.shadow_range[1]),
            c_void_p(self.step_t.data_ptr()), c_void_p(regl.data_ptr()), c_void_p(done.data_ptr()),
            c_void_p(seq.data_ptr()), c_int(1 if self.use_nvls else 0), c_int(self._grids[bi]),
            t["loss_parts"] if final else None, c_void_p(self.loss_sum.data_ptr() if final else 0),
            c_void_p(self._reg_extra_ptr() if final else 0), c_void_p(self.scalars.data_ptr() if final else 0),
            c_void_p(ring.data_ptr() if ring is not None else 0), c_int(ring.shape[0] if ring is not None else 0),
            c_longlong(self._local_range[0]), c_longlong(self._local_range[1]), c_int(1 if final else 0), st))
        return 2 if self._grids[bi] > 148 else 1

    def _post_wgrad_kernels(self, sides) -> int:
        """Early bucket: once every wgrad branch is done, exchange + update the MLP block on the first of them, beside
        dgrad_0 / emb_bwd on the main chain (the final join of the step waits for it)."""
        if not self._bucketed:
            return 0
        s0 = sides[0]
        for sd in sides[1:]:
            ev = torch.cuda.Event()
            ev.record(sd)
            s0.wait_event(ev)
        with torch.cuda.stream(s0):
            return self._launch_bucket(1, final=False)

    def _exchange_all(self) -> int:
        """Every bucket on the current stream (self-checks; a training step launches them on their own branches)."""
        n = 0
        if self._bucketed:
            n += self._launch_bucket(1, final=False)
        return n + self._optimizer_kernels()

    def _restore(self, snap):
        super()._restore(snap)
        self.reg_local.zero_()
        self.reg_local2.zero_()
        self.done2.zero_()
        torch.cuda.synchronize()
        dist.barrier(group=self.group)

    def _emb_backward(self, slot) -> int:
        if not self._routed:
            return super()._emb_backward(slot)
        L, cfg = self.layout, self.cfg
        _nlib.check(_nlib.load().dfm_emb_bwd_routed(
            c_void_p(slot["ids"].data_ptr()), c_void_p(slot["vals"].data_ptr()), c_void_p(self._w("fm_v").data_ptr()),
            c_void_p(self.Ssum.data_ptr()), c_void_p(self.dy.data_ptr()), c_void_p(self.dx0.data_ptr()),
            c_int(self.dx0.stride(0)), self._tables["G"], c_longlong(L.segments["fm_w"].offset),
            c_longlong(L.segments["fm_v"].offset), c_longlong(self._chunk), c_int(self.world), c_int(self.B),
            c_int(cfg.field_size), c_int(cfg.embedding_size), c_int(cfg.feature_size),
            c_void_p(torch.cuda.current_stream().cuda_stream)))
        return 1

    def compute_grads(self, ids, vals, labels):
        """Debug/test path. With owner routing the table gradients of a row live on the rank that owns it (already summed
        over ranks, nothing elsewhere), so the SUM over ranks of what this returns is still the global gradient; the peers'
        adds must have landed before the buffer is read."""
        if not self._routed:
            return super().compute_grads(ids, vals, labels)
        from collections import OrderedDict
        torch.cuda.synchronize(self.dev)
        dist.barrier(group=self.group)
        self.G.zero_()
        torch.cuda.synchronize(self.dev)
        dist.barrier(group=self.group)
        self.load_inputs(ids, vals, labels, 0)
        self._train_kernels(self.in_slots[0], with_optimizer=False)
        torch.cuda.synchronize(self.dev)
        dist.barrier(group=self.group)
        out = OrderedDict((n, self.layout.view(self.G, n).clone()) for n in self.layout.names(trainable_only=True))
        dist.barrier(group=self.group)
        self.G.zero_()
        torch.cuda.synchronize(self.dev)
        dist.barrier(group=self.group)
        return out

    def _reg_extra_ptr(self) -> int:
        """Device address of an L2 term this rank computed outside the flat buffer (row-sharded tables), or 0."""
        return 0

    def set_result_ring(self, ring) -> bool:
        """The fused kernel's last CTA stores (global log-loss sum, L2 term) of step t into ring[(t - 1) % len(ring)]."""
        if self._nccl or self._graphs:
            return False
        self._result_ring = ring
        return ring is not None

    def last_loss(self) -> torch.Tensor:
        if self._nccl:  # baseline path: a per-rank estimate (this rank's batch scaled to the global one)
            torch.add(self.loss_sum * (self.world / self.global_batch), self.reg_loss, out=self.loss_out)
            return self.loss_out
        # the kernel left (global log-loss sum, complete L2 term) in scalars: the reference's loss, mean over the GLOBAL batch
        torch.add(self.loss_sum / self.global_batch, self.reg_loss, out=self.loss_out)
        return self.loss_out

    # ---- state: slots are sharded; the chief assembles them through peer-mapped views (no collective needed)
    def named_slots(self):
        from collections import OrderedDict
        if self._nccl:
            return super().named_slots()
        L = self.layout
        n = L.trainable_numel
        full = []
        for i in range(len(SLOT_INIT[self.opt.name])):
            buf = torch.empty(n, device=self.dev, dtype=torch.float32)
            for r in range(self.world):
                pv = self.arena.peer_view("S%d" % i, r)
                for b, e, off in self._pieces(r):
                    if e > b:
                        buf[b:e].copy_(pv[off:off + e - b])
            full.append(buf)
        out = OrderedDict()
        for name in L.names(trainable_only=True):
            out[name] = [L.view(f, name) for f in full]
        return out

    def load_state_dict(self, sd, strict: bool = True):
        L = self.layout
        for name, t in self.named_parameters().items():
            if name in sd:
                t.copy_(sd[name].to(self.dev))
            elif strict:
                raise KeyError("missing parameter %s" % name)
        n = L.trainable_numel
        self._loaded_full = []
        for i, suffix in enumerate(SLOT_SUFFIXES[self.opt.name]):
            full = torch.full((n,), SLOT_INIT[self.opt.name][i], device=self.dev, dtype=torch.float32)
            self._loaded_full.append(full)
            for name in L.names(trainable_only=True):
                key = "%s/%s" % (name, suffix)
                if key in sd:
                    L.view(full, name).copy_(sd[key].to(self.dev))
            mine = self.S0 if i == 0 else self.S1
            if self._nccl:
                mine.copy_(full)
            else:
                for b, e, off in self._pieces(self.rank):
                    mine[off:off + e - b].copy_(full[b:e])
        if "global_step" in sd:
            self.step_t.copy_(sd["global_step"].to(self.dev).reshape(1))
            self.opt.step = int(self.step_t.item())
        self.refresh_shadow()
        self.G.zero_()

    def broadcast_state(self, src: int = 0):
        """Initial/restored state from rank `src` to every replica with P2P copies into the peers' mapped buffers
        (reference: BroadcastGlobalVariablesHook(0), HVD:372 — an NCCL broadcast of ~36 tensors per train() call)."""
        torch.cuda.synchronize(self.dev)
        dist.barrier(group=self.group)
        n = self.layout.trainable_numel
        if self.rank == src:
            # parameters: one pass of multicast (NVLS) or per-peer P2P stores from inside a kernel (bcast_state)
            lib = _nlib.load()
            n_p4 = self.P.numel() // 4 * 4  # bcast length: the whole padded buffer (incl. non-trainable BN statistics)
            _nlib.check(lib.dfm_bcast_f32(self._tables["P"], c_void_p(self.arena.mc("P")), c_longlong(n_p4),
                                          c_int(self.rank), c_int(self.world), c_int(1 if self.use_nvls else 0),
                                          c_void_p(torch.cuda.current_stream().cuda_stream)))
            for r in range(self.world):
                if r == src:
                    continue
                if self.P.numel() % 4:
                    self.arena.peer_view("P", r)[-(self.P.numel() % 4):].copy_(self.P[-(self.P.numel() % 4):])
                self.arena.peer_view("shadow", r).copy_(self.shadow)
        nslots = len(SLOT_INIT[self.opt.name])
        if self._nccl:  # the NCCL baseline arm keeps replicated slots and NCCL broadcasts
            step = self.step_t.clone()
            dist.broadcast(step, src, group=self.group)
            self.step_t.copy_(step)
            for i in range(nslots):
                dist.broadcast(self.S0 if i == 0 else self.S1, src, group=self.group)
        else:
            # slot slices and the step counter: src stores every rank's pieces straight into that rank's arena through the
            # peer mappings (only src loaded the checkpoint) - no NCCL on the start / resume path. The 8-byte step travels
            # in the first words of the "reg_parts2" scratch region (rewritten by the kernels before they read it).
            if self.rank == src:
                for i in range(nslots):
                    full = torch.empty(n, device=self.dev, dtype=torch.float32)
                    loaded = getattr(self, "_loaded_full", None)
                    full.copy_(loaded[i] if loaded else self._full_slot_for_broadcast(i))
                    for r in range(self.world):
                        pv = self.arena.peer_view("S%d" % i, r)
                        for b, e, off in self._pieces(r):
                            if e > b:
                                pv[off:off + e - b].copy_(full[b:e])
                for r in range(self.world):
                    self.arena.peer_view("reg_parts2", r).view(torch.int64)[0:1].copy_(self.step_t.view(torch.int64))
            torch.cuda.synchronize(self.dev)
            dist.barrier(group=self.group)
            box = self.arena.view("reg_parts2").view(torch.int64)
            self.step_t.copy_(box[0:1].view(self.step_t.dtype))
            box.zero_()
        self.opt.step = int(self.step_t.item())
        self._loaded_full = None
        torch.cuda.synchronize(self.dev)
        dist.barrier(group=self.group)

    def _full_slot_for_broadcast(self, i: int) -> torch.Tensor:
        n = self.layout.trainable_numel
        buf = torch.empty(n, device=self.dev, dtype=torch.float32)
        for r in range(self.world):
            pv = self.arena.peer_view("S%d" % i, r)
            for b, e, off in self._pieces(r):
                if e > b:
                    buf[b:e].copy_(pv[off:off + e - b])
        return buf
