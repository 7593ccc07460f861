"""FusedDeepFM — the B200 execution engine for the reference `model_fn` (PS:149-278 / HVD:141-273).

One training step is a fixed sequence of hand-written sm_100a kernels over persistent buffers, captured once in
a CUDA graph and replayed (no autograd, no allocator traffic, no host sync):

    emb_fm_fwd  ->  mlp_fwd x L (tcgen05, bias+ReLU+dropout epilogue)  ->  logit_loss (loss, dy, dZ_top, head grads)
    -> per layer, top down:  wgrad (tcgen05, MN-major operands, split-K atomics)  ||  dgrad (tcgen05, mask epilogue)
       || colsum (bias grads)            [wgrad/colsum run on a forked graph branch]
    -> emb_bwd (vector atomics into the dense grad buffer)
    -> flat optimizer (L2 decay + TF-1 update + grad zeroing + bf16 weight shadows + step counter)
       or, under data parallelism, the fused all-reduce + optimizer kernel (parallel/dp.py)

Layout: models/layout.py. Numerics oracle: models/oracle.py. Kernel ABI: include/deepfm_b200.h.
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Dict, Optional

import torch

from ..optim.tf1 import SLOT_INIT, SLOT_SUFFIXES, TF1Optimizer
from ..ops import kernels as K
from ..ops import lib as _native
from ..utils.auc import NUM_THRESHOLDS, thresholds
from .config import DeepFMConfig
from .layout import FlatLayout
from .params import init_params


def packed_slot_bytes(B: int, F: int) -> int:
    al = lambda x: (x + 255) & ~255
    return al(B * 4) + 2 * al(B * F * 4)


def make_packed_slot(B: int, F: int, device, pin: bool = False) -> Dict[str, torch.Tensor]:
    """One batch in the loader's slot layout; returns the raw buffer plus typed views."""
    al = lambda x: (x + 255) & ~255
    buf = torch.zeros(packed_slot_bytes(B, F), dtype=torch.uint8, device=device)
    if pin:
        buf = buf.pin_memory()
    o_ids = al(B * 4)
    o_vals = o_ids + al(B * F * 4)
    return dict(packed=buf, labels=buf[0:B * 4].view(torch.float32),
                ids=buf[o_ids:o_ids + B * F * 4].view(torch.int32).view(B, F),
                vals=buf[o_vals:o_vals + B * F * 4].view(torch.float32).view(B, F))


def pack_batch(ids: torch.Tensor, vals: torch.Tensor, labels: torch.Tensor, pin: bool = False, device="cpu"):
    B, F = ids.shape
    s = make_packed_slot(B, F, device, pin)
    s["ids"].copy_(ids.to(torch.int32))
    s["vals"].copy_(vals)
    s["labels"].copy_(labels.reshape(-1))
    return s["packed"]


class FusedDeepFM:
    def __init__(self, cfg: DeepFMConfig, batch_size: int, optimizer: Optional[TF1Optimizer] = None, seed: int = 0,
                 device: Optional[torch.device] = None, params: Optional[Dict[str, torch.Tensor]] = None,
                 global_batch: Optional[int] = None, row_offset: int = 0, use_graph: bool = True,
                 input_slots: int = 2):
        """`global_batch`/`row_offset`: under data parallelism the loss is the mean over the GLOBAL batch and the
        dropout mask of row r is that of global row row_offset+r, so DP-N is comparable to DP-1."""
        problems = cfg.fused_constraints()
        if problems:
            raise ValueError("this model cannot use the fused kernels (use --backend torch): " + "; ".join(problems))
        _native.require_gpu()
        self.cfg = cfg
        self.B = int(batch_size)
        self.global_batch = int(global_batch or batch_size)
        self.row_offset = int(row_offset)
        self.dev = device or torch.device("cuda", torch.cuda.current_device())
        self.opt = optimizer or TF1Optimizer("Adam", 5e-4)
        self.seed = int(seed)
        self.layout = self._make_layout(cfg)
        self.use_graph = use_graph
        L = self.layout
        F_, Kd = cfg.field_size, cfg.embedding_size
        dev = self.dev
        f32 = dict(device=dev, dtype=torch.float32)

        # ---- parameters / grads / slots (flat) ------------------------------------------------------------
        # P, G and the optimizer slots are carved from ONE allocation so a single persisting-L2 window can cover them
        nsl = len(SLOT_INIT[self.opt.name])
        self._state_arena = None
        self._state_off = 0
        if type(self)._alloc is FusedDeepFM._alloc:
            self._state_arena = torch.zeros(2 * L.total_numel + nsl * self._slot_numel() + 256, device=dev,
                                            dtype=torch.float32)
        self.P = self._alloc("P", L.total_numel, torch.float32)
        self.G = self._alloc("G", L.total_numel, torch.float32)
        inits = SLOT_INIT[self.opt.name]
        ns = self._slot_numel()
        self.S0 = self._alloc("S0", ns, torch.float32).fill_(inits[0]) if len(inits) > 0 else None
        self.S1 = self._alloc("S1", ns, torch.float32).fill_(inits[1]) if len(inits) > 1 else None
        self.shadow = self._alloc("shadow", L.shadow_range[1] - L.shadow_range[0], torch.bfloat16)
        src = params if params is not None else self._default_params(cfg, seed)
        for name in L.names():
            L.view(self.P, name).copy_(src[name].to(dev))
        self._init_params_src = src
        self.refresh_shadow()

        # ---- scalars ---------------------------------------------------------------------------------------
        self.step_t = torch.zeros(1, device=dev, dtype=torch.int64)
        self.scalars = torch.zeros(2, **f32)
        self.loss_sum = self.scalars[0:1]
        self.reg_loss = self.scalars[1:2]
        self.loss_out = torch.zeros(1, **f32)
        self.done_counter = torch.zeros(1, device=dev, dtype=torch.int32)

        # ---- activations -----------------------------------------------------------------------------------
        B = self.B
        bf = dict(device=dev, dtype=torch.bfloat16)
        # input slots use the native loader's packed slot layout (labels | ids | vals, 256-byte aligned parts) so a
        # whole batch moves host->device with ONE copy
        self.in_slots = [make_packed_slot(B, F_, dev) for _ in range(max(1, input_slots))]
        # Layer inputs carry a few extra columns (row stride rounded up to 128 bytes), the first of them all ones: the wgrad GEMM dW_i = dZ_i^T [X_i | 1] then yields
        # the bias gradient (column sums of dZ_i) as one more output column on the tensor core — no colsum kernel, no
        # atomics, nothing on the critical chain. Kernels see [B, width] views of the padded buffers.
        def act_buf(width):
            ld = (width + 1 + 63) // 64 * 64  # rows start on 128-byte lines: every TMA box row is one line, not two halves
            buf = torch.zeros(B, ld, **bf)
            buf[:, width] = 1.0
            return buf, buf[:, :width]
        self._x0_buf, self.x0 = act_buf(F_ * Kd)
        self.dx0 = torch.zeros(B, F_ * Kd, **bf)
        self.Ssum = torch.zeros(B, Kd, **f32)
        self.yfm = torch.zeros(B, **f32)
        self._act_bufs, self.acts = zip(*[act_buf(h) for h in cfg.deep_layers])
        self.acts = list(self.acts)
        self.dz = [torch.zeros(B, h, **bf) for h in cfg.deep_layers]
        self.bn = bool(cfg.batch_norm)
        if self.bn:  # ReLU outputs (pre-BN) and the saved batch statistics
            self.zbuf = [torch.zeros(B, h, **f32) for h in cfg.deep_layers]  # fp32: BN amplifies rounding of z
            self.bn_mean = [torch.zeros(h, **f32) for h in cfg.deep_layers]
            self.bn_rstd = [torch.zeros(h, **f32) for h in cfg.deep_layers]
        self.dy = torch.zeros(B, **f32)
        self.prob = torch.zeros(B, **f32)
        self.logit = torch.zeros(B, **f32)
        self.auc_thr = torch.from_numpy(thresholds()).to(dev)
        self.auc_hist = torch.zeros(2, NUM_THRESHOLDS + 1, **f32)

        # DFM_DETERMINISTIC=1: embedding gradients by sorted segment-sum (csrc/embed_det.cu) instead of fp32 atomics; with
        # the store-based wgrads this makes the big gradients bit-reproducible run to run (the bias / head column sums and
        # the loss scalar still accumulate a few hundred atomics)
        import os as _os
        self.det_emb = None
        if _os.environ.get("DFM_DETERMINISTIC", "0") == "1" and self.layout.include_embeddings \
                and type(self)._emb_backward is FusedDeepFM._emb_backward and Kd <= 64:
            self.det_emb = K.DetEmbBwd(B, F_, Kd, cfg.feature_size, dev)
        self._side = torch.cuda.Stream(device=dev)
        self._side2 = torch.cuda.Stream(device=dev)
        self._setup_split_optimizer()
        self._build_plans()
        self._graphs: Dict[int, torch.cuda.CUDAGraph] = {}
        self.kernels_per_step = 0
        self.host_step = 0
        self.l2_persist_bytes = self.enable_l2_persistence()

    # ------------------------------------------------------------------------------------------------ allocation
    def _make_layout(self, cfg) -> FlatLayout:
        return FlatLayout(cfg)

    def _default_params(self, cfg, seed):
        return init_params(cfg, seed)

    def _alloc(self, name: str, numel: int, dtype) -> torch.Tensor:
        """Persistent buffer allocation; parallel/dp.py overrides this to place P/G/shadow/slots in symmetric memory."""
        if name in ("P", "G", "S0", "S1") and getattr(self, "_state_arena", None) is not None:
            o = self._state_off
            self._state_off = (o + numel + 63) // 64 * 64
            return self._state_arena[o:o + numel]
        return torch.zeros(numel, device=self.dev, dtype=dtype)

    def enable_l2_persistence(self) -> int:
        """Ask the device to keep the training state (parameters, gradients, optimizer slots — re-read every step) in
        the persisting part of the 126 MB L2. Returns the bytes set aside (0 = unsupported or switched off)."""
        import ctypes, os
        if os.environ.get("DFM_L2_PERSIST", "1") == "0" or self._state_arena is None:
            return 0
        lib = _native.load()
        lib.dfm_set_l2_persist.restype = ctypes.c_longlong
        return int(lib.dfm_set_l2_persist(ctypes.c_void_p(self._state_arena.data_ptr()),
                                          ctypes.c_ulonglong(self._state_off * 4)))

    def _slot_numel(self) -> int:
        return self.layout.trainable_numel

    _wgrad_fork_default = "after_dgrad0"  # where the wgrad branches fork (data-parallel buckets fork one kernel earlier)

    def _post_wgrad_kernels(self, sides) -> int:
        """Hook: work that only needs the weight gradients (parallel/dp.py launches the MLP bucket of the gradient exchange
        here, on the wgrad branches). Returns the number of kernels it launched."""
        return 0

    def _wgrad_streams(self, n: int):
        st = getattr(self, "_wg_streams", None)
        if st is None or len(st) < n:
            st = self._wg_streams = [self._side] + [torch.cuda.Stream(device=self.dev) for _ in range(n - 1)]
        return st[:n]

    def _want_split(self) -> bool:
        """Split dense optimizer (csrc/split_opt.cu; needs the tables in the flat buffer). Off by default on one GPU
        (DFM_SPLIT_OPT=1 turns it on): measured on a B200 at batch 1024 the tail (MLP block + 20 K touched rows, random
        128-byte accesses at ~2 TB/s) is 16-17 us against 20 us for the single streaming sweep, and the concurrent
        untouched-row sweep slows the latency-bound forward/backward kernels by more than those 3-4 us (94.8 vs 92.0 us per
        step, profiles/step_timeline.md). It pays where the sweep is the smaller part: parallel/dp_sparse.py."""
        import os
        return self.layout.include_embeddings and os.environ.get("DFM_SPLIT_OPT", "0") == "1" \
            and type(self)._optimizer_kernels is FusedDeepFM._optimizer_kernels

    def _setup_split_optimizer(self) -> None:
        self.split = None
        if not self._want_split():
            return
        words = K.bitmap_words(self.cfg.feature_size)
        dev = self.dev
        self.bm_own = torch.zeros(words, device=dev, dtype=torch.int32)
        self.bm_union = torch.zeros(words, device=dev, dtype=torch.int32)
        self.reg_acc = torch.zeros(1, device=dev, dtype=torch.float32)
        lt = self.layout.trainable_numel
        self.split = K.SplitOptimizer(self.layout, self.cfg, self.opt, self.P[:lt], self.G[:lt], self.S0, self.S1,
                                      self.shadow, self.step_t, self.reg_acc, self.reg_loss, self.done_counter,
                                      self.bm_union, [self.bm_own], True, self.loss_sum,
                                      sweep_blocks=self._sweep_blocks())

    @staticmethod
    def _sweep_blocks() -> int:
        """Grid of the untouched-row sweep: one 8-warp CTA per SM. It runs beside the forward/backward kernels for most of
        the step and must leave each SM's registers, warps and bandwidth mostly to them (measured: with ~3 CTAs per SM the
        GEMM CTAs, one per SM by shared memory, queued behind it and the step got SLOWER than the single sweep)."""
        import os
        return int(os.environ.get("DFM_SWEEP_BLOCKS", "148"))

    def _split_side_kernels(self, slot) -> int:
        """Side branch of a split-optimizer step (forked at step start, joined before the tail): mark the rows this batch
        looks up, then update every other table row while forward/backward run."""
        K.touched_bitmap(slot["ids"], self.cfg.feature_size, self.bm_own)
        self.split.union_rowlist()
        self.split.sweep()
        return 3

    def _pre_step(self, side) -> int:
        """Hook at the start of a step (runs inside graph capture); returns #kernels it launched."""
        return 0

    # ------------------------------------------------------------------------------------------------ plans
    def _w(self, name):  # fp32 view
        return self.layout.view(self.P, name)

    def _g(self, name):
        return self.layout.view(self.G, name)

    def _wb(self, name):  # bf16 shadow view
        return self.layout.shadow_view(self.shadow, name)

    def _build_plans(self):
        cfg, B = self.cfg, self.B
        nl = len(cfg.deep_layers)
        self.fwd_train, self.fwd_eval, self.wgrad, self.dgrad = [], [], [], []
        # bias gradients ride the wgrad GEMMs (ones column), the head-weight gradient X_top^T dy is reduced in logit_loss
        self._fold_cols = (not self.bn) and cfg.deep_layers[-1] <= 1024
        x_in = self.x0
        for i, h in enumerate(cfg.deep_layers):
            d_in = x_in.shape[1]
            Wb = self._wb("mlp%d.weight" % i)
            bias = self._w("mlp%d.bias" % i)
            if self.bn:  # GEMM epilogue stops at ReLU; BN + dropout run in bn_fwd
                plan = K.GemmPlan("fwd", x_in, Wb, self.zbuf[i], B, h, d_in, bias=bias, relu=True).set_out_f32()
                self.fwd_train.append(plan)
                self.fwd_eval.append(plan)
            else:
                self.fwd_train.append(K.GemmPlan("fwd", x_in, Wb, self.acts[i], B, h, d_in, bias=bias, relu=True,
                                                 keep_prob=cfg.dropout[i], seed=self.seed, layer=i,
                                                 row_offset=self.row_offset, step_ptr=self.step_t))
                self.fwd_eval.append(K.GemmPlan("fwd", x_in, Wb, self.acts[i], B, h, d_in, bias=bias, relu=True))
            import os as _os
            gb = self._g("mlp%d.bias" % i) if self._fold_cols else None  # bias gradient = the ones column of [X | 1]
            if _os.environ.get("DFM_WGRAD_SPLITK", "0") == "1":  # split-K with red.add (fills the SMs; atomics)
                self.wgrad.append(K.GemmPlan("wgrad", self.dz[i], x_in, self._g("mlp%d.weight" % i), h, d_in, B, bias=gb))
            else:  # one K split per tile, plain stores: the weight gradient is written, not accumulated
                self.wgrad.append(K.GemmPlan("wgrad", self.dz[i], x_in, self._g("mlp%d.weight" % i), h, d_in, B, bias=gb,
                                             split_k=1).set_overwrite(True))
            if i == 0:
                self.dgrad.append(K.GemmPlan("dgrad", self.dz[0], Wb, self.dx0, B, d_in, h))
                import os as _os2
                if _os2.environ.get("DFM_WGRAD_FORK", self._wgrad_fork_default) != "before_dgrad0":
                    self.dgrad[0].set_exclusive(True)  # nothing runs beside it: whole-SM pipeline
            else:
                self.dgrad.append(K.GemmPlan("dgrad", self.dz[i], Wb, self.dz[i - 1], B, d_in, h,
                                             mask=self.acts[i - 1], mask_scale=1.0 / cfg.dropout[i - 1]))
                self.dgrad[i].set_exclusive(True)  # the wgrads are forked after dgrad_1: dgrad_i (i >= 1) runs alone
            x_in = self.acts[i]
            # timeline ids (csrc/common.cuh TR_*): forward 1+i, dgrad 5+i, wgrad 8+i
            if i < 3:
                self.fwd_train[i].set_trace_id(1 + i)
                self.dgrad[i].set_trace_id(5 + i)
                self.wgrad[i].set_trace_id(8 + i)
        self.n_layers = nl

    def refresh_shadow(self):
        a, b = self.layout.shadow_range
        K.cast_f32_bf16(self.P[a:b], self.shadow)

    # ------------------------------------------------------------------------------------------------ kernels
    _emb_zeroes_scalars = True  # the replicated emb_fm_fwd clears the step's loss accumulators itself

    def _emb_forward(self, slot, zero_scalars: bool = False) -> None:
        K.emb_fm_fwd(slot["ids"], slot["vals"], self._w("fm_w"), self._w("fm_v"), self.x0, self.Ssum, self.yfm,
                     self.scalars if zero_scalars else None)

    def _emb_backward(self, slot) -> int:
        if self.det_emb is not None:  # the lookups were sorted on the side branch at step start (_train_kernels)
            self.det_emb.run(slot["vals"], self._w("fm_v"), self.Ssum, self.dy, self.dx0, self._g("fm_w"), self._g("fm_v"))
            return 2
        K.emb_bwd(slot["ids"], slot["vals"], self._w("fm_v"), self.Ssum, self.dy, self.dx0, self._g("fm_w"),
                  self._g("fm_v"))
        return 1

    def _forward(self, slot, train: bool, rows: Optional[int] = None, zero_scalars: bool = False):
        cfg = self.cfg
        if zero_scalars and self._emb_zeroes_scalars:
            self._emb_forward(slot, True)
        else:
            if zero_scalars:
                self.scalars.zero_()
            self._emb_forward(slot)
        for i, p in enumerate(self.fwd_train if train else self.fwd_eval):
            p.run()
            if self.bn:
                K.bn_fwd(self.zbuf[i], self.acts[i], self._w("bn_%d.gamma" % i), self._w("bn_%d.beta" % i),
                         self._w("bn_%d.moving_mean" % i), self._w("bn_%d.moving_variance" % i), self.bn_mean[i],
                         self.bn_rstd[i], cfg.batch_norm_decay, train, cfg.dropout[i], self.seed, i, self.row_offset,
                         self.step_t)
        return 1 + self.n_layers * (2 if self.bn else 1)

    def _train_kernels(self, slot, with_optimizer: bool = True) -> int:
        """Enqueue one full training step on the current stream (+ a forked side branch). Returns #kernels."""
        cfg, nl = self.cfg, self.n_layers
        cur = torch.cuda.current_stream()
        n = 0
        pre = self._pre_step(self._side)
        split = self.split is not None and with_optimizer
        det_ev = None
        if self.det_emb is not None:  # fork: sort the lookups by row id (needs only the inputs) while the forward runs
            ev0 = torch.cuda.Event()
            ev0.record(cur)
            self._side2.wait_event(ev0)
            with torch.cuda.stream(self._side2):
                self.det_emb.sort(slot["ids"], slot["vals"])
                n += 4
                det_ev = torch.cuda.Event()
                det_ev.record(self._side2)
        if split:  # fork: touched-row bitmap + sweep of the untouched table rows overlap the whole forward/backward
            ev0 = torch.cuda.Event()
            ev0.record(cur)
            self._side2.wait_event(ev0)
            with torch.cuda.stream(self._side2):
                n += self._split_side_kernels(slot)
        n += self._forward(slot, True, zero_scalars=True)  # also clears loss_sum / reg_loss
        if pre:  # join the pre-step branch (e.g. gradient zeroing under data parallelism) before the first grad write
            evp = torch.cuda.Event()
            evp.record(self._side)
            cur.wait_event(evp)
        top = nl - 1
        K.logit_loss(self.acts[top], self._w("deep_out.weight"), self._w("deep_out.bias"), self._w("fm_bias"), self.yfm,
                     slot["labels"], 1.0 / self.global_batch, 1.0 / cfg.dropout[top], True, self.prob, self.logit,
                     self.dy, self.dz[top], self.loss_sum, self._g("deep_out.bias"), self._g("fm_bias"),
                     None, None if not self._fold_cols else self._g("deep_out.weight"))
        n += 1
        # Weight gradients: every wgrad_i only needs dz[i] (ready before dgrad_i starts) and nothing needs it before the
        # optimizer. They are forked late and side by side: after the whole dgrad chain, all L wgrads run at once on their
        # own branches (28 + 28 + 80 CTAs at 400-400-400: they fit the machine together) beside emb_bwd, which is bound by
        # its atomics and leaves the tensor cores idle. Forked as early as possible instead, each shared the SMs, the TMA
        # path and L2 with the dgrad chain and doubled its kernels' time (scripts/step_trace.py: dgrad 2.9 -> 5.9 us each;
        # DFM_WGRAD_FORK=before_dgrad0 starts them one kernel earlier: dgrad_0 3.7 -> 7.1 us).
        sides = self._wgrad_streams(nl)
        import os as _os
        fork_before_dgrad0 = _os.environ.get("DFM_WGRAD_FORK", self._wgrad_fork_default) == "before_dgrad0"

        def fork_wgrads():
            k = 0
            ev = torch.cuda.Event()
            ev.record(cur)
            for j in range(top, -1, -1):
                sides[j].wait_event(ev)
                with torch.cuda.stream(sides[j]):
                    self.wgrad[j].run()
                    k += 1
                    if not self._fold_cols:  # batch-norm path / very wide top layer: column sums in their own pass
                        K.colsum_bf16(self.dz[j], self._g("mlp%d.bias" % j))
                        k += 1
                        if j == top:  # g_w_out = X_top^T dy
                            K.colsum_bf16(self.acts[top], self._g("deep_out.weight"), self.dy)
                            k += 1
            return k

        for i in range(top, -1, -1):
            if self.bn:  # dz[i] holds d(BN output); turn it into d(pre-activation) in place (BN + ReLU backward)
                K.bn_bwd(self.dz[i], self.zbuf[i], self._w("bn_%d.gamma" % i), self.bn_mean[i], self.bn_rstd[i],
                         self._g("bn_%d.gamma" % i), self._g("bn_%d.beta" % i))
                n += 1
            if i == 0 and fork_before_dgrad0:  # dz[0] (and every dz above it) is final here
                n += fork_wgrads()
                if with_optimizer:
                    n += self._post_wgrad_kernels(sides)
            self.dgrad[i].run()
            n += 1
        if not fork_before_dgrad0:
            n += fork_wgrads()
            if with_optimizer:
                n += self._post_wgrad_kernels(sides)
        if det_ev is not None:
            cur.wait_event(det_ev)
        n += self._emb_backward(slot)
        for sd in sides:  # join
            ev = torch.cuda.Event()
            ev.record(sd)
            cur.wait_event(ev)
        if split:
            ev2 = torch.cuda.Event()
            ev2.record(self._side2)
            cur.wait_event(ev2)
        if with_optimizer:
            n += self._optimizer_kernels()
        return n

    def compute_grads(self, ids, vals, labels) -> "OrderedDict[str, torch.Tensor]":
        """Debug/test path: forward+backward WITHOUT the optimizer; returns copies of the data gradients
        (the whole-table L2 term is applied inside the optimizer sweep and is not part of these)."""
        self.G.zero_()
        self.load_inputs(ids, vals, labels, 0)
        self._train_kernels(self.in_slots[0], with_optimizer=False)
        torch.cuda.synchronize()
        out = OrderedDict((n, self.layout.view(self.G, n).clone()) for n in self.layout.names(trainable_only=True))
        self.G.zero_()
        return out

    def _optimizer_kernels(self) -> int:
        o, L = self.opt, self.layout
        lt = L.trainable_numel
        decay = L.decay_range if self.cfg.l2_reg > 0 else (0, 0)
        ring = getattr(self, "_result_ring", None)
        if self.split is not None:
            self.split.tail(publish=(self.scalars, ring) if ring is not None else None)
            return 1
        K.flat_opt_step(self.P[:lt], self.G[:lt], self.S0, self.S1, o.name, o.lr, o.beta1, o.beta2, o.eps, o.momentum,
                        1.0, decay, self.cfg.l2_reg, L.shadow_range, self.shadow, self.step_t, self.reg_loss,
                        self.done_counter, publish=(self.scalars, ring) if ring is not None else None)
        return 1

    def set_result_ring(self, ring: Optional[torch.Tensor]) -> bool:
        """Have the optimizer kernel store each step's (loss sum, L2 term) into `ring[(step - 1) % len(ring)]`, a pinned
        HOST tensor [R, 2] - the per-step result read-back then needs no D2H copy in the stream. Must be called before
        the first training step (the pointer is baked into the captured graphs). Returns False where the step does not
        end with the local flat optimizer (dp / sharded subclasses publish through their own D2H copy)."""
        if type(self)._optimizer_kernels is not FusedDeepFM._optimizer_kernels or self._graphs:
            return False
        self._result_ring = ring
        return ring is not None

    # ------------------------------------------------------------------------------------------------ public API
    def load_inputs(self, ids: torch.Tensor, vals: torch.Tensor, labels: Optional[torch.Tensor], slot_idx: int = 0,
                    non_blocking: bool = True):
        """Copy one batch (host pinned or device; ids int32 or int64) into input slot `slot_idx`."""
        s = self.in_slots[slot_idx]
        n = ids.shape[0]
        if ids.dtype != torch.int32:
            ids = ids.to(torch.int32)
        s["ids"][:n].copy_(ids, non_blocking=non_blocking)
        s["vals"][:n].copy_(vals, non_blocking=non_blocking)
        if labels is not None:
            s["labels"][:n].copy_(labels.reshape(-1), non_blocking=non_blocking)
        if n < self.B:  # pad: id 0 with value 0 contributes nothing
            s["ids"][n:].zero_()
            s["vals"][n:].zero_()
            s["labels"][n:].zero_()
        return n

    def load_packed(self, packed: torch.Tensor, slot_idx: int = 0, non_blocking: bool = True):
        """Copy one full batch in packed slot layout (see make_packed_slot / the native loader) with a single copy."""
        self.in_slots[slot_idx]["packed"].copy_(packed, non_blocking=non_blocking)
        return self.B

    def train_step_slot(self, slot_idx: int = 0) -> torch.Tensor:
        """Run one optimizer step on the batch already resident in input slot `slot_idx`.
        Returns the device scalar `loss_sum` holder; call `last_loss()` for the reference-defined loss."""
        if self.use_graph:
            g = self._graphs.get(slot_idx)
            if g is None:
                # warm-up outside capture (loads modules, sets func attributes) on a state snapshot, then capture
                snap = self._snapshot()
                self._train_kernels(self.in_slots[slot_idx])
                torch.cuda.synchronize()
                self._restore(snap)
                g = torch.cuda.CUDAGraph()
                # capture on a high-priority stream: the kernel nodes of the critical chain inherit it, the forked side
                # branches (wgrad, untouched-row sweep) keep the default priority and yield SMs to the chain
                if getattr(self, "_cap_stream", None) is None:
                    self._cap_stream = torch.cuda.Stream(device=self.dev, priority=-1)
                with torch.cuda.graph(g, stream=self._cap_stream):
                    self.kernels_per_step = self._train_kernels(self.in_slots[slot_idx])
                self._graphs[slot_idx] = g
            g.replay()
        else:
            self.kernels_per_step = self._train_kernels(self.in_slots[slot_idx])
        self.host_step += 1
        return self.loss_sum

    def _snapshot(self):
        return [t.clone() if t is not None else None for t in (self.P, self.S0, self.S1, self.step_t, self.shadow)]

    def _restore(self, snap):
        for dst, src in zip((self.P, self.S0, self.S1, self.step_t, self.shadow), snap):
            if dst is not None:
                dst.copy_(src)
        self.G.zero_()
        self.done_counter.zero_()
        if self.split is not None:
            self.bm_own.zero_()
            self.reg_acc.zero_()
        torch.cuda.synchronize()

    def train_step(self, ids, vals, labels, slot_idx: int = 0) -> torch.Tensor:
        self.load_inputs(ids, vals, labels, slot_idx)
        return self.train_step_slot(slot_idx)

    def last_loss(self) -> torch.Tensor:
        """mean log-loss over the global batch + L2 terms, as the reference defines `loss` (PS:244-246)."""
        torch.add(self.loss_sum / self.global_batch, self.reg_loss, out=self.loss_out)
        return self.loss_out

    @torch.no_grad()
    def predict(self, ids, vals) -> torch.Tensor:
        n = self.load_inputs(ids, vals, None, 0)
        s = self.in_slots[0]
        self._forward(s, False)
        top = self.n_layers - 1
        K.logit_loss(self.acts[top], self._w("deep_out.weight"), self._w("deep_out.bias"), self._w("fm_bias"), self.yfm,
                     None, 1.0, 1.0, False, self.prob, self.logit)
        return self.prob[:n]

    @torch.no_grad()
    def predict_slot(self, n: int, slot_idx: int = 0) -> torch.Tensor:
        """Forward pass over the batch already resident in input slot `slot_idx` (rows >= n are padding), replayed from a
        CUDA graph captured on first use: emb_fm_fwd -> mlp_fwd x L -> logit head, one graph launch per call. The serving
        path (serving/predictor.py) keeps one engine per batch-size bucket so a 1-row request does not pay for 4096."""
        g = self.__dict__.get("_pgraphs", {}).get(slot_idx)
        if g is None:
            s = self.in_slots[slot_idx]
            top = self.n_layers - 1

            def fwd():
                self._forward(s, False)
                K.logit_loss(self.acts[top], self._w("deep_out.weight"), self._w("deep_out.bias"), self._w("fm_bias"),
                             self.yfm, None, 1.0, 1.0, False, self.prob, self.logit)
            fwd()  # warm-up outside capture (module load, function attributes)
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                fwd()
            self.__dict__.setdefault("_pgraphs", {})[slot_idx] = g
        g.replay()
        return self.prob[:n]

    @torch.no_grad()
    def eval_step(self, ids, vals, labels, update_auc: bool = True):
        """Forward + loss (+ AUC histogram). Returns (sum of per-example log-loss [1], probs [n])."""
        n = self.load_inputs(ids, vals, labels, 0)
        s = self.in_slots[0]
        self._forward(s, False)
        top = self.n_layers - 1
        self.loss_sum.zero_()
        x = self.acts[top][:n]
        K.logit_loss(x, self._w("deep_out.weight"), self._w("deep_out.bias"), self._w("fm_bias"), self.yfm[:n],
                     s["labels"][:n], 1.0, 1.0, False, self.prob[:n], self.logit[:n], None, None, self.loss_sum)
        if update_auc:
            K.auc_hist(self.prob[:n], s["labels"][:n], self.auc_thr, self.auc_hist)
        return self.loss_sum.clone(), self.prob[:n]

    def l2_penalty(self) -> torch.Tensor:
        if not self.layout.include_embeddings:
            return torch.zeros((), device=self.dev)
        return 0.5 * self.cfg.l2_reg * (self._w("fm_w").pow(2).sum() + self._w("fm_v").pow(2).sum())

    # ------------------------------------------------------------------------------------------------ state
    @property
    def global_step(self) -> int:
        return int(self.step_t.item())

    def named_parameters(self) -> "OrderedDict[str, torch.Tensor]":
        return OrderedDict((n, self.layout.view(self.P, n)) for n in self.layout.names())

    def named_slots(self) -> "OrderedDict[str, List[torch.Tensor]]":
        out = OrderedDict()
        for n in self.layout.names(trainable_only=True):
            sl = []
            if self.S0 is not None:
                sl.append(self.layout.view(self.S0, n))
            if self.S1 is not None:
                sl.append(self.layout.view(self.S1, n))
            out[n] = sl
        return out

    def state_dict(self) -> Dict[str, torch.Tensor]:
        sd = OrderedDict()
        for n, t in self.named_parameters().items():
            sd[n] = t.detach().clone()
        for n, sl in self.named_slots().items():
            for suffix, t in zip(SLOT_SUFFIXES[self.opt.name], sl):
                sd["%s/%s" % (n, suffix)] = t.detach().clone()
        sd["global_step"] = self.step_t.detach().clone()
        return sd

    def load_state_dict(self, sd: Dict[str, torch.Tensor], strict: bool = True):
        for n, t in self.named_parameters().items():
            if n in sd:
                t.copy_(sd[n].to(self.dev))
            elif strict:
                raise KeyError("missing parameter %s" % n)
        for n, sl in self.named_slots().items():
            for suffix, t in zip(SLOT_SUFFIXES[self.opt.name], sl):
                key = "%s/%s" % (n, suffix)
                if key in sd:
                    t.copy_(sd[key].to(self.dev))
        if "global_step" in sd:
            self.step_t.copy_(sd["global_step"].to(self.dev).reshape(1))
            self.opt.step = int(self.step_t.item())
        self.refresh_shadow()
        self.G.zero_()
