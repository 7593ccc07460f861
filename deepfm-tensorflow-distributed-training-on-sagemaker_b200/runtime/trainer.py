"""Trainer — the public training/eval/predict API (the role `tf.estimator.Estimator` plays in the reference:
`.train / .evaluate / .predict`, PS:435-449 / HVD:365-413), bound to one GPU (one process per GPU).

    tr = Trainer(cfg, batch_size, TF1Optimizer("Adam", 5e-4), world_size=W, rank=r)
    tr.train_batch(ids, vals, labels)        # host tensors (pinned): async H2D on a copy stream, double-buffered
    tr.train_batch_device(ids, vals, labels) # tensors already on the GPU
    tr.drain() -> [loss, ...]                # losses of the steps since the last drain (async D2H ring)
    tr.evaluate(batches) -> {"auc", "loss", "global_step"};  tr.predict(batches) -> probs

Backends: 'fused' = sm_100a kernels + CUDA graph (+ fused all-reduce/optimizer kernel when world_size > 1);
'torch' = plain PyTorch fp32 (CPU-capable; gloo); 'proxy' = NCCL + cuBLAS + torch.optim stand-in for the
reference's Horovod build (baseline only).
"""
from __future__ import annotations

from typing import Dict, Iterable, List, Optional

import torch

from ..models import DeepFMConfig, TorchDeepFM
from ..optim.tf1 import TF1Optimizer
from ..utils.auc import StreamingAUC, auc_from_hist


class _CopyTicket:
    """What `train_packed` returns on the native step path: completes when the pinned source buffer of that step has
    been copied out (same `query()` / `synchronize()` surface as the torch Event of the Python path)."""
    __slots__ = ("_lib", "_h", "_slot")

    def __init__(self, lib, handle, slot):
        self._lib, self._h, self._slot = lib, handle, slot

    def query(self) -> bool:
        return self._lib.dfm_driver_copy_done(self._h, self._slot) != 0

    def synchronize(self) -> None:
        from ..ops import lib as _nlib
        _nlib.check(self._lib.dfm_driver_copy_wait(self._h, self._slot))


class Trainer:
    def __init__(self, cfg: DeepFMConfig, batch_size: int, optimizer: Optional[TF1Optimizer] = None, seed: int = 0,
                 device=None, backend: str = "auto", world_size: int = 1, rank: int = 0, use_graph: bool = True,
                 comm: str = "auto", params: Optional[Dict[str, torch.Tensor]] = None, process_group=None,
                 strategy: str = "dp"):
        self.cfg, self.B = cfg, batch_size
        self.world, self.rank = world_size, rank
        self.opt = optimizer or TF1Optimizer("Adam", 5e-4)
        if backend == "auto":
            backend = "fused" if torch.cuda.is_available() else "torch"
            why_not = cfg.fused_constraints() if backend == "fused" else []
            if why_not:  # shapes the sm_100a kernels do not take: run the same model on stock torch ops instead
                import sys as _sys
                print("WARNING:deepfm_b200: --backend auto falls back to torch ops: " + "; ".join(why_not), file=_sys.stderr)
                backend = "torch"
        self.backend = backend
        self.comm_name = "none"
        gb = batch_size * world_size
        ro = rank * batch_size
        if backend == "fused":
            self.dev = device or torch.device("cuda", torch.cuda.current_device())
            if strategy == "sharded":
                from ..parallel.sharded import ShardedDeepFM
                self._ensure_group(world_size, rank)
                self.model = ShardedDeepFM(cfg, batch_size, self.opt, seed, self.dev, world_size, rank, comm=comm,
                                           use_graph=use_graph, params=params, process_group=process_group)
                self.comm_name = self.model.comm_name
            elif world_size > 1:
                import os as _os0
                # default: the dense ZeRO-1 kernel (parallel/dp.py: fused reduce-scatter + sharded optimizer + all-gather);
                # DFM_DP_EXCHANGE=sparse selects the touched-row exchange + replicated split optimizer
                # (parallel/dp_sparse.py), which ships fewer bytes but pays random 128-byte row accesses in its tail:
                # measured slower at this table size (2 GPUs: 141 vs 126 us/step), kept for tables where a batch touches a
                # small fraction of the rows
                dense = comm == "nccl" or _os0.environ.get("DFM_DP_EXCHANGE", "dense") != "sparse"
                if dense:
                    from ..parallel.dp import FusedDataParallel as _DP
                else:
                    from ..parallel.dp_sparse import SparseDataParallel as _DP
                self.model = _DP(cfg, batch_size, self.opt, seed, self.dev, world_size, rank, comm=comm,
                                 use_graph=use_graph, params=params, process_group=process_group)
                self.comm_name = self.model.comm_name
            else:
                from ..models.fused import FusedDeepFM
                self.model = FusedDeepFM(cfg, batch_size, self.opt, seed, self.dev, params=params, use_graph=use_graph)
        elif backend == "proxy":
            from ..parallel.proxy_baseline import ProxyDeepFM
            self.dev = device or torch.device("cuda", torch.cuda.current_device())
            self.model = ProxyDeepFM(cfg, batch_size, self.opt, seed, self.dev, world_size, rank)
            self.comm_name = "nccl-allreduce" if world_size > 1 else "none"
        else:
            self.dev = torch.device(device if device is not None else "cpu")
            self.model = TorchDeepFM(cfg, batch_size, self.opt, seed, self.dev, params=params,
                                     global_batch=gb if world_size > 1 else None, row_offset=ro)
            if world_size > 1:
                from ..parallel.dp import install_torch_allreduce
                install_torch_allreduce(self.model, process_group)
                self.comm_name = "torch.distributed.all_reduce"
        import os as _os
        self.nvtx = bool(_os.environ.get("DFM_NVTX"))  # NVTX range per step for nsys/ncu timelines
        self._cuda = self.dev.type == "cuda"
        self._nslots = len(getattr(self.model, "in_slots", [])) or 1
        self._k = 0
        self._pending: List[int] = []
        if self._cuda:
            self._copy_stream = torch.cuda.Stream(device=self.dev)
            self._slot_free = [None] * self._nslots
            self._ring_n = 4096
            self._ring = torch.zeros(self._ring_n, 2, dtype=torch.float32).pin_memory()
            self._ring_pos = 0
        self._host_losses: List[float] = []
        # result publication: the optimizer kernel stores (loss sum, L2 term) of step t into ring[(t - 1) % ring_n]
        # itself, so the per-step D2H read-back is 8 bytes written by the kernel instead of a copy op between two graph
        # launches (which cost ~10 us/step of engine hand-offs). _pending then holds the model's step index at issue.
        self._publish = False
        if self._cuda and backend == "fused" and _os.environ.get("DFM_PUBLISH", "1") != "0" \
                and hasattr(self.model, "set_result_ring"):
            self._publish = bool(self.model.set_result_ring(self._ring))
        # native step driver (csrc/step_driver.cu): one C call per host-fed step once the slot's graph exists
        self._drv = None
        self.native_steps = 0
        # default on (DFM_NATIVE_STEP=0 keeps the Python-issued sequence); needs the step captured in a CUDA graph
        want = _os.environ.get("DFM_NATIVE_STEP", "1")
        if self._cuda and backend == "fused" and want != "0" and getattr(self.model, "use_graph", False):
            import ctypes
            from ..ops import lib as _nlib
            lib = _nlib.load()
            lib.dfm_driver_steps.restype = ctypes.c_ulonglong
            h = ctypes.c_void_p()
            _nlib.check(lib.dfm_driver_create(ctypes.byref(h), ctypes.c_int(self._nslots)))
            self._drv = (lib, h)

    def close(self) -> None:
        if self._drv is not None:
            if self._cuda:
                torch.cuda.synchronize(self.dev)
            lib, h = self._drv
            self._drv = None
            lib.dfm_driver_destroy(h)
        if hasattr(self.model, "close"):
            self.model.close()

    def _graph_exec(self, slot: int, g) -> int:
        cache = self.__dict__.setdefault("_exec_cache", {})
        e = cache.get(slot)
        if e is None or e[0] is not g:
            e = cache[slot] = (g, int(g.raw_cuda_graph_exec()))
        return e[1]

    def _native_step(self, packed: torch.Tensor):
        """train_packed through dfm_driver_step; None if this step has to take the Python path (graph not captured yet,
        pageable source, ring full)."""
        import ctypes
        m = self.model
        slot = self._k % self._nslots
        g = m._graphs.get(slot)
        if g is None or not packed.is_pinned() or self._ring_pos == self._ring_n:
            return None
        lib, h = self._drv
        dst = m.in_slots[slot]["packed"]
        nbytes = packed.numel() * packed.element_size()
        if nbytes != dst.numel() * dst.element_size() or not packed.is_contiguous():
            raise ValueError("train_packed: buffer does not match the packed slot layout (%d bytes)" % nbytes)
        cs = self._copy_stream
        if self._slot_free[slot] is not None:  # the slot was last used through the Python path
            cs.wait_event(self._slot_free[slot])
            self._slot_free[slot] = None
        main = torch.cuda.current_stream(self.dev)
        ring = self._ring[self._ring_pos]
        sc = m.scalars
        nscal = 0 if self._publish else min(sc.numel(), ring.numel()) * 4
        from ..ops import lib as _nlib
        _nlib.check(lib.dfm_driver_step(
            h, ctypes.c_int(slot), ctypes.c_void_p(packed.data_ptr()), ctypes.c_void_p(dst.data_ptr()),
            ctypes.c_ulonglong(nbytes), ctypes.c_void_p(self._graph_exec(slot, g)), ctypes.c_void_p(cs.cuda_stream),
            ctypes.c_void_p(main.cuda_stream), ctypes.c_void_p(sc.data_ptr()), ctypes.c_void_p(ring.data_ptr()),
            ctypes.c_ulonglong(nscal)))
        self._k += 1
        self._pending.append(m.host_step if self._publish else 0)
        m.host_step += 1
        self._ring_pos += 1
        self.native_steps += 1
        return _CopyTicket(lib, h, slot)

    @staticmethod
    def _ensure_group(world_size: int, rank: int):
        """Symmetric memory needs a process group even for a single rank."""
        import os
        if not torch.distributed.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29517")
            torch.distributed.init_process_group("nccl", rank=rank, world_size=world_size,
                                                 device_id=torch.device("cuda", torch.cuda.current_device()))

    # ------------------------------------------------------------------------------------------- training
    @property
    def kernels_per_step(self) -> int:
        return getattr(self.model, "kernels_per_step", 0)

    @property
    def global_step(self) -> int:
        return self.model.global_step

    def last_loss(self) -> torch.Tensor:
        return self.model.last_loss()

    def train_batch_device(self, ids, vals, labels):
        """One optimizer step on a batch that already lives on this trainer's device."""
        if self.backend == "fused":
            slot = self._k % self._nslots
            self._k += 1
            self.model.load_inputs(ids, vals, labels, slot)
            return self.model.train_step_slot(slot)
        return self.model.train_step(ids, vals, labels)

    def train_packed_device(self, packed: torch.Tensor):
        """One optimizer step on a device-resident batch in packed slot layout (one d2d copy + graph replay)."""
        slot = self._k % self._nslots
        self._k += 1
        self.model.load_packed(packed, slot)
        return self.model.train_step_slot(slot)

    def train_packed(self, packed: torch.Tensor):
        """One optimizer step on a HOST batch in packed slot layout (what the native loader produces): a single
        pinned async H2D copy on the copy stream, graph replay, async D2H of the loss scalars."""
        if self._drv is not None and not self.nvtx:
            t = self._native_step(packed)
            if t is not None:
                return t
        return self._train_host(lambda slot: self.model.load_packed(packed, slot, non_blocking=True))

    def train_raw(self, raw: torch.Tensor, rows: int):
        """One optimizer step on a HOST batch of SERIALIZED Examples (raw loader slot: offsets | record bytes): one pinned
        H2D copy of the raw bytes, decoded on the GPU by example_parse straight into the input slot, then the step."""
        from ..ops import kernels as K
        if rows != self.B:
            raise ValueError("train_raw needs full batches (drop_remainder=True)")
        if not hasattr(self, "_raw_dev"):
            self._raw_dev = [torch.zeros(raw.numel(), dtype=torch.uint8, device=self.dev) for _ in range(self._nslots)]
            self._parse_err = torch.zeros(1, dtype=torch.int32, device=self.dev)
            self._raw_off = ((self.B + 1) * 4 + 255) & ~255
        F = self.cfg.field_size

        def load(slot):
            d = self._raw_dev[slot]
            d.copy_(raw, non_blocking=True)
            s = self.model.in_slots[slot]
            K.example_parse(d[self._raw_off:], d[: (self.B + 1) * 4].view(torch.int32), self.B, F, s["labels"], s["ids"],
                            s["vals"], self._parse_err)
        return self._train_host(load)

    def check_parse_errors(self) -> None:
        if hasattr(self, "_parse_err"):
            e = int(self._parse_err.item())
            if e:
                raise ValueError("GPU Example parser rejected record %d of a batch (schema mismatch or malformed proto)" % (e - 1))

    def train_batch(self, ids, vals, labels):
        """One optimizer step on a HOST batch. On CUDA: pinned-memory async H2D on a side stream into the next input
        slot (overlapping the previous step), graph replay, async D2H of the step's loss scalars."""
        if self.backend != "fused":
            loss = self.model.train_step(ids.to(self.dev, non_blocking=True), vals.to(self.dev, non_blocking=True),
                                         labels.to(self.dev, non_blocking=True))
            if self._cuda:
                self._push_loss(torch.stack([loss.reshape(()), torch.zeros((), device=self.dev)]), already_mean=True)
            else:
                self._host_losses.append(float(loss))
            return None
        return self._train_host(lambda slot: self.model.load_inputs(ids, vals, labels, slot, non_blocking=True))

    def _train_host(self, load_fn):
        if self.nvtx:
            torch.cuda.nvtx.range_push("deepfm_b200.step")
        try:
            return self._train_host_impl(load_fn)
        finally:
            if self.nvtx:
                torch.cuda.nvtx.range_pop()

    def _train_host_impl(self, load_fn):
        main = torch.cuda.current_stream(self.dev)
        slot = self._k % self._nslots
        self._k += 1
        cs = self._copy_stream
        if self._publish and self._ring_pos == self._ring_n:
            self.drain(keep=True)  # before issuing: the new step's record would land on the oldest unread one
        if self._slot_free[slot] is not None:
            cs.wait_event(self._slot_free[slot])  # the step that last read this slot has finished
        elif self.native_steps:
            cs.wait_stream(main)  # the slot's last reader was issued by the native driver (its events live there)
        with torch.cuda.stream(cs):
            load_fn(slot)
            ev = torch.cuda.Event()
            ev.record(cs)
        main.wait_event(ev)
        issued = getattr(self.model, "host_step", 0)
        self.model.train_step_slot(slot)
        done = torch.cuda.Event()
        done.record(main)
        self._slot_free[slot] = done
        if self._publish:
            self._pending.append(issued)
            self._ring_pos += 1
        else:
            self._push_loss(self.model.scalars, already_mean=False)
        return ev  # completes when the host buffer has been copied out (the caller may then reuse it)

    def _push_loss(self, scalars: torch.Tensor, already_mean: bool):
        if self._ring_pos == self._ring_n:
            self.drain(keep=True)
        self._ring[self._ring_pos].copy_(scalars, non_blocking=True)
        self._pending.append(1 if already_mean else 0)
        self._ring_pos += 1

    def drain(self, keep: bool = False) -> List[float]:
        """Wait for outstanding steps; return the per-step losses (mean log-loss + L2 terms, PS:244-246)."""
        if self._cuda:
            torch.cuda.synchronize(self.dev)
            self.check_parse_errors()
            reg_now = None
            # the fused engine reports a log-loss SUM: over the global batch where the model exchanges it (the fused dp
            # kernels), else over this rank's batch (single GPU; the NCCL baseline path: a per-rank estimate)
            denom = float(self.model.global_batch) if getattr(self.model, "loss_is_global", False) else float(self.B)
            # published results sit at ring[(t - 1) % n] for optimizer step t = step index at issue + 1 + (device step
            # counter - host step count), the last term being the global_step restored from a checkpoint
            shift = (self.model.global_step - self.model.host_step) if (self._publish and self._pending) else 0
            for i, mean_flag in enumerate(self._pending):
                if self._publish:
                    pos = (mean_flag + shift) % self._ring_n
                    a, b, mean_flag = float(self._ring[pos, 0]), float(self._ring[pos, 1]), 0
                else:
                    a, b = float(self._ring[i, 0]), float(self._ring[i, 1])
                if reg_now is not None:
                    b = reg_now
                # fused: (local sum of log-loss, L2 term). Under DP the local sum is scaled to the local batch (an
                # unbiased per-rank estimate of the global mean; the exact mean needs the other ranks' sums)
                self._host_losses.append(a if mean_flag else a / denom + b)
            self._pending = []
            self._ring_pos = 0
        if keep:
            return []
        out, self._host_losses = self._host_losses, []
        return out

    # ------------------------------------------------------------------------------------------- eval / predict
    def evaluate(self, batches: Iterable, distributed: bool = False) -> Dict[str, float]:
        """`Estimator.evaluate` (PS:443-444): mean loss (incl. the L2 terms, as the reference's EVAL loss does) and
        streaming AUC with tf.metrics.auc defaults. The reference evaluates on the chief only (HVD:393); pass
        `distributed=True` when EVERY rank calls this on its own shard: loss sums and the 2x201 confusion histogram
        are then combined with one small all-reduce."""
        auc = StreamingAUC()
        fused = self.backend == "fused"
        if fused:
            self.model.auc_hist.zero_()
        tot = torch.zeros(1, device=self.dev, dtype=torch.float64)
        n = 0
        for ids, vals, labels in batches:
            ids, vals, labels = ids.to(self.dev), vals.to(self.dev), labels.to(self.dev)
            loss_sum, probs = self.model.eval_step(ids, vals, labels)
            tot += loss_sum.double()
            n += ids.shape[0]
            if not fused:
                auc.update(labels, probs)
        hist = self.model.auc_hist.double().cpu() if fused else auc.hist
        stats = torch.cat([tot.cpu(), torch.tensor([float(n)], dtype=torch.float64), hist.reshape(-1)])
        if distributed and self.world > 1 and torch.distributed.is_initialized():
            s = stats.to(self.dev) if torch.distributed.get_backend() == "nccl" else stats
            torch.distributed.all_reduce(s)
            stats = s.cpu()
        tot_f, n_f = float(stats[0]), float(stats[1])
        hist = stats[2:].reshape(2, -1)
        return {"loss": tot_f / max(n_f, 1.0) + float(self.model.l2_penalty()), "auc": auc_from_hist(hist),
                "global_step": self.global_step, "examples": int(n_f)}

    def predict(self, batches: Iterable):
        """`Estimator.predict(predict_keys="prob")` (PS:445-446): yields a 1-D tensor of probabilities per batch."""
        for b in batches:
            ids, vals = b[0].to(self.dev), b[1].to(self.dev)
            yield self.model.predict(ids, vals).detach().float().cpu()

    # ------------------------------------------------------------------------------------------- state
    def broadcast_state(self, src: int = 0):
        """Make every rank's parameters / slots / step equal to rank `src`'s (HVD:372)."""
        if self.world > 1 and torch.distributed.is_initialized():
            if hasattr(self.model, "broadcast_state"):
                self.model.broadcast_state(src)
            else:
                from ..parallel.dp import broadcast_model
                broadcast_model(self.model, src)

    def state_dict(self):
        return self.model.state_dict()

    def load_state_dict(self, sd, strict: bool = True):
        self.model.load_state_dict(sd, strict)
