"""Estimator — the train / evaluate / predict / export driver with the reference's Estimator semantics
(PS:435-467, HVD:365-431): a `model_dir` that is restored from on entry, checkpoints written by the chief only,
`evaluate` returning {auc, loss, global_step}, `predict` yielding probabilities, `export_savedmodel` writing a
servable with the raw serving signature (feat_ids [None,F] int64, feat_vals [None,F] float32 -> prob)."""
from __future__ import annotations

import collections
import os
import sys
import time
from typing import Callable, Iterable, Optional

from ..models import DeepFMConfig
from ..optim.tf1 import TF1Optimizer
from ..parallel.dist import DistEnv, agree_min
from ..utils.checkpoint import latest_checkpoint, load_checkpoint
from .hooks import CheckpointSaverHook, LoggingHook
from .trainer import Trainer


def _parse_fault(spec: str, rank: int):
    """Fault injection for the failure-detection / resume tests (the reference has none).
    DFM_FAULT_INJECT="<rank>:<step>[:<marker file>]": that rank dies with exit code 17 right after optimizer step
    <step>. With a marker file the fault fires once per job: the file is created first, and a process that finds it
    (the restarted attempt) runs normally."""
    if not spec:
        return None
    parts = spec.split(":", 2)
    try:
        r, step = int(parts[0]), int(parts[1])
    except (ValueError, IndexError):
        raise SystemExit("DFM_FAULT_INJECT must be <rank>:<step>[:<marker file>], got %r" % spec)
    marker = parts[2] if len(parts) > 2 else ""
    if r != rank or (marker and os.path.exists(marker)):
        return None
    return (step, marker)


def _fire_fault(fault, step: int, rank: int) -> None:
    if fault[1]:
        with open(fault[1], "w") as f:
            f.write("rank %d died at step %d\n" % (rank, step))
    sys.stderr.write("ERROR:deepfm_b200: injected fault: rank %d exits at step %d\n" % (rank, step))
    sys.stderr.flush()
    os._exit(17)


class Estimator:
    def __init__(self, cfg: DeepFMConfig, batch_size: int, optimizer: TF1Optimizer, model_dir: Optional[str],
                 env: Optional[DistEnv] = None, backend: str = "auto", seed: int = 0, log_steps: int = 100,
                 save_checkpoints_secs: int = 600, save_checkpoints_steps: int = 0, keep_checkpoint_max: int = 5,
                 metrics_file: Optional[str] = None, comm: str = "auto", strategy: str = "dp", ps_staleness: int = 4,
                 out=sys.stdout):
        self.cfg, self.B, self.opt = cfg, batch_size, optimizer
        self.env = env or DistEnv()
        # like the reference, only the chief owns a model_dir (HVD:365-368); other ranks get state by broadcast
        self.model_dir = model_dir if (model_dir and self.env.is_chief) else None
        self.shared_model_dir = model_dir
        self.out = out
        self.strategy = strategy
        if strategy == "ps":
            from ..parallel.ps import AsyncPSTrainer
            self.trainer = AsyncPSTrainer(cfg, batch_size, optimizer, seed, self.env, backend=backend,
                                          staleness=ps_staleness)
        else:
            self.trainer = Trainer(cfg, batch_size, optimizer, seed, self.env.device, backend, self.env.world_size,
                                   self.env.rank, comm=comm, strategy=strategy)
        if self.env.is_chief:  # the transport actually selected (reference: NCCL_DEBUG=INFO in the launcher, NB-HVD:90)
            print("INFO:deepfm_b200:communication: %s" % getattr(self.trainer, "comm_name", "none"), file=out)
        gb = batch_size * self.env.world_size
        mf = metrics_file or (os.path.join(model_dir, "metrics.jsonl") if model_dir else None)
        self.log = LoggingHook(log_steps, gb, mf if self.env.is_chief else None, out, enabled=self.env.is_chief,
                               event_dir=model_dir if self.env.is_chief else None)
        self.saver = CheckpointSaverHook(self.model_dir or "", save_checkpoints_secs, save_checkpoints_steps,
                                         keep_checkpoint_max, enabled=self.env.is_chief and bool(model_dir))
        self._restored = False
        self._fault = _parse_fault(os.environ.get("DFM_FAULT_INJECT", ""), self.env.rank)

    # ------------------------------------------------------------------------------------------ state
    def restore_latest(self) -> Optional[str]:
        """Estimator behaviour: every train/evaluate/predict call starts from the latest checkpoint in model_dir.
        The chief reads it; with several ranks the state is then broadcast from rank 0 (the reference does this
        with BroadcastGlobalVariablesHook(0), HVD:372)."""
        prefix = latest_checkpoint(self.model_dir) if self.model_dir else None
        if prefix:
            names = self.trainer.parameter_names() if hasattr(self.trainer, "parameter_names") else \
                list(self.trainer.model.named_parameters().keys())
            self.trainer.load_state_dict(load_checkpoint(prefix, names, self.opt.name), strict=False)
            print("INFO:deepfm_b200:Restoring parameters from %s" % prefix, file=self.out)
        if self.env.world_size > 1 and hasattr(self.trainer, "broadcast_state"):
            self.trainer.broadcast_state()
        self._restored = True
        return prefix

    # ------------------------------------------------------------------------------------------ train
    def train(self, input_fn: Callable[[], Iterable], max_steps: int = 0, steps_hint: Optional[int] = None) -> int:
        """One `Estimator.train` call: consume `input_fn()` to exhaustion (or `max_steps`)."""
        self.restore_latest()
        tr = self.trainer
        step = tr.global_step
        self.log.begin(step)
        fused = getattr(tr, "backend", "") == "fused"
        inflight = collections.deque()
        n_local = 0
        limit = agree_min(self.env, steps_hint) if steps_hint is not None else None
        batches = input_fn()
        if limit is None and self.env.world_size > 1 and self.strategy != "ps" and self.env.backend != "none":
            # no step count known up front (pipe mode: a FIFO cannot be counted): agree block by block instead
            batches = self._agreed_batches(batches)
        for batch in batches:
            if limit is not None and n_local >= limit:
                batch.release()
                break
            if batch.rows != self.B:  # drop_remainder=True upstream; a short batch would change the graph shape
                batch.release()
                continue
            if fused and batch.raw is not None:
                ev = tr.train_raw(batch.raw, batch.rows)
            elif fused and batch.packed is not None:
                ev = tr.train_packed(batch.packed)
            else:
                ev = tr.train_batch(batch.ids, batch.vals, batch.labels)
            inflight.append((batch, ev))
            while inflight and (len(inflight) > 3 or (inflight[0][1] is not None and inflight[0][1].query())):
                b, e = inflight.popleft()
                if e is not None:
                    e.synchronize()
                b.release()
            step += 1
            n_local += 1
            if self._fault is not None and step >= self._fault[0]:
                _fire_fault(self._fault, step, self.env.rank)
            self.log.after_step(step, lambda: self._current_loss())
            if self.saver.due(step):
                self.saver.save(tr, self.out)
            if max_steps and step >= max_steps:
                break
        for b, e in inflight:
            if e is not None:
                e.synchronize()
            b.release()
        tail = tr.drain()
        if tail:
            self._last_loss = tail[-1]
        if self.env.is_chief and self.saver.enabled:
            self.saver.save(tr, self.out)
        print("INFO:deepfm_b200:Loss for final step: %.8g." % self._current_loss(), file=self.out)
        return n_local

    def _agreed_batches(self, batches: Iterable, block: int = 3):
        """Synchronous strategies on several ranks without a step count known up front (pipe mode, HVD:396-405): the
        ranks agree every `block` steps on how many of the next `block` full batches ALL of them hold, and stop together
        at the first short block. A rank whose FIFO holds fewer records can therefore never leave the others waiting in
        a collective (the Horovod failure described at PDF p22-23)."""
        it = iter(batches)
        while True:
            buf = []
            while len(buf) < block:
                b = next(it, None)
                if b is None:
                    break
                if b.rows != self.B:  # drop_remainder: short batches are not steps
                    b.release()
                    continue
                buf.append(b)
            n = agree_min(self.env, len(buf))
            for b in buf[n:]:
                b.release()
            for b in buf[:n]:
                yield b
            if n < block:
                return

    def _current_loss(self) -> float:
        tr = self.trainer
        losses = tr.drain()
        if losses:
            self._last_loss = losses[-1]
        return float(getattr(self, "_last_loss", float("nan")))

    # ------------------------------------------------------------------------------------------ eval / predict
    def evaluate(self, input_fn: Callable[[], Iterable]) -> dict:
        if not self._restored:
            self.restore_latest()
        t0 = time.time()
        m = self.trainer.evaluate(self._tuples(input_fn()))
        m["eval_seconds"] = time.time() - t0
        self.log.log_eval(m)
        return m

    def predict(self, input_fn: Callable[[], Iterable]):
        if not self._restored:
            self.restore_latest()
        for probs in self.trainer.predict(self._tuples(input_fn())):
            for p in probs.tolist():
                yield {"prob": p}

    @staticmethod
    def _tuples(batches):
        for b in batches:
            t = (b.ids.clone(), b.vals.clone(), b.labels.clone())
            b.release()
            yield t

    def export_savedmodel(self, export_dir_base: str) -> Optional[str]:
        if not self.env.is_chief:
            return None
        if not self._restored:
            self.restore_latest()
        from ..serving.export import export_servable
        params = {k: v.detach().float().cpu() for k, v in self.trainer.state_dict().items() if "/" not in k
                  and k != "global_step"}
        path = export_servable(export_dir_base, self.cfg, params, self.trainer.global_step)
        print("INFO:deepfm_b200:SavedModel written to: %s" % path, file=self.out)
        return path
