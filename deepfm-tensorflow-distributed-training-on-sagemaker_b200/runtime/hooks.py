"""Training hooks: what tf.estimator's default hooks do for the reference — loss/step logging with
steps-per-second, periodic checkpoints (every 600 s, keep 5), JSONL metrics — plus `log_steps`, which the reference
declares but never reads (PS:47)."""
from __future__ import annotations

import json
import os
import sys
import time
from typing import Optional


class LoggingHook:
    def __init__(self, every_steps: int, global_batch: int, metrics_file: Optional[str] = None, out=sys.stdout,
                 enabled: bool = True, event_dir: Optional[str] = None):
        """`event_dir`: also write TensorBoard event files the way the Estimator does - training scalars into
        `<event_dir>/events.out.tfevents.*`, evaluation metrics into `<event_dir>/eval/` (utils/tfevents.py)."""
        self.every = max(1, int(every_steps))
        self.gb = global_batch
        self.out, self.enabled = out, enabled
        self.metrics_file = metrics_file
        self.event_dir = event_dir if enabled else None
        self._ev_train = self._ev_eval = None
        self._t0 = None
        self._s0 = 0
        if metrics_file and enabled:
            os.makedirs(os.path.dirname(os.path.abspath(metrics_file)), exist_ok=True)

    def _events(self, which: str):
        if not self.event_dir:
            return None
        from ..utils.tfevents import EventFileWriter
        if which == "train":
            if self._ev_train is None:
                self._ev_train = EventFileWriter(self.event_dir)
            return self._ev_train
        if self._ev_eval is None:
            self._ev_eval = EventFileWriter(os.path.join(self.event_dir, "eval"))
        return self._ev_eval

    def close(self):
        for w in (self._ev_train, self._ev_eval):
            if w is not None:
                w.close()
        self._ev_train = self._ev_eval = None

    def begin(self, step: int):
        self._t0, self._s0 = time.time(), step

    def after_step(self, step: int, get_loss):
        """`get_loss()` is only called when a line is due: reading the loss forces a device sync."""
        if not self.enabled or step % self.every != 0:
            return
        loss = float(get_loss())
        now = time.time()
        dt = max(now - (self._t0 or now), 1e-9)
        sps = (step - self._s0) / dt
        print("INFO:deepfm_b200:loss = %.8g, step = %d (%.3f sec)" % (loss, step, dt), file=self.out)
        print("INFO:deepfm_b200:global_step/sec: %.4g  examples/sec: %.6g" % (sps, sps * self.gb), file=self.out)
        self.out.flush()
        if self.metrics_file:
            with open(self.metrics_file, "a") as f:
                f.write(json.dumps({"step": step, "loss": loss, "steps_per_sec": sps, "examples_per_sec": sps * self.gb,
                                    "step_time_ms": 1e3 / max(sps, 1e-9), "time": now}) + "\n")
        ev = self._events("train")
        if ev is not None:
            ev.add_scalars({"loss": loss, "global_step/sec": sps, "examples/sec": sps * self.gb}, step, now)
            ev.flush()
        self._t0, self._s0 = now, step

    def log_eval(self, metrics: dict):
        if not self.enabled:
            return
        print("INFO:deepfm_b200:Saving dict for global step %d: %s" % (
            metrics.get("global_step", -1), ", ".join("%s = %.6g" % (k, v) for k, v in sorted(metrics.items()))),
            file=self.out)
        self.out.flush()
        if self.metrics_file:
            with open(self.metrics_file, "a") as f:
                f.write(json.dumps(dict(metrics, eval=True, time=time.time())) + "\n")
        ev = self._events("eval")
        if ev is not None:
            ev.add_scalars({k: float(v) for k, v in metrics.items() if k in ("auc", "loss")},
                           int(metrics.get("global_step", 0)))
            ev.flush()


class CheckpointSaverHook:
    def __init__(self, model_dir: str, save_secs: int = 600, save_steps: int = 0, keep: int = 5, enabled: bool = True):
        self.dir, self.secs, self.steps, self.keep, self.enabled = model_dir, save_secs, save_steps, keep, enabled
        self._last_t = time.time()
        self._last_step = -1

    def due(self, step: int) -> bool:
        if not self.enabled or not self.dir:
            return False
        if self.steps > 0:
            return step % self.steps == 0 and step != self._last_step
        return (time.time() - self._last_t) >= self.secs

    def save(self, trainer, out=sys.stdout) -> Optional[str]:
        if not self.enabled or not self.dir:
            return None
        from ..utils.checkpoint import save_checkpoint
        sd = trainer.state_dict()
        step = int(sd["global_step"].reshape(-1)[0])
        if step == self._last_step:
            return None
        prefix = save_checkpoint(self.dir, sd, trainer.opt, self.keep)
        print("INFO:deepfm_b200:Saving checkpoints for %d into %s." % (step, prefix), file=out)
        self._last_t, self._last_step = time.time(), step
        return prefix
