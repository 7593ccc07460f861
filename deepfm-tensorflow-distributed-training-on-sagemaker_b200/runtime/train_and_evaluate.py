"""`tf.estimator.train_and_evaluate` semantics for the parameter-server mode (reference: PS:439-442 —
`TrainSpec(input_fn, max_steps)` + `EvalSpec(input_fn, steps=None, start_delay_secs=1000, throttle_secs=1200)`; README-EN:38
makes train_and_evaluate mandatory there; PDF p34 shows the master saving a checkpoint, then evaluating it).

While the workers keep training asynchronously, the chief evaluates every NEW checkpoint that appears in `model_dir`: not
before `start_delay_secs` after the start, at most once per `throttle_secs`, and once more on the final checkpoint when
training ends. Evaluation never touches the live parameter store: it loads the checkpoint files into its own scoring engine
(serving.Predictor: the fused sm_100a forward on a GPU, plain PyTorch otherwise), so it measures exactly what a restart or an
export would see, and it cannot stall the workers.
"""
from __future__ import annotations

import threading
import time
from typing import Callable, Iterable, List, Optional

import torch

from ..models import DeepFMConfig
from ..models.params import native_names
from ..utils.auc import StreamingAUC
from ..utils.checkpoint import latest_checkpoint, load_checkpoint


def evaluate_checkpoint(cfg: DeepFMConfig, prefix: str, opt_name: str, batches: Iterable, scorer=None, backend: str = "auto",
                        max_batch: int = 4096):
    """{loss, auc, global_step, examples} of checkpoint `prefix` over `batches` of (ids, vals, labels); returns
    (metrics, scorer) so the caller can reuse the engine for the next checkpoint."""
    from ..serving.predictor import Predictor
    names = [n for n, _, _ in native_names(cfg)]
    sd = load_checkpoint(prefix, names, opt_name)
    step = int(sd["global_step"].reshape(-1)[0]) if "global_step" in sd else -1
    params = {n: sd[n] for n in names if n in sd}
    if scorer is None:
        scorer = Predictor(cfg, params, max_batch=max_batch, backend=backend)
    else:
        scorer.set_params(params)
    auc, tot, n = StreamingAUC(), 0.0, 0
    for b in batches:
        ids, vals, labels = b[0], b[1], b[2]
        p = scorer.predict({"feat_ids": ids, "feat_vals": vals})["prob"].double().clamp(1e-7, 1 - 1e-7)
        lab = torch.as_tensor(labels).double().reshape(-1).cpu()
        tot += float(-(lab * p.log() + (1 - lab) * (1 - p).log()).sum())
        auc.update(lab.float(), p.float())
        n += int(lab.numel())
        if hasattr(b, "release"):
            b.release()
    loss = tot / max(n, 1)
    if cfg.l2_mode == "exact" and cfg.l2_reg > 0:  # the reference's EVAL loss carries the whole-table L2 terms (PS:244-246)
        loss += 0.5 * cfg.l2_reg * float(params["fm_w"].double().pow(2).sum() + params["fm_v"].double().pow(2).sum())
    return {"loss": loss, "auc": auc.result(), "global_step": step, "examples": n}, scorer


class ThrottledCheckpointEvaluator(threading.Thread):
    def __init__(self, cfg: DeepFMConfig, model_dir: str, opt_name: str, eval_input_fn: Callable[[], Iterable],
                 on_metrics: Callable[[dict], None], start_delay_secs: float = 1000.0, throttle_secs: float = 1200.0,
                 backend: str = "auto", poll_secs: float = 1.0):
        super().__init__(name="deepfm-evaluator", daemon=True)
        self.cfg, self.model_dir, self.opt_name = cfg, model_dir, opt_name
        self.eval_input_fn, self.on_metrics = eval_input_fn, on_metrics
        self.start_delay, self.throttle, self.backend, self.poll = float(start_delay_secs), float(throttle_secs), backend, poll_secs
        self._stop_ev = threading.Event()
        self._scorer = None
        self._last_prefix: Optional[str] = None
        self.results: List[dict] = []
        self.error: Optional[BaseException] = None

    def _evaluate_latest(self) -> bool:
        prefix = latest_checkpoint(self.model_dir)
        if not prefix or prefix == self._last_prefix:
            return False
        t0 = time.time()
        m, self._scorer = evaluate_checkpoint(self.cfg, prefix, self.opt_name, self.eval_input_fn(), self._scorer, self.backend)
        m["eval_seconds"] = time.time() - t0
        self._last_prefix = prefix
        self.results.append(m)
        self.on_metrics(m)
        return True

    def run(self) -> None:
        try:
            if self._stop_ev.wait(self.start_delay):  # training ended before the first evaluation was due
                return
            while not self._stop_ev.is_set():
                evaluated = self._evaluate_latest()
                # throttle: after an evaluation wait the full period; with nothing new, poll for the next checkpoint
                if self._stop_ev.wait(self.throttle if evaluated else self.poll):
                    return
        except BaseException as e:  # surfaced by finish()
            self.error = e

    def finish(self, final_eval: bool = True) -> List[dict]:
        """Stop the loop; evaluate the final checkpoint if it has not been evaluated yet (train_and_evaluate does)."""
        self._stop_ev.set()
        self.join()
        if self.error is not None:
            raise self.error
        if final_eval:
            self._evaluate_latest()
        return self.results
