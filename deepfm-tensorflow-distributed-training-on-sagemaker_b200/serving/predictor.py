"""Inference entrypoint for an exported servable (the role TensorFlow Serving plays for the reference's SavedModel).

    from deepfm_b200.serving import Predictor
    p = Predictor.load("/opt/ml/model")                     # newest timestamped export
    out = p.predict({"feat_ids": ids, "feat_vals": vals})   # -> {"prob": float32 [n]}

On a B200 the fused forward kernels run (emb_fm_fwd -> tcgen05 MLP -> logit head); elsewhere plain PyTorch.
CLI:  python -m deepfm_b200.serving.predictor --model_dir DIR --input te.tfrecords --output pred.txt
"""
from __future__ import annotations

import argparse
import json
import os
from typing import Dict

import torch

from ..models import DeepFMConfig, TorchDeepFM
from ..models.params import native_names
from ..utils.checkpoint import tf_to_native
from ..utils.tf_bundle import read_bundle
from .export import latest_export


class Predictor:
    """`buckets`: batch sizes for which a graph-captured fused forward is kept (one engine each, built on first use). A
    request of n rows runs on the smallest bucket >= n (chunks of `max_batch` beyond that): inputs go through ONE pinned
    host buffer and one H2D copy, the forward is one CUDA-graph launch, the probabilities come back with one D2H copy into
    pinned memory and a single event wait."""

    def __init__(self, cfg: DeepFMConfig, params: Dict[str, torch.Tensor], max_batch: int = 4096, backend: str = "auto",
                 buckets=(64, 512)):
        self.cfg, self.max_batch = cfg, max_batch
        use_fused = backend == "fused" or (backend == "auto" and torch.cuda.is_available() and not cfg.fused_constraints())
        self.fused = bool(use_fused)
        if use_fused:
            self._params = params
            self._buckets = sorted({int(b) for b in buckets if 0 < int(b) < max_batch} | {int(max_batch)})
            self._engines = {}
            self.dev = torch.device("cuda", torch.cuda.current_device())
            self.model = self._engine(self._buckets[-1])[0]
        else:
            self.dev = torch.device("cuda" if (torch.cuda.is_available() and backend != "torch-cpu") else "cpu")
            self.model = TorchDeepFM(cfg, max_batch, device=self.dev, params=params)

    def _engine(self, B: int):
        e = self._engines.get(B)
        if e is None:
            from ..models.fused import FusedDeepFM, make_packed_slot
            from ..optim.tf1 import TF1Optimizer
            m = FusedDeepFM(self.cfg, B, TF1Optimizer("GD", 0.0), params=self._params, use_graph=False, input_slots=1)
            host_in = make_packed_slot(B, self.cfg.field_size, "cpu", pin=True)
            host_out = torch.zeros(B, dtype=torch.float32).pin_memory()
            e = self._engines[B] = (m, host_in, host_out, torch.cuda.Event())
        return e

    def set_params(self, params: Dict[str, torch.Tensor]) -> None:
        """Replace the weights of every engine built so far (and of those built later) — e.g. a newer checkpoint."""
        self._params = params if self.fused else None
        if self.fused:
            for m, _, _, _ in self._engines.values():
                m.load_state_dict(params, strict=False)
            torch.cuda.synchronize()
        else:
            self.model.load_state_dict(params, strict=False)

    def _predict_fused(self, ids: torch.Tensor, vals: torch.Tensor) -> torch.Tensor:
        n = ids.shape[0]
        B = next(b for b in self._buckets if b >= n)
        m, hin, hout, ev = self._engine(B)
        hin["ids"][:n].copy_(ids)  # int64 -> int32 on the way into the pinned slot
        hin["vals"][:n].copy_(vals)
        if n < B:  # padding rows: id 0 with value 0
            hin["ids"][n:].zero_()
            hin["vals"][n:].zero_()
        m.load_packed(hin["packed"], 0, non_blocking=True)
        p = m.predict_slot(n, 0)
        hout[:n].copy_(p, non_blocking=True)
        ev.record()
        ev.synchronize()
        return hout[:n].clone()

    @classmethod
    def load(cls, model_dir: str, max_batch: int = 4096, backend: str = "auto") -> "Predictor":
        path = model_dir if os.path.exists(os.path.join(model_dir, "config.json")) else latest_export(model_dir)
        meta = json.load(open(os.path.join(path, "config.json")))
        cfg = DeepFMConfig(**meta["model"])
        names = [n for n, _, _ in native_names(cfg)]
        params = tf_to_native(read_bundle(os.path.join(path, "variables", "variables")), names, "GD")
        params.pop("global_step", None)
        return cls(cfg, params, max_batch, backend)

    @torch.no_grad()
    def predict(self, inputs: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        ids = torch.as_tensor(inputs["feat_ids"]).reshape(-1, self.cfg.field_size)
        vals = torch.as_tensor(inputs["feat_vals"], dtype=torch.float32).reshape(-1, self.cfg.field_size)
        out = []
        for s in range(0, ids.shape[0], self.max_batch):
            if self.fused:
                out.append(self._predict_fused(ids[s:s + self.max_batch], vals[s:s + self.max_batch]))
                continue
            p = self.model.predict(ids[s:s + self.max_batch].to(self.dev), vals[s:s + self.max_batch].to(self.dev))
            out.append(p.detach().float().cpu().clone())
        return {"prob": torch.cat(out) if out else torch.zeros(0)}


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description="Score TFRecords with an exported DeepFM servable")
    ap.add_argument("--model_dir", required=True)
    ap.add_argument("--input", required=True, nargs="+")
    ap.add_argument("--output", default="pred.txt")
    ap.add_argument("--batch_size", type=int, default=4096)
    a = ap.parse_args(argv)
    from ..data.pipeline import InputPipeline
    p = Predictor.load(a.model_dir, a.batch_size)
    n = 0
    with open(a.output, "w") as fo:
        for b in InputPipeline(a.input, p.cfg.field_size, a.batch_size, drop_remainder=False, pin_memory=False):
            probs = p.predict({"feat_ids": b.ids.clone(), "feat_vals": b.vals.clone()})["prob"]
            b.release()
            for v in probs.tolist():
                fo.write("%f\n" % v)
            n += probs.numel()
    print("wrote %d predictions to %s" % (n, a.output))
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
