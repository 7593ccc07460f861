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
    def __init__(self, cfg: DeepFMConfig, params: Dict[str, torch.Tensor], max_batch: int = 4096, backend: str = "auto"):
        self.cfg, self.max_batch = cfg, max_batch
        use_fused = backend == "fused" or (backend == "auto" and torch.cuda.is_available() and not cfg.batch_norm
                                           and not cfg.fused_constraints())
        if use_fused:
            from ..models.fused import FusedDeepFM
            self.model = FusedDeepFM(cfg, max_batch, params=params, use_graph=False, input_slots=1)
            self.dev = self.model.dev
        else:
            self.dev = torch.device("cuda" if (torch.cuda.is_available() and backend != "torch-cpu") else "cpu")
            self.model = TorchDeepFM(cfg, max_batch, device=self.dev, params=params)

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
