"""Servable export (reference: `export_savedmodel(servable_model_dir, raw serving receiver)`, PS:451-467 /
HVD:415-431). Emitting a TensorFlow GraphDef is out of scope; the artefact keeps everything a serving
system needs, with the reference's signature names:

    <servable_model_dir>/<unix-timestamp>/
        variables/variables.{index,data-00000-of-00001}   inference variables, TF names, TF tensor-bundle format
        signature.json    {"serving_default": inputs feat_ids [None,F] int64, feat_vals [None,F] float32 -> prob}
        config.json       DeepFMConfig + global_step

Load it with `deepfm_b200.serving.Predictor`.
"""
from __future__ import annotations

import dataclasses
import json
import os
import time
from typing import Dict

import torch

from ..models.config import DeepFMConfig
from ..utils.checkpoint import native_to_tf
from ..utils.tf_bundle import write_bundle


def signature(cfg: DeepFMConfig) -> dict:
    return {"serving_default": {
        "method_name": "tensorflow/serving/predict",
        "inputs": {"feat_ids": {"dtype": "int64", "shape": [None, cfg.field_size]},
                   "feat_vals": {"dtype": "float32", "shape": [None, cfg.field_size]}},
        "outputs": {"prob": {"dtype": "float32", "shape": [None]}}}}


def export_servable(export_dir_base: str, cfg: DeepFMConfig, params: Dict[str, torch.Tensor], global_step: int = 0) -> str:
    ts = int(time.time())
    path = os.path.join(export_dir_base, str(ts))
    while os.path.exists(path):
        ts += 1
        path = os.path.join(export_dir_base, str(ts))
    os.makedirs(os.path.join(path, "variables"))
    write_bundle(os.path.join(path, "variables", "variables"), native_to_tf(params, "GD"))
    with open(os.path.join(path, "signature.json"), "w") as f:
        json.dump(signature(cfg), f, indent=1)
    with open(os.path.join(path, "config.json"), "w") as f:
        json.dump({"model": dataclasses.asdict(cfg), "global_step": int(global_step), "framework": "deepfm_b200"}, f,
                  indent=1)
    return path


def latest_export(export_dir_base: str) -> str:
    subs = [d for d in os.listdir(export_dir_base) if d.isdigit()]
    if not subs:
        raise FileNotFoundError("no exported model under %s" % export_dir_base)
    return os.path.join(export_dir_base, max(subs, key=int))
