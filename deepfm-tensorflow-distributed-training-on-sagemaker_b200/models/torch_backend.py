"""TorchDeepFM — the plain-PyTorch backend (any device, fp32): oracle forward + autograd + TF-1 optimizers.

Used (a) on CPU boxes — tests, the gloo/PS plumbing config — (b) for `--batch_norm` models, and (c) as the
numerical referee of FusedDeepFM. Same public interface as FusedDeepFM.
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Dict, Optional

import torch

from ..optim.tf1 import SLOT_SUFFIXES, TF1Optimizer
from .config import DeepFMConfig
from .oracle import oracle_forward, oracle_grads
from .params import init_params


class TorchDeepFM:
    def __init__(self, cfg: DeepFMConfig, batch_size: int = 0, optimizer: Optional[TF1Optimizer] = None, seed: int = 0,
                 device="cpu", params: Optional[Dict[str, torch.Tensor]] = None, global_batch: Optional[int] = None,
                 row_offset: int = 0):
        self.cfg, self.B = cfg, batch_size
        self.dev = torch.device(device)
        self.opt = optimizer or TF1Optimizer("Adam", 5e-4)
        self.seed = seed
        self.global_batch = global_batch
        self.row_offset = row_offset
        src = params if params is not None else init_params(cfg, seed)
        self.params = OrderedDict((k, v.detach().clone().to(self.dev)) for k, v in src.items())
        self.trainable = [k for k in self.params if not k.endswith(("moving_mean", "moving_variance"))]
        self.slots = self.opt.init_slots({k: self.params[k] for k in self.trainable})
        self._loss = torch.zeros((), device=self.dev)
        self.grad_hook = None  # parallel wrappers install an all-reduce / PS push here
        self.kernels_per_step = 0

    # -- training ------------------------------------------------------------------------------------------------
    def compute_grads(self, ids, vals, labels):
        ids, vals, labels = ids.to(self.dev), vals.to(self.dev), labels.to(self.dev)
        r = oracle_grads(self.params, self.cfg, ids, vals, labels, True, self.seed, self.opt.step, self.row_offset)
        if self.global_batch and self.global_batch != ids.shape[0]:
            # mean over the GLOBAL batch: data-gradient scaled by local/global, L2 part unchanged
            sc = ids.shape[0] / float(self.global_batch)
            l2 = self.cfg.l2_reg
            for k, g in r["grads"].items():
                if k in ("fm_w", "fm_v") and self.cfg.l2_mode == "exact":
                    r["grads"][k] = (g - l2 * self.params[k]) * sc  # data part only; L2 re-added after the reduce
                else:
                    r["grads"][k] = g * sc
            r["l2_deferred"] = True
        return r

    def train_step(self, ids, vals, labels, slot_idx: int = 0) -> torch.Tensor:
        r = self.compute_grads(ids, vals, labels)
        grads = r["grads"]
        if self.grad_hook is not None:
            grads = self.grad_hook(grads)
        if r.get("l2_deferred"):
            for k in ("fm_w", "fm_v"):
                grads[k] = grads[k] + self.cfg.l2_reg * self.params[k]
        self.opt.apply_all(self.params, grads, self.slots)
        for k, v in r["bn_updates"].items():
            self.params[k].copy_(v)
        self._loss = r["loss"]
        return self._loss

    def last_loss(self) -> torch.Tensor:
        return self._loss

    # -- inference -----------------------------------------------------------------------------------------------
    @torch.no_grad()
    def predict(self, ids, vals) -> torch.Tensor:
        return torch.sigmoid(oracle_forward(self.params, self.cfg, ids.to(self.dev), vals.to(self.dev), False))

    @torch.no_grad()
    def eval_step(self, ids, vals, labels, update_auc: bool = True):
        ids, vals, labels = ids.to(self.dev), vals.to(self.dev), labels.to(self.dev).float().reshape(-1)
        y = oracle_forward(self.params, self.cfg, ids, vals, False)
        loss_sum = torch.nn.functional.binary_cross_entropy_with_logits(y, labels, reduction="sum").reshape(1)
        return loss_sum, torch.sigmoid(y)

    def l2_penalty(self) -> torch.Tensor:
        return 0.5 * self.cfg.l2_reg * (self.params["fm_w"].pow(2).sum() + self.params["fm_v"].pow(2).sum())

    # -- state ---------------------------------------------------------------------------------------------------
    @property
    def global_step(self) -> int:
        return self.opt.step

    def named_parameters(self):
        return self.params

    def named_slots(self):
        return self.slots

    def state_dict(self):
        sd = OrderedDict((k, v.detach().clone()) for k, v in self.params.items())
        for k, sl in self.slots.items():
            for suffix, t in zip(SLOT_SUFFIXES[self.opt.name], sl):
                sd["%s/%s" % (k, suffix)] = t.detach().clone()
        sd["global_step"] = torch.tensor([self.opt.step], dtype=torch.int64)
        return sd

    def load_state_dict(self, sd, strict: bool = True):
        for k in self.params:
            if k in sd:
                self.params[k].copy_(sd[k].to(self.dev))
            elif strict:
                raise KeyError("missing parameter %s" % k)
        for k, sl in self.slots.items():
            for suffix, t in zip(SLOT_SUFFIXES[self.opt.name], sl):
                key = "%s/%s" % (k, suffix)
                if key in sd:
                    t.copy_(sd[key].to(self.dev))
        if "global_step" in sd:
            self.opt.step = int(sd["global_step"].reshape(-1)[0])
