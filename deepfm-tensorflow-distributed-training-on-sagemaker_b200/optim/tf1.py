"""Optimizers with TensorFlow-1 update rules (reference: PS:260-267 / HVD:252-259).

    Adam      tf.train.AdamOptimizer(lr, 0.9, 0.999, 1e-8):  lr_t = lr*sqrt(1-b2^t)/(1-b1^t);
              m = b1 m + (1-b1) g;  v = b2 v + (1-b2) g^2;  w -= lr_t * m / (sqrt(v) + eps)
              (epsilon OUTSIDE the bias-corrected sqrt: differs from torch.optim.Adam)
    Adagrad   tf.train.AdagradOptimizer(lr, initial_accumulator_value=1e-8): a += g^2; w -= lr*g/sqrt(a)
    Momentum  tf.train.MomentumOptimizer(lr, 0.95): acc = 0.95 acc + g; w -= lr*acc
    ftrl      tf.train.FtrlOptimizer(lr): lr_power=-0.5, initial accumulator 0.1, l1=l2=0
    GD        advertised by the reference's help text (PS:51) but missing there (NameError at PS:269);
              implemented here as plain SGD.

This is the reference implementation (pure torch, any device).  The fused sm_100a
kernels in `csrc/optim.cu` implement the same rules and are tested against it.
Slot names follow TF (`<var>/Adam`, `<var>/Adam_1`, `<var>/Adagrad`, `<var>/Momentum`,
`<var>/Ftrl`, `<var>/Ftrl_1`) so checkpoints interchange with the reference's.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional

import torch

OPT_IDS = {"Adam": 0, "Adagrad": 1, "Momentum": 2, "ftrl": 3, "GD": 4}
_ALIASES = {"adam": "Adam", "adagrad": "Adagrad", "momentum": "Momentum", "ftrl": "ftrl", "gd": "GD", "sgd": "GD"}


def canonical_name(name: str) -> str:
    if name in OPT_IDS:
        return name
    k = name.lower()
    if k in _ALIASES:
        return _ALIASES[k]
    raise ValueError("unknown optimizer %r (expected one of %s)" % (name, sorted(OPT_IDS)))


SLOT_SUFFIXES = {"Adam": ["Adam", "Adam_1"], "Adagrad": ["Adagrad"], "Momentum": ["Momentum"],
                 "ftrl": ["Ftrl", "Ftrl_1"], "GD": []}
SLOT_INIT = {"Adam": [0.0, 0.0], "Adagrad": [1e-8], "Momentum": [0.0], "ftrl": [0.1, 0.0], "GD": []}


class TF1Optimizer:
    def __init__(self, name: str = "Adam", learning_rate: float = 5e-4, beta1: float = 0.9, beta2: float = 0.999,
                 epsilon: float = 1e-8, momentum: float = 0.95):
        self.name = canonical_name(name)
        self.lr = float(learning_rate)
        self.beta1, self.beta2, self.eps, self.momentum = beta1, beta2, epsilon, momentum
        self.step = 0  # number of applied updates (== TF global_step for a single optimizer)

    # -- state ---------------------------------------------------------------------------------
    @property
    def num_slots(self) -> int:
        return len(SLOT_SUFFIXES[self.name])

    def init_slots(self, params: Dict[str, torch.Tensor]) -> Dict[str, List[torch.Tensor]]:
        return {k: [torch.full_like(v, init) for init in SLOT_INIT[self.name]] for k, v in params.items()}

    def adam_lr_t(self, step: Optional[int] = None) -> float:
        t = self.step + 1 if step is None else step
        return self.lr * math.sqrt(1.0 - self.beta2 ** t) / (1.0 - self.beta1 ** t)

    # -- updates -------------------------------------------------------------------------------
    @torch.no_grad()
    def apply_dense(self, p: torch.Tensor, g: torch.Tensor, slots: List[torch.Tensor], t: Optional[int] = None) -> None:
        """In-place update of one variable. `t` = 1-based step index (defaults to self.step+1)."""
        n = self.name
        if n == "Adam":
            m, v = slots
            m.mul_(self.beta1).add_(g, alpha=1 - self.beta1)
            v.mul_(self.beta2).addcmul_(g, g, value=1 - self.beta2)
            p.addcdiv_(m, v.sqrt().add_(self.eps), value=-self.adam_lr_t(t))
        elif n == "Adagrad":
            (a,) = slots
            a.addcmul_(g, g)
            p.addcdiv_(g, a.sqrt(), value=-self.lr)
        elif n == "Momentum":
            (acc,) = slots
            acc.mul_(self.momentum).add_(g)
            p.add_(acc, alpha=-self.lr)
        elif n == "ftrl":
            a, lin = slots
            a_new = a + g * g
            sigma = (a_new.sqrt() - a.sqrt()) / self.lr
            lin.add_(g - sigma * p)
            p.copy_(-lin / (a_new.sqrt() / self.lr))
            a.copy_(a_new)
        elif n == "GD":
            p.add_(g, alpha=-self.lr)

    @torch.no_grad()
    def apply_sparse(self, p: torch.Tensor, rows: torch.Tensor, g_rows: torch.Tensor, slots: List[torch.Tensor],
                     t: Optional[int] = None) -> None:
        """TF sparse-apply semantics: duplicate ids are segment-summed first, then only the touched
        rows (and their slot rows) are updated.  `rows` [n] int64, `g_rows` [n, ...]."""
        u, inv = torch.unique(rows, return_inverse=True)
        g = torch.zeros((u.numel(),) + tuple(g_rows.shape[1:]), dtype=g_rows.dtype, device=g_rows.device)
        g.index_add_(0, inv, g_rows)
        pr = p[u]
        sl = [s[u] for s in slots]
        self.apply_dense(pr, g, sl, t)
        p[u] = pr
        for s, r in zip(slots, sl):
            s[u] = r

    def apply_all(self, params: Dict[str, torch.Tensor], grads: Dict[str, torch.Tensor],
                  slots: Dict[str, List[torch.Tensor]]) -> None:
        t = self.step + 1
        for k, g in grads.items():
            self.apply_dense(params[k], g, slots[k], t)
        self.step = t

    # -- checkpoint helpers --------------------------------------------------------------------
    def slot_names(self, var_tf_name: str) -> List[str]:
        return ["%s/%s" % (var_tf_name, s) for s in SLOT_SUFFIXES[self.name]]

    def scalar_state(self) -> Dict[str, float]:
        """TF keeps beta1_power/beta2_power as non-slot variables for Adam (value = beta^(t+1) after t steps)."""
        if self.name != "Adam":
            return {}
        return {"beta1_power": self.beta1 ** (self.step + 1), "beta2_power": self.beta2 ** (self.step + 1)}
