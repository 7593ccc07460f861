"""fp32 CPU/GPU oracle: a plain-PyTorch restatement of the reference `model_fn`
(PS:149-278 == HVD:141-273).  Every CUDA kernel and fused path is tested against this.

    y_w  = sum_f fm_w[ids] * vals                                   (PS:177-179)
    e    = fm_v[ids] * vals[..., None]                               (PS:181-184)
    y_v  = 0.5 * sum_k ((sum_f e)^2 - sum_f e^2)                     (PS:185-187)
    x    = reshape(e, [B, F*K]); per layer relu(xW+b) -> [BN] -> dropout(keep) in TRAIN (PS:203-218)
    y_d  = x W_out + b_out                                           (PS:221-223)
    y    = fm_bias + y_w + y_v + y_d ; prob = sigmoid(y)             (PS:228-232)
    loss = mean(sigmoid_xent(y, label)) + l2*0.5*|fm_w|^2 + l2*0.5*|fm_v|^2   (PS:244-246)

The MLP `weights_regularizer` is registered but never added to the loss in the reference
(PS:214,222 vs PS:244) — reproduced here by not regularising the MLP.
"""
from __future__ import annotations

from typing import Dict, Optional

import torch
import torch.nn.functional as F

from ..utils.philox import dropout_keep_mask
from .config import DeepFMConfig


def oracle_forward(params: Dict[str, torch.Tensor], cfg: DeepFMConfig, ids: torch.Tensor, vals: torch.Tensor,
                   train: bool = False, seed: int = 0, step: int = 0, row_offset: int = 0,
                   bn_updates: Optional[dict] = None, compute_dtype: torch.dtype = torch.float32):
    """Returns logits y [B].  `compute_dtype=bfloat16` emulates the kernels' rounding points
    (bf16 deep-input / activations / weights, fp32 accumulation) for tight kernel tests."""
    ids = ids.reshape(-1, cfg.field_size).long()
    vals = vals.reshape(-1, cfg.field_size).float()
    B = ids.shape[0]
    y_w = (params["fm_w"][ids] * vals).sum(1)
    e = params["fm_v"][ids] * vals.unsqueeze(-1)
    y_v = 0.5 * (e.sum(1).pow(2) - e.pow(2).sum(1)).sum(1)
    x = e.reshape(B, cfg.deep_input_dim)
    lowp = compute_dtype != torch.float32
    if lowp:
        x = x.to(compute_dtype).float()
    for i, h in enumerate(cfg.deep_layers):
        w = params["mlp%d.weight" % i]
        if lowp:
            w = w.to(compute_dtype).float()
        x = F.relu(x @ w.t() + params["mlp%d.bias" % i])
        if cfg.batch_norm:
            x = _batch_norm(params, cfg, i, x, train, bn_updates)
        if train and cfg.dropout[i] < 1.0:
            keep = torch.from_numpy(dropout_keep_mask(B, h, cfg.dropout[i], seed, step, i, row_offset)).to(x.device)
            x = torch.where(keep, x / cfg.dropout[i], torch.zeros_like(x))
        if lowp:
            x = x.to(compute_dtype).float()
    y_d = x @ params["deep_out.weight"] + params["deep_out.bias"]
    return params["fm_bias"] + y_w + y_v + y_d


def _batch_norm(params, cfg, i, x, train, bn_updates, eps: float = 1e-3):
    """`tf.contrib.layers.batch_norm(decay, center, scale, updates_collections=None)` after ReLU
    (PS:215-216, 288-292). contrib default epsilon = 0.001; moving stats updated in place
    (collected in `bn_updates` so callers can apply them outside autograd)."""
    g, b = params["bn_%d.gamma" % i], params["bn_%d.beta" % i]
    if train:
        mean = x.mean(0)
        var = x.var(0, unbiased=False)
        if bn_updates is not None:
            d = cfg.batch_norm_decay
            bn_updates["bn_%d.moving_mean" % i] = (d * params["bn_%d.moving_mean" % i] + (1 - d) * mean).detach()
            bn_updates["bn_%d.moving_variance" % i] = (d * params["bn_%d.moving_variance" % i] + (1 - d) * var).detach()
    else:
        mean, var = params["bn_%d.moving_mean" % i], params["bn_%d.moving_variance" % i]
    return (x - mean) * torch.rsqrt(var + eps) * g + b


def oracle_loss(params, cfg: DeepFMConfig, logits: torch.Tensor, labels: torch.Tensor,
                touched_ids: Optional[torch.Tensor] = None):
    """Mean sigmoid cross-entropy + the two FM L2 terms (PS:244-246).
    `l2_mode='touched'`: the L2 terms cover only rows present in `touched_ids` (unique)."""
    data = F.binary_cross_entropy_with_logits(logits, labels.float().reshape(-1), reduction="mean")
    if cfg.l2_mode == "exact" or touched_ids is None:
        reg = 0.5 * cfg.l2_reg * (params["fm_w"].pow(2).sum() + params["fm_v"].pow(2).sum())
    else:
        u = torch.unique(touched_ids.reshape(-1).long())
        reg = 0.5 * cfg.l2_reg * (params["fm_w"][u].pow(2).sum() + params["fm_v"][u].pow(2).sum())
    return data + reg, data


def oracle_grads(params, cfg: DeepFMConfig, ids, vals, labels, train=True, seed=0, step=0, row_offset=0,
                 compute_dtype=torch.float32):
    """Autograd gradients of the full reference loss w.r.t. every trainable variable."""
    leaves = {k: v.detach().clone().requires_grad_(not k.startswith("bn_") or k.endswith(("beta", "gamma")))
              for k, v in params.items()}
    bn_updates: dict = {}
    y = oracle_forward(leaves, cfg, ids, vals, train, seed, step, row_offset, bn_updates, compute_dtype)
    loss, data_loss = oracle_loss(leaves, cfg, y, labels, ids)
    loss.backward()
    grads = {k: (v.grad if v.grad is not None else torch.zeros_like(v)) for k, v in leaves.items() if v.requires_grad}
    return {"loss": loss.detach(), "data_loss": data_loss.detach(), "logits": y.detach(), "grads": grads,
            "bn_updates": bn_updates}
