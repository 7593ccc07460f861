"""Parameter container for DeepFM with the reference's variable names.

Native layout vs the reference's TF variables:
    fm_bias [1], fm_w [N], fm_v [N,K]                      identical
    Deep-part/mlp{i}/weights  TF [in,out]   <->  native `mlp{i}.weight` [out,in]  (K-major for the
                                                 forward tcgen05 GEMM; transposed on checkpoint I/O)
    Deep-part/mlp{i}/biases   [out]
    Deep-part/deep_out/{weights [in,1], biases [1]}   <->  `deep_out.weight` [in], `deep_out.bias` [1]
    Deep-part/bn_{i}/{beta,gamma,moving_mean,moving_variance}  (only with --batch_norm)
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Tuple

import torch

from .config import DeepFMConfig
from .init import glorot_normal_, xavier_uniform_


def native_names(cfg: DeepFMConfig):
    """Ordered (name, shape, trainable) for every model variable."""
    out = [("fm_bias", (1,), True), ("fm_w", (cfg.feature_size,), True),
           ("fm_v", (cfg.feature_size, cfg.embedding_size), True)]
    d = cfg.deep_input_dim
    for i, h in enumerate(cfg.deep_layers):
        out.append(("mlp%d.weight" % i, (h, d), True))
        out.append(("mlp%d.bias" % i, (h,), True))
        if cfg.batch_norm:
            out.append(("bn_%d.beta" % i, (h,), True))
            out.append(("bn_%d.gamma" % i, (h,), True))
            out.append(("bn_%d.moving_mean" % i, (h,), False))
            out.append(("bn_%d.moving_variance" % i, (h,), False))
        d = h
    out.append(("deep_out.weight", (d,), True))
    out.append(("deep_out.bias", (1,), True))
    return out


def tf_name(native: str) -> Tuple[str, bool]:
    """native name -> (TF checkpoint variable name, needs_transpose_or_reshape)."""
    if native in ("fm_bias", "fm_w", "fm_v"):
        return native, False
    scope, leaf = native.split(".")
    if scope.startswith("bn_"):
        return "Deep-part/%s/%s" % (scope, leaf), False
    leaf_tf = {"weight": "weights", "bias": "biases"}[leaf]
    return "Deep-part/%s/%s" % (scope, leaf_tf), leaf == "weight"


def init_params(cfg: DeepFMConfig, seed: int = 0, device="cpu") -> "OrderedDict[str, torch.Tensor]":
    """Random-init every variable with the reference's initialisers (fp32 master copies)."""
    gen = torch.Generator().manual_seed(seed)
    p: "OrderedDict[str, torch.Tensor]" = OrderedDict()
    for name, shape, _ in native_names(cfg):
        t = torch.zeros(shape, dtype=torch.float32)
        if name in ("fm_w", "fm_v"):
            glorot_normal_(t, gen)
        elif name.endswith(".weight"):
            if name.startswith("deep_out"):
                xavier_uniform_(t, shape[0], 1, gen)
            else:
                xavier_uniform_(t, shape[1], shape[0], gen)
        elif name.endswith(".gamma") or name.endswith(".moving_variance"):
            t.fill_(1.0)
        p[name] = t.to(device)
    return p


def to_tf_layout(name: str, t: torch.Tensor) -> torch.Tensor:
    """native tensor -> tensor with the reference's TF shape."""
    if name.startswith("deep_out") and name.endswith("weight"):
        return t.reshape(-1, 1).contiguous()
    if name.endswith(".weight"):
        return t.t().contiguous()
    return t


def from_tf_layout(name: str, t: torch.Tensor) -> torch.Tensor:
    if name.startswith("deep_out") and name.endswith("weight"):
        return t.reshape(-1).contiguous()
    if name.endswith(".weight"):
        return t.t().contiguous()
    return t
