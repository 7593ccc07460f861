"""TF-1 initialisers, restated (SURVEY 2.2 variable table).

* `fm_w`, `fm_v`: `tf.glorot_normal_initializer()` (PS:167-168) = VarianceScaling(scale=1,
  mode=fan_avg, distribution=truncated_normal): stddev = sqrt(1/fan_avg)/0.87962566103423978,
  samples beyond 2 sigma are redrawn.  1-D `fm_w`: fan_in = fan_out = N.  `fm_v` [N,K]: fan_in=N, fan_out=K.
* MLP weights: `tf.contrib.layers.fully_connected` default = Xavier *uniform*, limit sqrt(6/(in+out)); biases zero.
* `fm_bias`: constant 0 (PS:166).
"""
from __future__ import annotations

import math

import torch

_TRUNC_STD = 0.87962566103423978


def glorot_normal_(t: torch.Tensor, gen: torch.Generator) -> torch.Tensor:
    if t.dim() == 1:
        fan_in = fan_out = t.shape[0]
    else:
        fan_in, fan_out = t.shape[0], t.shape[1]
    std = math.sqrt(1.0 / ((fan_in + fan_out) / 2.0)) / _TRUNC_STD
    with torch.no_grad():
        x = torch.empty(t.shape, dtype=torch.float32)
        x.normal_(0.0, 1.0, generator=gen)
        bad = x.abs() > 2.0
        while bool(bad.any()):
            x[bad] = torch.empty(int(bad.sum()), dtype=torch.float32).normal_(0.0, 1.0, generator=gen)
            bad = x.abs() > 2.0
        t.copy_(x * std)
    return t


def xavier_uniform_(t: torch.Tensor, fan_in: int, fan_out: int, gen: torch.Generator) -> torch.Tensor:
    lim = math.sqrt(6.0 / (fan_in + fan_out))
    with torch.no_grad():
        x = torch.empty(t.shape, dtype=torch.float32)
        x.uniform_(-lim, lim, generator=gen)
        t.copy_(x)
    return t
