"""Flat parameter layout shared by the fused engine, the fused optimizers and the comm kernels.

All trainable variables of the reference `model_fn` live in ONE fp32 buffer so that the optimizer
(reference: 11 `ApplyAdam` launches, PS:269) and the gradient all-reduce (reference: Horovod's fusion buffer,
HVD:262) are single sweeps:

    [ fm_bias | fm_w | fm_v | mlp0.weight ... mlpL.weight | mlp0.bias ... | deep_out.weight | deep_out.bias | bn.* ]
      `--- decay range (whole-table L2) ---'  `--- shadow range (bf16 GEMM operands) ---'

Every segment starts on a 64-element (256-byte) boundary; padding stays zero under every optimizer.
Gradients and optimizer slots use the same offsets in their own flat buffers.
"""
from __future__ import annotations

from collections import OrderedDict
from dataclasses import dataclass
from typing import List, Tuple

from .config import DeepFMConfig

ALIGN = 64


def _al(x: int) -> int:
    return (x + ALIGN - 1) // ALIGN * ALIGN


@dataclass
class Segment:
    name: str
    shape: Tuple[int, ...]
    offset: int
    numel: int
    trainable: bool = True


class FlatLayout:
    def __init__(self, cfg: DeepFMConfig, include_embeddings: bool = True):
        """`include_embeddings=False`: fm_w / fm_v live elsewhere (row-sharded tables, parameter server); the flat
        buffer then holds fm_bias and the MLP only and the decay range is empty."""
        self.cfg = cfg
        self.include_embeddings = include_embeddings
        segs: "OrderedDict[str, Segment]" = OrderedDict()
        off = 0

        def add(name, shape, trainable=True):
            nonlocal off
            n = 1
            for s in shape:
                n *= s
            segs[name] = Segment(name, tuple(shape), off, n, trainable)
            off = _al(off + n)

        add("fm_bias", (1,))
        if include_embeddings:
            add("fm_w", (cfg.feature_size,))
            add("fm_v", (cfg.feature_size, cfg.embedding_size))
            self.decay_range = (segs["fm_w"].offset, off)
        else:
            self.decay_range = (0, 0)
        d = cfg.deep_input_dim
        shadow_begin = off
        for i, h in enumerate(cfg.deep_layers):
            add("mlp%d.weight" % i, (h, d))
            d = h
        self.shadow_range = (shadow_begin, off)
        for i, h in enumerate(cfg.deep_layers):
            add("mlp%d.bias" % i, (h,))
        add("deep_out.weight", (d,))
        add("deep_out.bias", (1,))
        if cfg.batch_norm:
            for i, h in enumerate(cfg.deep_layers):
                add("bn_%d.beta" % i, (h,))
                add("bn_%d.gamma" % i, (h,))
        self.trainable_numel = off
        # non-trainable state (BN moving statistics) lives after the optimised range
        if cfg.batch_norm:
            for i, h in enumerate(cfg.deep_layers):
                add("bn_%d.moving_mean" % i, (h,), trainable=False)
                add("bn_%d.moving_variance" % i, (h,), trainable=False)
        self.total_numel = off
        self.segments = segs

    def view(self, flat, name: str):
        s = self.segments[name]
        return flat[s.offset:s.offset + s.numel].view(s.shape)

    def shadow_view(self, shadow, name: str):
        s = self.segments[name]
        o = s.offset - self.shadow_range[0]
        return shadow[o:o + s.numel].view(s.shape)

    def names(self, trainable_only: bool = False) -> List[str]:
        return [k for k, s in self.segments.items() if s.trainable or not trainable_only]
