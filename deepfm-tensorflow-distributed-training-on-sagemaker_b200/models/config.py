"""Model hyper-parameters (reference: `model_params` dict HVD:343-352, parsed at HVD:143-155)."""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Sequence, Union


def _int_list(v: Union[str, Sequence[int]]) -> List[int]:
    if isinstance(v, str):
        return [int(x) for x in v.split(",") if x.strip()]
    return [int(x) for x in v]


def _float_list(v: Union[str, Sequence[float]]) -> List[float]:
    if isinstance(v, str):
        return [float(x) for x in v.split(",") if x.strip()]
    return [float(x) for x in v]


@dataclass
class DeepFMConfig:
    """`dropout` holds KEEP probabilities, as in the reference (`tf.nn.dropout(keep_prob=)`, PS:218)."""
    field_size: int = 39
    feature_size: int = 117581
    embedding_size: int = 32
    deep_layers: List[int] = field(default_factory=lambda: [256, 128, 64])
    dropout: List[float] = field(default_factory=lambda: [0.5, 0.5, 0.5])
    l2_reg: float = 1e-4
    batch_norm: bool = False
    batch_norm_decay: float = 0.9
    # 'exact' reproduces the reference's whole-table l2_loss (dense decay of every row each
    # step); 'touched' regularises only rows present in the batch (needed once the
    # table no longer fits a per-step sweep: sharded / parameter-server configs).
    l2_mode: str = "exact"

    def __post_init__(self):
        self.deep_layers = _int_list(self.deep_layers)
        self.dropout = _float_list(self.dropout)
        if len(self.dropout) < len(self.deep_layers):
            raise ValueError("need one dropout keep-prob per deep layer")
        if self.l2_mode not in ("exact", "touched"):
            raise ValueError("l2_mode must be 'exact' or 'touched'")
        if self.field_size <= 0 or self.feature_size <= 0 or self.embedding_size <= 0:
            raise ValueError("field_size, feature_size and embedding_size must be positive (got %d, %d, %d)"
                             % (self.field_size, self.feature_size, self.embedding_size))
        if not self.deep_layers or min(self.deep_layers) <= 0:
            raise ValueError("deep_layers needs at least one positive width, got %r" % (self.deep_layers,))
        if any(not (0.0 < p <= 1.0) for p in self.dropout):
            raise ValueError("dropout holds KEEP probabilities in (0, 1], got %r" % (self.dropout,))
        if not (0.0 <= self.batch_norm_decay < 1.0) or self.l2_reg < 0:
            raise ValueError("batch_norm_decay must be in [0, 1) and l2_reg >= 0")

    def fused_constraints(self) -> List[str]:
        """Why this configuration cannot run on the fused sm_100a kernels ([] = it can). The kernels move bf16 rows with
        TMA (16-byte row strides) and embedding rows with 128-bit accesses; the torch backend has no such limits."""
        problems = []
        if self.embedding_size not in (4, 8, 16, 32, 64, 128):
            problems.append("embedding_size must be one of 4, 8, 16, 32, 64, 128 (got %d)" % self.embedding_size)
        if self.deep_input_dim % 8:
            problems.append("field_size * embedding_size must be a multiple of 8 (got %d)" % self.deep_input_dim)
        bad = [h for h in self.deep_layers if h % 8]
        if bad:
            problems.append("deep layer widths must be multiples of 8 (got %r)" % (bad,))
        return problems

    @property
    def deep_input_dim(self) -> int:
        return self.field_size * self.embedding_size

    @property
    def layer_dims(self):
        """[(in, out)] for mlp0..mlpL-1 then deep_out."""
        dims, d = [], self.deep_input_dim
        for h in self.deep_layers:
            dims.append((d, h))
            d = h
        dims.append((d, 1))
        return dims

    def num_parameters(self) -> int:
        n = 1 + self.feature_size * (1 + self.embedding_size)
        for i, o in self.layer_dims:
            n += i * o + o
        if self.batch_norm:
            n += sum(2 * h for h in self.deep_layers)
        return n
