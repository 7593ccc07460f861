"""Synthetic Criteo-shaped CTR data (there is no network for the real dataset).

Shape follows the reference's run configuration (NB-HVD:94-103): 39 fields = 13 numeric (fixed id per field,
value in [0,1]) + 26 categorical (one id per field out of that field's slice of the vocabulary, value 1.0),
`feature_size` total ids. Categorical ids follow a power-law so duplicate-id handling is exercised.
Labels come from a hidden linear+pairwise teacher so training has something to learn (AUC rises above 0.5).
"""
from __future__ import annotations

import os
from typing import List, Tuple

import numpy as np
import torch

from ..utils.tfrecord import TFRecordWriter, make_ctr_example


class SyntheticCTR:
    def __init__(self, field_size: int = 39, feature_size: int = 117581, num_numeric: int = 13, seed: int = 0,
                 zipf_a: float = 1.2, positive_rate: float = 0.26, teacher_seed: int = 12345):
        """`seed` drives the SAMPLING only; the hidden teacher depends on `teacher_seed`, so train / validation / test
        sets drawn with different seeds share one ground truth."""
        self.F, self.N = field_size, feature_size
        self.num_numeric = min(num_numeric, field_size)
        self.seed = seed
        self.zipf_a = zipf_a
        n_cat = field_size - self.num_numeric
        # vocabulary slices: ids [0, num_numeric) are the numeric fields; the rest is split across categorical fields
        rest = feature_size - self.num_numeric
        base = rest // max(n_cat, 1)
        self.cat_lo = np.array([self.num_numeric + i * base for i in range(n_cat)], dtype=np.int64)
        self.cat_sz = np.array([base] * n_cat, dtype=np.int64)
        if n_cat:
            self.cat_sz[-1] = rest - base * (n_cat - 1)
        rng = np.random.default_rng(teacher_seed)
        self.teacher_w = rng.normal(0, 1.0, feature_size).astype(np.float32)
        self.bias = float(np.log(positive_rate / (1 - positive_rate)))

    def numpy_batch(self, B: int, index: int) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
        rng = np.random.default_rng((self.seed, index))
        ids = np.empty((B, self.F), dtype=np.int64)
        vals = np.ones((B, self.F), dtype=np.float32)
        nn = self.num_numeric
        ids[:, :nn] = np.arange(nn)[None, :]
        vals[:, :nn] = rng.random((B, nn), dtype=np.float32)
        n_cat = self.F - nn
        if n_cat:
            u = rng.random((B, n_cat))
            # inverse-CDF power law over each field's slice: rank = size * u^a  (a > 1 -> skewed toward rank 0)
            rank = np.floor(self.cat_sz[None, :] * np.power(u, self.zipf_a * 2.0)).astype(np.int64)
            rank = np.minimum(rank, self.cat_sz[None, :] - 1)
            ids[:, nn:] = self.cat_lo[None, :] + rank
        logit = self.bias + 0.35 * (self.teacher_w[ids] * vals).sum(1)
        p = 1.0 / (1.0 + np.exp(-logit))
        labels = (rng.random(B) < p).astype(np.float32)
        return ids, vals, labels

    def batch(self, B: int, index: int, ids_dtype=torch.int32):
        ids, vals, labels = self.numpy_batch(B, index)
        return torch.from_numpy(ids).to(ids_dtype), torch.from_numpy(vals), torch.from_numpy(labels)


def write_synthetic_tfrecords(out_dir: str, prefix: str, num_files: int, rows_per_file: int, field_size: int = 39,
                              feature_size: int = 117581, seed: int = 0, native: bool = True) -> List[str]:
    """Writes `<prefix>NNN.tfrecords` files in the converter's Example schema (TOOL:25-32). Returns the paths.
    The trainers glob `tr*.tfrecords`, `va*.tfrecords`, `te*.tfrecords` (HVD:320-324)."""
    os.makedirs(out_dir, exist_ok=True)
    gen = SyntheticCTR(field_size, feature_size, seed=seed)
    paths = []
    encode = None
    if native:
        try:
            from .native import available, encode_ctr_batch
            if available():
                encode = encode_ctr_batch
        except Exception:
            pass
    import zlib
    for fi in range(num_files):
        path = os.path.join(out_dir, "%s%03d.tfrecords" % (prefix, fi))
        ids, vals, labels = gen.numpy_batch(rows_per_file, zlib.crc32(("%s/%d" % (prefix, fi)).encode()))
        if encode is not None:  # columns -> framed records in one native call per 64 Ki rows
            with open(path, "wb") as f:
                for r0 in range(0, rows_per_file, 65536):
                    f.write(encode(labels[r0:r0 + 65536], ids[r0:r0 + 65536], vals[r0:r0 + 65536]))
        else:
            with TFRecordWriter(path) as w:
                for r in range(rows_per_file):
                    w.write(make_ctr_example(float(labels[r]), ids[r].tolist(), vals[r].tolist()))
        paths.append(path)
    return paths
