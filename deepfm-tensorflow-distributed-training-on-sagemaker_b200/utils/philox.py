"""Counter-based Philox4x32-10 shared by the CUDA kernels and the CPU oracle.

The reference applies `tf.nn.dropout(keep_prob=...)` (PS:218 / HVD:210) with TF's
own Philox stream layout, which cannot be reproduced outside TF.  Both sides of
this framework therefore use ONE counter-based generator so a kernel and the
oracle draw identical masks:

    counter = (elem // 4, layer, step_lo, step_hi)     key = (seed_lo, seed_hi)
    u32     = philox4x32_10(counter, key)[elem % 4]
    keep    = u32 < floor(keep_prob * 2**32)            (keep_prob >= 1 keeps all)

`elem = row * width + col` of the layer's output matrix.  The CUDA twin of this
file is `csrc/philox.cuh`.
"""
import numpy as np

_M0 = np.uint64(0xD2511F53)
_M1 = np.uint64(0xCD9E8D57)
_W0 = 0x9E3779B9
_W1 = 0xBB67AE85
_MASK = np.uint64(0xFFFFFFFF)
_S32 = np.uint64(32)


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Vectorised Philox4x32-10. All inputs uint32 arrays/scalars; returns 4 uint32 arrays."""
    c0 = np.asarray(c0, dtype=np.uint64)
    c1 = np.broadcast_to(np.asarray(c1, dtype=np.uint64), c0.shape)
    c2 = np.broadcast_to(np.asarray(c2, dtype=np.uint64), c0.shape)
    c3 = np.broadcast_to(np.asarray(c3, dtype=np.uint64), c0.shape)
    k0 = int(k0) & 0xFFFFFFFF
    k1 = int(k1) & 0xFFFFFFFF
    for _ in range(10):
        p0 = _M0 * c0
        p1 = _M1 * c2
        hi0, lo0 = p0 >> _S32, p0 & _MASK
        hi1, lo1 = p1 >> _S32, p1 & _MASK
        c0, c1, c2, c3 = (hi1 ^ c1 ^ np.uint64(k0)), lo1, (hi0 ^ c3 ^ np.uint64(k1)), lo0
        k0 = (k0 + _W0) & 0xFFFFFFFF
        k1 = (k1 + _W1) & 0xFFFFFFFF
    return (c0.astype(np.uint32), c1.astype(np.uint32), c2.astype(np.uint32), c3.astype(np.uint32))


def keep_threshold(keep_prob: float) -> int:
    """uint32 threshold; 0xFFFFFFFF + 'keep all' flag handled by callers for keep_prob>=1."""
    if keep_prob >= 1.0:
        return 0xFFFFFFFF
    return int(np.floor(float(keep_prob) * 4294967296.0)) & 0xFFFFFFFF


def dropout_keep_mask(rows: int, width: int, keep_prob: float, seed: int, step: int, layer: int,
                      row_offset: int = 0) -> np.ndarray:
    """Boolean [rows, width] keep-mask identical to the kernels' epilogue mask.

    `row_offset` lets a data-parallel rank reproduce its slice of a global batch.
    """
    if keep_prob >= 1.0:
        return np.ones((rows, width), dtype=bool)
    elem = (np.arange(rows, dtype=np.uint64)[:, None] + np.uint64(row_offset)) * np.uint64(width) \
        + np.arange(width, dtype=np.uint64)[None, :]
    ctr = (elem >> np.uint64(2)) & _MASK
    lane = (elem & np.uint64(3)).astype(np.int64)
    r = philox4x32_10(ctr, layer, step & 0xFFFFFFFF, (step >> 32) & 0xFFFFFFFF,
                      seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    stacked = np.stack(r, axis=-1)
    u = np.take_along_axis(stacked, lane[..., None], axis=-1)[..., 0]
    return u < np.uint32(keep_threshold(keep_prob))
