"""Python binding of the C++ TensorFlow tensor-bundle codec (csrc/io/tensor_bundle.cc)."""
from __future__ import annotations

import ctypes
from ctypes import c_char_p, c_int, c_int64, c_uint64, c_void_p
from typing import Dict, Optional

import numpy as np
import torch

from ..data import native as _n

DT_FLOAT, DT_INT32, DT_INT64 = 1, 3, 9
_NP = {DT_FLOAT: np.float32, DT_INT32: np.int32, DT_INT64: np.int64}
_DT = {np.dtype(np.float32): DT_FLOAT, np.dtype(np.int32): DT_INT32, np.dtype(np.int64): DT_INT64}


def _lib():
    lib = _n.load()
    if not getattr(lib, "_bundle_ready", False):
        lib.dfmio_bundle_last_error.restype = c_char_p
        lib.dfmio_bundle_writer_create.restype = c_void_p
        lib.dfmio_bundle_writer_create.argtypes = [c_char_p]
        lib.dfmio_bundle_writer_create_sharded.restype = c_void_p
        lib.dfmio_bundle_writer_create_sharded.argtypes = [c_char_p, c_int]
        lib.dfmio_bundle_writer_add.argtypes = [c_void_p, c_char_p, c_int, c_int, c_void_p, c_void_p, c_uint64]
        lib.dfmio_bundle_writer_add_partitioned.argtypes = [c_void_p, c_char_p, c_int, c_int, c_void_p, c_void_p, c_uint64,
                                                            c_int]
        lib.dfmio_bundle_writer_finish.argtypes = [c_void_p]
        lib.dfmio_bundle_reader_open.restype = c_void_p
        lib.dfmio_bundle_reader_open.argtypes = [c_char_p]
        lib.dfmio_bundle_reader_count.argtypes = [c_void_p]
        lib.dfmio_bundle_reader_entry.argtypes = [c_void_p, c_int, c_char_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p]
        lib.dfmio_bundle_reader_read.argtypes = [c_void_p, c_int, c_void_p, c_uint64, c_int]
        lib.dfmio_bundle_reader_close.argtypes = [c_void_p]
        lib._bundle_ready = True
    return lib


def _err(lib):
    m = lib.dfmio_bundle_last_error()
    return m.decode() if m else "?"


def write_bundle(prefix: str, tensors: Dict[str, torch.Tensor], num_shards: int = 1,
                 partitioned: Optional[Dict[str, int]] = None) -> None:
    """Write `<prefix>.index` + `<prefix>.data-SSSSS-of-NNNNN` holding `tensors` (keys = TF variable names).
    `num_shards` > 1 spreads the variables round-robin over that many data files - the layout the reference's
    parameter-server jobs leave behind (sharded Saver, one file per PS task); `read_bundle` reads any shard count.
    `partitioned` = {variable name: number of parts}: those variables are stored the way a TF-1 job with
    `tf.fixed_size_partitioner(num_shards)` (the deck's advice for big tables, PDF p32) stores them - a full-shape entry
    listing the slices plus one entry per slice along axis 0; `read_bundle` reassembles them."""
    lib = _lib()
    h = lib.dfmio_bundle_writer_create_sharded(prefix.encode(), int(num_shards))
    if not h:
        raise IOError(_err(lib))
    for name in sorted(tensors):
        a = np.array(tensors[name].detach().cpu().numpy(), order="C", copy=True)  # keeps 0-d scalars 0-d
        if a.dtype not in _DT:
            a = a.astype(np.float32)
        dims = (c_int64 * max(a.ndim, 1))(*a.shape)
        parts = (partitioned or {}).get(name, 0)
        if parts > 1:
            rc = lib.dfmio_bundle_writer_add_partitioned(h, name.encode(), _DT[a.dtype], a.ndim, dims, a.ctypes.data,
                                                         a.nbytes, int(parts))
        else:
            rc = lib.dfmio_bundle_writer_add(h, name.encode(), _DT[a.dtype], a.ndim, dims, a.ctypes.data, a.nbytes)
        if rc != 0:
            raise IOError(_err(lib))
    if lib.dfmio_bundle_writer_finish(h) != 0:
        raise IOError(_err(lib))


def read_bundle(prefix: str, verify_crc: bool = True) -> Dict[str, torch.Tensor]:
    lib = _lib()
    h = lib.dfmio_bundle_reader_open(prefix.encode())
    if not h:
        raise IOError(_err(lib))
    out = {}
    try:
        name = ctypes.create_string_buffer(1024)
        for i in range(lib.dfmio_bundle_reader_count(h)):
            dtype, ndim, nbytes = c_int(), c_int(), c_uint64()
            dims = (c_int64 * 8)()
            lib.dfmio_bundle_reader_entry(h, i, name, 1024, ctypes.byref(dtype), ctypes.byref(ndim), dims,
                                          ctypes.byref(nbytes))
            if dtype.value not in _NP:
                continue  # not a numeric tensor we know (e.g. string tensors in foreign checkpoints)
            shape = tuple(dims[k] for k in range(ndim.value))
            a = np.empty(shape, dtype=_NP[dtype.value])
            if a.nbytes != nbytes.value:
                raise IOError("size mismatch for %s" % name.value.decode())
            if lib.dfmio_bundle_reader_read(h, i, a.ctypes.data, a.nbytes, 1 if verify_crc else 0) != 0:
                raise IOError(_err(lib))
            out[name.value.decode()] = torch.from_numpy(a)
    finally:
        lib.dfmio_bundle_reader_close(h)
    return out


def ordered_code_slice_key(name, extents) -> bytes:
    """Pure-Python derivation of TF's EncodeTensorNameSlice (OrderedCode, tensorflow/core/lib/strings/ordered_code.h): the
    number 0, the escaped name, the rank, then (start, length) per dimension as signed increasing numbers; (-1, -1) = the
    whole axis. The C++ codec has its own encoder (csrc/io/tensor_bundle.cc slice_key); the tests hold the two against
    each other, and `inspect_checkpoint` can use this one to explain the keys of a partitioned variable."""
    raw = name if isinstance(name, (bytes, bytearray)) else name.encode()

    def num_inc(v):
        b = v.to_bytes((v.bit_length() + 7) // 8, "big") if v else b""
        return bytes([len(b)]) + b

    def signed_inc(v):
        x = ~v if v < 0 else v
        n = 1
        while x.bit_length() > 7 * n - 1:
            n += 1
        out = bytearray((v & ((1 << 80) - 1)).to_bytes(10, "big")[10 - n:])
        for i in range(n):  # n leading one bits mark the length; xor keeps negatives below positives
            out[i // 8] ^= 0x80 >> (i % 8)
        return bytes(out)

    key = num_inc(0) + bytes(raw).replace(b"\x00", b"\x00\xff").replace(b"\xff", b"\xff\x00") + b"\x00\x01"
    key += num_inc(len(extents))
    for start, length in extents:
        key += signed_inc(start) + signed_inc(length)
    return key
