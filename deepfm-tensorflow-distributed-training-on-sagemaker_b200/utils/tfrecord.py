"""Pure-Python TFRecord framing + `tf.train.Example` codec (no TensorFlow).

Wire format (what `tf.python_io.TFRecordWriter` at TOOL:12,33
and `tf.data.TFRecordDataset` at HVD:93 / PS:108 produce and consume):

    record  = u64 length (LE) | u32 masked_crc32c(length) | data | u32 masked_crc32c(data)
    masked  = ((crc >> 15) | (crc << 17)) + 0xa282ead8   (mod 2**32)
    Example = { features(1): Features{ feature(1): map<string, Feature> } }
    Feature = oneof bytes_list(1) | float_list(2) | int64_list(3)   (packed values)

Schema used by the trainers (TOOL:25-32, HVD:78-83): "label" float x1,
"ids" int64 xF, "values" float xF.

This module is the slow, obviously-correct implementation: it is the oracle for
the C++ reader/parser in `csrc/io/` and the fallback when the native library is
not built.
"""
from __future__ import annotations

import struct
from typing import BinaryIO, Dict, Iterable, Iterator, List, Sequence, Tuple, Union

import numpy as np

# ----------------------------------------------------------------------------- crc32c
_POLY = 0x82F63B78


def _make_table() -> List[int]:
    tbl = []
    for i in range(256):
        c = i
        for _ in range(8):
            c = (c >> 1) ^ _POLY if c & 1 else c >> 1
        tbl.append(c)
    return tbl


_TABLE = _make_table()


def crc32c(data: bytes, crc: int = 0) -> int:
    """Castagnoli CRC-32C (software table). The C++ library uses the SSE4.2 instruction."""
    c = crc ^ 0xFFFFFFFF
    tbl = _TABLE
    for b in data:
        c = tbl[(c ^ b) & 0xFF] ^ (c >> 8)
    return c ^ 0xFFFFFFFF


def masked_crc32c(data: bytes) -> int:
    c = crc32c(data)
    return ((((c >> 15) | (c << 17)) & 0xFFFFFFFF) + 0xA282EAD8) & 0xFFFFFFFF


def unmask_crc(m: int) -> int:
    r = (m - 0xA282EAD8) & 0xFFFFFFFF
    return ((r >> 17) | (r << 15)) & 0xFFFFFFFF


# ----------------------------------------------------------------------------- framing
class TFRecordError(IOError):
    pass


def frame_record(data: bytes) -> bytes:
    hdr = struct.pack("<Q", len(data))
    return hdr + struct.pack("<I", masked_crc32c(hdr)) + data + struct.pack("<I", masked_crc32c(data))


class TFRecordWriter:
    """Minimal stand-in for `tf.python_io.TFRecordWriter` (TOOL:12)."""

    def __init__(self, path_or_file: Union[str, BinaryIO]):
        self._own = isinstance(path_or_file, str)
        self._f = open(path_or_file, "wb") if self._own else path_or_file

    def write(self, data: bytes) -> None:
        self._f.write(frame_record(data))

    def flush(self) -> None:
        self._f.flush()

    def close(self) -> None:
        if self._own:
            self._f.close()
        else:
            self._f.flush()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


def read_records(path_or_file: Union[str, BinaryIO], check_crc: bool = True) -> Iterator[bytes]:
    """Yield raw record payloads. Works on regular files, FIFOs and stdin-like streams
    (it only ever calls `read`, never `seek`) — the pipe-mode property of HVD:105."""
    own = isinstance(path_or_file, str)
    f = open(path_or_file, "rb") if own else path_or_file
    try:
        while True:
            hdr = _read_exact(f, 12, allow_eof=True)
            if hdr is None:
                return
            (length,) = struct.unpack("<Q", hdr[:8])
            if check_crc and struct.unpack("<I", hdr[8:])[0] != masked_crc32c(hdr[:8]):
                raise TFRecordError("corrupted record length CRC")
            body = _read_exact(f, length + 4, allow_eof=False)
            data = body[:length]
            if check_crc and struct.unpack("<I", body[length:])[0] != masked_crc32c(data):
                raise TFRecordError("corrupted record data CRC")
            yield data
    finally:
        if own:
            f.close()


def _read_exact(f: BinaryIO, n: int, allow_eof: bool):
    buf = b""
    while len(buf) < n:
        chunk = f.read(n - len(buf))
        if not chunk:
            if allow_eof and not buf:
                return None
            raise TFRecordError("truncated record")
        buf += chunk
    return buf


# ----------------------------------------------------------------------------- protobuf bits
def _varint(n: int) -> bytes:
    if n < 0:
        n += 1 << 64
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        if n:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _read_varint(buf: bytes, pos: int) -> Tuple[int, int]:
    shift = 0
    val = 0
    while True:
        b = buf[pos]
        pos += 1
        val |= (b & 0x7F) << shift
        if not b & 0x80:
            return val, pos
        shift += 7
        if shift > 63:
            raise ValueError("varint too long")


def _ld(field: int, payload: bytes) -> bytes:
    return _varint((field << 3) | 2) + _varint(len(payload)) + payload


def encode_example(feature: Dict[str, Tuple[str, Sequence]]) -> bytes:
    """feature: name -> ("float" | "int64" | "bytes", values). Keys are emitted sorted,
    as protobuf's deterministic map serialisation does."""
    entries = b""
    for key in sorted(feature):
        kind, vals = feature[key]
        if kind == "float":
            packed = np.asarray(vals, dtype="<f4").tobytes()
            feat = _ld(2, _ld(1, packed))
        elif kind == "int64":
            packed = b"".join(_varint(int(v)) for v in vals)
            feat = _ld(3, _ld(1, packed))
        elif kind == "bytes":
            feat = _ld(1, b"".join(_ld(1, bytes(v)) for v in vals))
        else:
            raise ValueError(kind)
        entries += _ld(1, _ld(1, key.encode()) + _ld(2, feat))
    return _ld(1, entries)


def make_ctr_example(label: float, ids: Sequence[int], values: Sequence[float]) -> bytes:
    """The exact Example the reference converter writes (TOOL:25-32)."""
    return encode_example({"label": ("float", [label]), "ids": ("int64", ids), "values": ("float", values)})


def decode_example(buf: bytes) -> Dict[str, Tuple[str, list]]:
    """Generic Example decoder (accepts packed and unpacked repeated scalars)."""
    out: Dict[str, Tuple[str, list]] = {}
    pos = 0
    while pos < len(buf):
        tag, pos = _read_varint(buf, pos)
        if tag != (1 << 3 | 2):
            pos = _skip(buf, pos, tag & 7)
            continue
        n, pos = _read_varint(buf, pos)
        feats, pos = buf[pos:pos + n], pos + n
        p = 0
        while p < len(feats):
            t, p = _read_varint(feats, p)
            if t != (1 << 3 | 2):
                p = _skip(feats, p, t & 7)
                continue
            m, p = _read_varint(feats, p)
            entry, p = feats[p:p + m], p + m
            key, val = _decode_map_entry(entry)
            out[key] = val
    return out


def _skip(buf: bytes, pos: int, wt: int) -> int:
    if wt == 0:
        _, pos = _read_varint(buf, pos)
    elif wt == 1:
        pos += 8
    elif wt == 2:
        n, pos = _read_varint(buf, pos)
        pos += n
    elif wt == 5:
        pos += 4
    else:
        raise ValueError("unsupported wire type %d" % wt)
    return pos


def _decode_map_entry(entry: bytes):
    key, val, p = "", ("float", []), 0
    while p < len(entry):
        t, p = _read_varint(entry, p)
        n, p = _read_varint(entry, p)
        body, p = entry[p:p + n], p + n
        if t >> 3 == 1:
            key = body.decode()
        elif t >> 3 == 2:
            val = _decode_feature(body)
    return key, val


def _decode_feature(body: bytes):
    p = 0
    kind, vals = "float", []
    while p < len(body):
        t, p = _read_varint(body, p)
        n, p = _read_varint(body, p)
        lst, p = body[p:p + n], p + n
        fld = t >> 3
        q = 0
        if fld == 2:
            kind = "float"
            while q < len(lst):
                tt, q = _read_varint(lst, q)
                if tt & 7 == 2:
                    m, q = _read_varint(lst, q)
                    vals.extend(np.frombuffer(lst[q:q + m], dtype="<f4").tolist())
                    q += m
                else:
                    vals.append(struct.unpack("<f", lst[q:q + 4])[0])
                    q += 4
        elif fld == 3:
            kind = "int64"
            while q < len(lst):
                tt, q = _read_varint(lst, q)
                if tt & 7 == 2:
                    m, q = _read_varint(lst, q)
                    end = q + m
                    while q < end:
                        v, q = _read_varint(lst, q)
                        vals.append(v - (1 << 64) if v >= (1 << 63) else v)
                else:
                    v, q = _read_varint(lst, q)
                    vals.append(v - (1 << 64) if v >= (1 << 63) else v)
        elif fld == 1:
            kind = "bytes"
            while q < len(lst):
                tt, q = _read_varint(lst, q)
                m, q = _read_varint(lst, q)
                vals.append(lst[q:q + m])
                q += m
    return kind, vals


def parse_ctr_batch(records: Iterable[bytes], field_size: int):
    """Python twin of the vectorised `tf.parse_example` decode (HVD:76-89, PS:79-92):
    a batch of serialized Examples -> (label [B] f32, ids [B,F] i64, values [B,F] f32).
    Every record must carry exactly `field_size` ids and values (FixedLenFeature)."""
    labels, ids, vals = [], [], []
    for rec in records:
        ex = decode_example(rec)
        for k in ("label", "ids", "values"):
            if k not in ex:
                raise ValueError("Example is missing feature %r" % k)
        if len(ex["label"][1]) != 1:
            raise ValueError("label must have exactly 1 value")
        if len(ex["ids"][1]) != field_size or len(ex["values"][1]) != field_size:
            raise ValueError("expected %d ids/values per record, got %d/%d"
                             % (field_size, len(ex["ids"][1]), len(ex["values"][1])))
        labels.append(ex["label"][1][0])
        ids.append(ex["ids"][1])
        vals.append(ex["values"][1])
    return (np.asarray(labels, dtype=np.float32),
            np.asarray(ids, dtype=np.int64).reshape(-1, field_size),
            np.asarray(vals, dtype=np.float32).reshape(-1, field_size))
