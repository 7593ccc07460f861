"""TensorBoard event files (`events.out.tfevents.*`), written without TensorFlow.

The reference's Estimator drops scalar summaries (loss, global_step/sec) into `model_dir` and evaluation metrics into
`model_dir/eval` as TFRecord-framed `Event` protos ("event files in model_dir", 5.5: "summaries to event
files"); TensorBoard pointed at a `model_dir` written by this framework shows the same curves.

Wire format (tensorflow/core/util/event.proto, framework/summary.proto), only what scalars need:
    Event   { double wall_time = 1; int64 step = 2; string file_version = 3; Summary summary = 5; }
    Summary { repeated Value value = 1; }      Value { string tag = 1; float simple_value = 2; }
"""
from __future__ import annotations

import os
import socket
import struct
import time
from typing import Dict, Iterator, List, Optional, Tuple

from .tfrecord import TFRecordWriter, _ld, _read_varint, _skip, _varint, read_records


def _event(wall_time: float, step: int, file_version: Optional[str] = None,
           scalars: Optional[Dict[str, float]] = None) -> bytes:
    out = b"\x09" + struct.pack("<d", wall_time)                          # field 1, 64-bit
    if step:
        out += b"\x10" + _varint(step & 0xFFFFFFFFFFFFFFFF)               # field 2, varint
    if file_version is not None:
        out += _ld(3, file_version.encode())
    if scalars:
        summary = b"".join(_ld(1, _ld(1, tag.encode()) + b"\x15" + struct.pack("<f", float(v)))   # Value: tag, simple_value
                           for tag, v in scalars.items())
        out += _ld(5, summary)
    return out


class EventFileWriter:
    """Appends scalar summaries to `<logdir>/events.out.tfevents.<time>.<host>`; first record is the version header."""

    def __init__(self, logdir: str, filename_suffix: str = ""):
        os.makedirs(logdir, exist_ok=True)
        self.path = os.path.join(logdir, "events.out.tfevents.%010d.%s%s" % (int(time.time()), socket.gethostname(),
                                                                             filename_suffix))
        self._w = TFRecordWriter(self.path)
        self._w.write(_event(time.time(), 0, file_version="brain.Event:2"))
        self._w.flush()

    def add_scalars(self, scalars: Dict[str, float], step: int, wall_time: Optional[float] = None) -> None:
        self._w.write(_event(time.time() if wall_time is None else wall_time, int(step), scalars=scalars))

    def add_scalar(self, tag: str, value: float, step: int) -> None:
        self.add_scalars({tag: value}, step)

    def flush(self) -> None:
        self._w.flush()

    def close(self) -> None:
        self._w.close()


def read_events(path: str) -> Iterator[Tuple[float, int, Optional[str], List[Tuple[str, float]]]]:
    """(wall_time, step, file_version, [(tag, value), ...]) per record - the inverse of the writer, for tests and tools."""
    for rec in read_records(path):
        wall, step, version, scalars, pos = 0.0, 0, None, [], 0
        while pos < len(rec):
            key, pos = _read_varint(rec, pos)
            field, wt = key >> 3, key & 7
            if field == 1 and wt == 1:
                wall = struct.unpack_from("<d", rec, pos)[0]
                pos += 8
            elif field == 2 and wt == 0:
                step, pos = _read_varint(rec, pos)
            elif field in (3, 5) and wt == 2:
                n, pos = _read_varint(rec, pos)
                body = rec[pos:pos + n]
                pos += n
                if field == 3:
                    version = body.decode()
                else:
                    scalars += _summary_values(body)
            else:
                pos = _skip(rec, pos, wt)
        yield wall, step, version, scalars


def _summary_values(body: bytes) -> List[Tuple[str, float]]:
    out, pos = [], 0
    while pos < len(body):
        key, pos = _read_varint(body, pos)
        if key >> 3 == 1 and key & 7 == 2:
            n, pos = _read_varint(body, pos)
            val, vpos, tag, simple = body[pos:pos + n], 0, "", None
            pos += n
            while vpos < len(val):
                k, vpos = _read_varint(val, vpos)
                if k >> 3 == 1 and k & 7 == 2:
                    m, vpos = _read_varint(val, vpos)
                    tag = val[vpos:vpos + m].decode()
                    vpos += m
                elif k >> 3 == 2 and k & 7 == 5:
                    simple = struct.unpack_from("<f", val, vpos)[0]
                    vpos += 4
                else:
                    vpos = _skip(val, vpos, k & 7)
            if simple is not None:
                out.append((tag, simple))
        else:
            pos = _skip(body, pos, key & 7)
    return out
