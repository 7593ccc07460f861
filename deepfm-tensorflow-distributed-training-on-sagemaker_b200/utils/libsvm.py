"""libsvm -> TFRecord converter (reference: tools/libsvm_to_tfrecord.py, TOOL:5-37).

Same behaviour — one `Example{label, ids, values}` per `label id:val id:val ...`
line — minus the reference's hard-coded paths (TOOL:43-44) and unused TF session
(TOOL:40).  Adds what the reference leaves to "upstream": optional padding /
truncation to exactly `field_size` entries, because the trainers parse with
FixedLenFeature([field_size]) (HVD:81-82) and fail on anything else.

CLI:  python -m deepfm_b200.utils.libsvm in.libsvm out.tfrecords [--field_size 39]
"""
from __future__ import annotations

import argparse
from typing import Iterable, List, Optional, Tuple

from .tfrecord import TFRecordWriter, make_ctr_example


def parse_libsvm_line(line: str) -> Tuple[float, List[int], List[float]]:
    """`label id:val id:val ...` split on single spaces (TOOL:16,21)."""
    toks = line.rstrip("\r\n").split(" ")
    label = float(toks[0])
    ids, vals = [], []
    for tok in toks[1:]:
        if not tok:
            continue
        i, v = tok.split(":")
        ids.append(int(i))
        vals.append(float(v))
    return label, ids, vals


def fit_to_fields(ids: List[int], vals: List[float], field_size: int, pad_id: int = 0):
    """Pad with (pad_id, 0.0) — a zero value contributes nothing to any DeepFM term — or truncate."""
    if len(ids) > field_size:
        return ids[:field_size], vals[:field_size]
    n = field_size - len(ids)
    return ids + [pad_id] * n, vals + [0.0] * n


def convert_lines(lines: Iterable[str], writer: TFRecordWriter, field_size: Optional[int] = None,
                  pad_id: int = 0) -> int:
    n = 0
    for line in lines:
        if not line.strip():
            continue
        label, ids, vals = parse_libsvm_line(line)
        if field_size is not None:
            ids, vals = fit_to_fields(ids, vals, field_size, pad_id)
        writer.write(make_ctr_example(label, ids, vals))
        n += 1
    return n


def convert_tfrecords(input_filename: str, output_filename: str, field_size: Optional[int] = None,
                      pad_id: int = 0, use_native: bool = True, threads: int = 0) -> int:
    """Same name and positional arguments as the reference function (TOOL:5). With the host IO library built, the
    multi-threaded C++ converter does the work (byte-identical output, ~100x the rate of the Python loop below)."""
    if use_native:
        try:
            from ..data import native
            if native.available():
                return native.libsvm_convert_native(input_filename, output_filename, field_size, pad_id, threads)
        except (ImportError, OSError):
            pass
    with open(input_filename, "r") as fin, TFRecordWriter(output_filename) as w:
        return convert_lines(fin, w, field_size, pad_id)


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description="libsvm -> TFRecord (DeepFM CTR schema)")
    ap.add_argument("input")
    ap.add_argument("output")
    ap.add_argument("--field_size", type=int, default=None,
                    help="pad/truncate every row to exactly this many id:value pairs")
    ap.add_argument("--pad_id", type=int, default=0)
    ap.add_argument("--python", action="store_true", help="use the pure-Python converter")
    ap.add_argument("--threads", type=int, default=0, help="converter threads (0 = one per hardware thread)")
    a = ap.parse_args(argv)
    n = convert_tfrecords(a.input, a.output, a.field_size, a.pad_id, not a.python, a.threads)
    print("converted %d rows: %s -> %s" % (n, a.input, a.output))
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
