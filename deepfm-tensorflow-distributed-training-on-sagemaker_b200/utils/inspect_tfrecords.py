"""Check TFRecord training files against the trainer's schema before a job is started:

    python -m deepfm_b200.utils.inspect_tfrecords data/train/tr*.tfrecords --field_size 39 [--feature_size 117581]

Every record is CRC-checked and parsed with the fixed schema the trainers use (label float x1, ids int64 x field_size,
values float x field_size; TOOL:25-32, HVD:78-83). Reports the record count, label mean, id range and the share of
zero-valued (padding) entries. With --feature_size, ids outside [0, feature_size) are counted: the reference's gather
would fail on them mid-job; the kernels here treat them as padding, which is safe but silently drops the feature.
Exit status 1 if a file is corrupt, does not match the schema, or holds out-of-range ids.
"""
from __future__ import annotations

import argparse
import sys
from typing import List, Optional

import torch

from ..data.pipeline import InputPipeline


def inspect(paths: List[str], field_size: int, feature_size: Optional[int] = None, batch_size: int = 4096,
            threads: int = 4) -> dict:
    n = 0
    lab_sum = 0.0
    id_min, id_max = None, None
    zero_vals = 0
    out_of_range = 0
    for b in InputPipeline(paths, field_size, batch_size, 1, (1, 0), False, threads, False):
        ids, vals, labels = b.ids, b.vals, b.labels
        n += b.rows
        lab_sum += float(labels.double().sum())
        lo, hi = int(ids.min()), int(ids.max())
        id_min = lo if id_min is None else min(id_min, lo)
        id_max = hi if id_max is None else max(id_max, hi)
        zero_vals += int((vals == 0).sum())
        if feature_size is not None:
            out_of_range += int(((ids < 0) | (ids >= feature_size)).sum())
        b.release()
    return {"files": len(paths), "records": n, "label_mean": lab_sum / max(n, 1), "id_min": id_min, "id_max": id_max,
            "zero_value_share": zero_vals / max(n * field_size, 1), "ids_out_of_range": out_of_range,
            "feature_size": feature_size}


def main(argv=None, out=sys.stdout) -> int:
    ap = argparse.ArgumentParser(description="validate TFRecord files against the DeepFM input schema")
    ap.add_argument("paths", nargs="+")
    ap.add_argument("--field_size", type=int, required=True)
    ap.add_argument("--feature_size", type=int, default=None)
    a = ap.parse_args(argv)
    try:
        r = inspect(a.paths, a.field_size, a.feature_size)
    except (IOError, ValueError) as e:
        print("FAILED: %s" % e, file=out)
        return 1
    for k, v in r.items():
        print("%-18s %s" % (k, "%.6g" % v if isinstance(v, float) else v), file=out)
    if r["ids_out_of_range"]:
        print("FAILED: %d ids outside [0, %d)" % (r["ids_out_of_range"], a.feature_size), file=out)
        return 1
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
