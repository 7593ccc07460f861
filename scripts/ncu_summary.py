#!/usr/bin/env python
"""Turn an `.ncu-rep` captured with `ncu --set full` into the markdown table kept under profiles/.

    python scripts/ncu_summary.py gpurun_out/prof_step.ncu-rep > profiles/step_kernels_ncu.md   (table only)

Reads the report through `ncu -i <rep> --page raw --csv` (B200_PROFILING.md recipe); columns are looked up by metric
name, values are converted with the unit row, so ncu's per-column auto-scaling (byte / Kbyte / Mbyte) cannot skew a row.
"""
from __future__ import annotations

import csv
import io
import json
import os
import re
import subprocess
import sys

COLS = [
    ("dur_us", "gpu__time_duration.sum", "time"),
    ("grid", "Grid Size", "dims"),
    ("block", "Block Size", "dims"),
    ("regs", "launch__registers_per_thread", "num"),
    ("smem_KB", "launch__shared_mem_per_block_allocated", "bytes_kb"),
    ("dram_rd_MB", "dram__bytes_read.sum", "bytes_mb"),
    ("dram_wr_MB", "dram__bytes_write.sum", "bytes_mb"),
    ("dram_%", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "num"),
    ("L2_hit_%", "lts__t_sector_hit_rate.pct", "num"),
    ("warps_active_%", "sm__warps_active.avg.pct_of_peak_sustained_active", "num"),
    ("tensor_pipe_%", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "num"),
    ("warp_insts", "smsp__inst_executed.sum", "num"),
]
BYTE_SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
TIME_SCALE = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "usecond": 1.0, "nsecond": 1e-3, "msecond": 1e3, "second": 1e6}


def short_name(full: str) -> str:
    m = re.match(r"(?:void\s+)?(?:dfm::)?([A-Za-z0-9_]+(?:<[^>]*>)?)", full.replace("(anonymous namespace)::", ""))
    return m.group(1) if m else full[:48]


def fmt(x: float) -> str:
    if x != x:
        return "-"
    if abs(x) >= 1e5:
        return "%.3g" % x
    return ("%.3g" % x) if abs(x) < 100 else ("%.0f" % x)


def main() -> int:
    if len(sys.argv) < 2:
        print(__doc__)
        return 2
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    head, units, body = rows[0], rows[1], rows[2:]
    idx = {}
    for i, name in enumerate(head):
        idx.setdefault(name, i)
        if ".TriageCompute." in name:  # section-prefixed duplicates of plain metric names
            idx.setdefault(name.split(".TriageCompute.", 1)[1], i)
    out_rows = []
    for r in body:
        rec = {"kernel": short_name(r[idx["Kernel Name"]])}
        for label, metric, kind in COLS:
            i = idx.get(metric)
            if i is None or r[i] in ("", "n/a"):
                rec[label] = float("nan") if kind != "dims" else "-"
                continue
            if kind == "dims":
                dims = [int(v) for v in re.findall(r"\d+", r[i])]
                n = 1
                for v in dims:
                    n *= v
                rec[label] = n
                continue
            v = float(r[i].replace(",", ""))
            u = units[i]
            if kind == "time":
                v *= TIME_SCALE.get(u, 1.0)
            elif kind == "bytes_mb":
                v *= BYTE_SCALE.get(u, 1.0) / 1e6
            elif kind == "bytes_kb":
                v *= BYTE_SCALE.get(u.split("/")[0], 1.0) / 1e3
            rec[label] = v
        out_rows.append(rec)

    peaks = {}
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peaks = json.load(open(p))
    hbm = float(peaks.get("hbm_gbs", 6555.2))
    labels = ["kernel"] + [c[0] for c in COLS] + ["dram_GB/s", "frac_of_copy_peak"]
    print("| " + " | ".join(labels) + " |")
    print("|" + "---|" * len(labels))
    tot = 0.0
    for rec in out_rows:
        dur = rec["dur_us"]
        tot += dur if dur == dur else 0.0
        gbs = (rec["dram_rd_MB"] + rec["dram_wr_MB"]) * 1e6 / (dur * 1e-6) / 1e9 if dur == dur and dur > 0 else float("nan")
        cells = [rec["kernel"]]
        for label, _, kind in COLS:
            v = rec[label]
            cells.append(str(v) if kind == "dims" else fmt(v))
        cells += [fmt(gbs), fmt(gbs / hbm)]
        print("| " + " | ".join(cells) + " |")
    print()
    print("%d kernels, %.1f us summed (serialised, cold caches, under ncu replay: shares, not absolutes); "
          "copy peak used for the last column: %.1f GB/s (MEASURED_PEAKS.json)." % (len(out_rows), tot, hbm))
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
