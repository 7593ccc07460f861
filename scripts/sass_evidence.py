#!/usr/bin/env python
"""Regenerate profiles/sass_evidence.md from the built library (no GPU needed):

    python scripts/sass_evidence.py > profiles/sass_evidence.md

Counts, per kernel, the SASS mnemonics that prove the Blackwell paths are really in the binary (B200_PROFILING.md):
tcgen05 (UTCHMMA / UTCBAR / LDTM / UTCATOMSWS), TMA (UTMALDG), mbarrier (SYNCS), PDL (ACQBULK), vector reductions
(REDG), NVLS multimem (LDGMC / multimem stores), thread-block clusters (UCGABAR_*) and system-scope fences.
"""
from __future__ import annotations

import collections
import os
import re
import subprocess
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(REPO, "deepfm-tensorflow-distributed-training-on-sagemaker_b200", "ops", "_lib", "libdeepfm_b200.so")
GROUPS = [  # (label, regex on the opcode)
    ("UTCHMMA", r"^UTCHMMA"), ("UTCBAR", r"^UTCBAR"), ("LDTM", r"^LDTM"), ("UTCATOMSWS", r"^UTCATOMSWS"),
    ("UTMALDG", r"^UTMALDG"), ("SYNCS", r"^SYNCS"), ("ELECT", r"^ELECT"), ("ACQBULK", r"^ACQBULK"),
    ("REDG", r"^REDG?\."), ("ATOMG", r"^ATOMG"), ("LDGMC", r"^LDGMC"), ("STG.SYS", r"^STG\.E\.(128|64)\.STRONG\.SYS"),
    ("UCGABAR", r"^UCGABAR"), ("MEMBAR.SYS", r"^MEMBAR\.(SC|ALL)\.SYS"), ("MATCH", r"^MATCH"),
]
SAMPLES = ["UTCHMMA", "UTMALDG", "LDTM", "UTCBAR", "LDGMC", "UCGABAR_ARV", "ACQBULK", "REDG.E.ADD.F32x4", "MATCH.ANY"]


def main() -> int:
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    per = collections.OrderedDict()
    samples = collections.OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            per[cur] = collections.Counter()
            continue
        m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Za-z0-9_.]+)", line)
        if not m or cur is None:
            continue
        op = m.group(1)
        for label, rx in GROUPS:
            if re.search(rx, op):
                per[cur][label] += 1
        for s in SAMPLES:
            if op.startswith(s) and len(samples.setdefault(s, [])) < 2:
                samples[s].append(line.rstrip())
    names = subprocess.run(["c++filt"], input="\n".join(per), capture_output=True, text=True).stdout.splitlines()
    print("# SASS evidence (`cuobjdump -sass ops/_lib/libdeepfm_b200.so`, sm_100a; regenerate with `python scripts/sass_evidence.py`)\n")
    print("Per-kernel counts of Blackwell-specific SASS mnemonics: UTCHMMA = tcgen05.mma, UTCBAR = tcgen05.commit, LDTM = tcgen05.ld,")
    print("UTCATOMSWS = TMEM alloc/dealloc, UTMALDG = TMA cp.async.bulk.tensor, SYNCS = mbarrier, ACQBULK = griddepcontrol.wait (PDL),")
    print("REDG = red.global.add (v4.f32 in the gradient kernels), LDGMC = NVLS multimem.ld_reduce, UCGABAR = cluster barrier")
    print("(barrier.cluster arrive/wait around the DSMEM exchange), MEMBAR.SYS = system-scope fences, MATCH = match.any,")
    print("STG.SYS = system-scope vector stores (multimem.st compiles to STG.E.128.STRONG.SYS on the multicast address; the")
    print("P2P flavour stores to peer addresses with plain STG.E.128 and orders them with the final MEMBAR.SYS).\n")
    print("```")
    for mangled, nice in zip(per, names):
        c = per[mangled]
        if not c:
            continue
        short = re.sub(r"\(.*", "", nice).replace("void ", "").replace("dfm::", "").replace("(anonymous namespace)::", "")
        print("%-46s %s" % (short[:46], "  ".join("%s=%d" % (k, c[k]) for k, _ in GROUPS if c[k])))
    print("```\n")
    # ---- static resources per kernel (cuobjdump -res-usage): registers, spill stack, static shared memory
    res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True, check=True).stdout
    rows, fn = [], None
    for line in res.splitlines():
        m = re.search(r"Function (\S+):", line)
        if m:
            fn = m.group(1)
            continue
        m = re.search(r"REG:(\d+) STACK:(\d+) SHARED:(\d+) LOCAL:(\d+)", line)
        if m and fn:
            rows.append((fn,) + tuple(int(x) for x in m.groups()))
            fn = None
    nice = subprocess.run(["c++filt"], input="\n".join(r[0] for r in rows), capture_output=True, text=True).stdout.splitlines()
    print("Static resources (`cuobjdump -res-usage`; STACK > 0 = spill / local arrays, SHARED = static shared memory, the GEMMs")
    print("add 189-200 KB of dynamic shared memory at launch):\n")
    print("```")
    print("%-46s %5s %6s %7s %6s" % ("kernel", "REG", "STACK", "SHARED", "LOCAL"))
    for r, n in zip(rows, nice):
        short = re.sub(r"\(.*", "", n).replace("void ", "").replace("dfm::", "").replace("(anonymous namespace)::", "")
        print("%-46s %5d %6d %7d %6d" % (short[:46], r[1], r[2], r[3], r[4]))
    print("```\n\nSample lines:\n```")
    for s, lines in samples.items():
        for ln in lines:
            print(re.sub(r"\s+/\* 0x[0-9a-f]+ \*/\s*$", "", ln))
    print("```")
    return 0


if __name__ == "__main__":
    sys.exit(main())
