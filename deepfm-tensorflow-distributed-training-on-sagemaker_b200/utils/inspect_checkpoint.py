"""List / print the variables of a checkpoint (TensorFlow tensor-bundle format), like TF's `inspect_checkpoint` tool:

    python -m deepfm_b200.utils.inspect_checkpoint <model_dir | checkpoint prefix | servable dir> [--tensor NAME] [--stats]

Works on checkpoints written by this framework and on ones written by the reference's TF-1 Estimator (same format and
variable names: fm_bias, fm_w, fm_v, Deep-part/mlp0/weights, ..., <var>/Adam, beta1_power, global_step)."""
from __future__ import annotations

import argparse
import os
import sys
from typing import Optional

from .checkpoint import latest_checkpoint
from .tf_bundle import read_bundle


def resolve_prefix(path: str) -> Optional[str]:
    """model_dir -> latest checkpoint prefix; servable dir -> variables/variables; a prefix stays a prefix."""
    if os.path.exists(path + ".index"):
        return path
    if os.path.isdir(path):
        if os.path.exists(os.path.join(path, "variables", "variables.index")):
            return os.path.join(path, "variables", "variables")
        subs = sorted((d for d in os.listdir(path) if d.isdigit()), key=int)
        if subs and os.path.exists(os.path.join(path, subs[-1], "variables", "variables.index")):
            return os.path.join(path, subs[-1], "variables", "variables")
        return latest_checkpoint(path)
    return None


def main(argv=None, out=sys.stdout) -> int:
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("path")
    ap.add_argument("--tensor", default="", help="print this variable's values")
    ap.add_argument("--stats", action="store_true", help="add min / max / mean / l2 norm per variable")
    a = ap.parse_args(argv)
    prefix = resolve_prefix(a.path)
    if prefix is None:
        print("no checkpoint found at %s" % a.path, file=sys.stderr)
        return 1
    sd = read_bundle(prefix)
    print("checkpoint: %s  (%d variables, %d elements)" % (prefix, len(sd), sum(t.numel() for t in sd.values())), file=out)
    if a.tensor:
        if a.tensor not in sd:
            print("variable %r not in checkpoint" % a.tensor, file=sys.stderr)
            return 1
        print(a.tensor, tuple(sd[a.tensor].shape), sd[a.tensor].dtype, file=out)
        print(sd[a.tensor], file=out)
        return 0
    for name in sorted(sd):
        t = sd[name]
        line = "%-48s %-10s %s" % (name, str(t.dtype).replace("torch.", ""), list(t.shape))
        if a.stats and t.numel() and t.is_floating_point():
            f = t.float()
            line += "  min %.4g max %.4g mean %.4g l2 %.4g" % (f.min(), f.max(), f.mean(), f.norm())
        elif t.numel() == 1:
            line += "  = %s" % t.reshape(-1)[0].item()
        print(line, file=out)
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
