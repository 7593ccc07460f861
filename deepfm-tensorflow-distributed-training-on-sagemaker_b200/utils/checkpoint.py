"""Checkpoints in the reference's own on-disk format (TF tensor bundle, Estimator file naming).

The reference saves through `tf.estimator.Estimator(model_dir=...)` (PS:435, HVD:365-368): files
`model.ckpt-<global_step>.{index,data-00000-of-00001}` plus a `checkpoint` text proto naming the latest one,
keep-5 rotation, restore-latest on start. This module writes/reads exactly that, with the reference's variable
names (`fm_bias`, `fm_w`, `fm_v`, `Deep-part/mlp{i}/{weights,biases}`, `Deep-part/deep_out/...`,
optimizer slots `<var>/Adam`, `<var>/Adam_1`, ..., `beta1_power`, `beta2_power`, `global_step`), with the intent that
a checkpoint produced here loads into the reference's `model_fn` and vice versa. That interchange is UNVERIFIED against
a TensorFlow-written file (none can be produced offline): the format follows SURVEY Appendix A and round-trips its own
output, sharded data files and partitioned variables (`slices`) included. MLP weights are stored `[in,out]` as TF does
and transposed to the native `[out,in]` on load.
"""
from __future__ import annotations

import glob
import os
import re
from typing import Dict, List, Optional

import torch

from ..models.params import from_tf_layout, tf_name, to_tf_layout
from ..optim.tf1 import SLOT_SUFFIXES
from .tf_bundle import read_bundle, write_bundle

_CKPT_RE = re.compile(r"model\.ckpt-(\d+)\.index$")


def native_to_tf(sd: Dict[str, torch.Tensor], opt_name: str, beta_powers: Optional[Dict[str, float]] = None):
    out = {}
    for k, v in sd.items():
        if k == "global_step":
            out["global_step"] = v.reshape(()).to(torch.int64)
            continue
        if "/" in k:  # optimizer slot of a native variable
            var, suffix = k.rsplit("/", 1)
            out["%s/%s" % (tf_name(var)[0], suffix)] = to_tf_layout(var, v.float())
        else:
            out[tf_name(k)[0]] = to_tf_layout(k, v.float())
    for k, val in (beta_powers or {}).items():
        out[k] = torch.tensor(val, dtype=torch.float32)
    return out


def tf_to_native(tf_sd: Dict[str, torch.Tensor], native_names: List[str], opt_name: str):
    """Inverse mapping for the variables this model has; unknown TF variables are ignored."""
    out = {}
    for n in native_names:
        t = tf_name(n)[0]
        if t in tf_sd:
            out[n] = from_tf_layout(n, tf_sd[t])
        for suffix in SLOT_SUFFIXES[opt_name]:
            key = "%s/%s" % (t, suffix)
            if key in tf_sd:
                out["%s/%s" % (n, suffix)] = from_tf_layout(n, tf_sd[key])
    if "global_step" in tf_sd:
        out["global_step"] = tf_sd["global_step"].reshape(1).to(torch.int64)
    return out


def all_checkpoints(model_dir: str) -> List[int]:
    if not os.path.isdir(model_dir):
        return []
    steps = []
    for f in os.listdir(model_dir):
        m = _CKPT_RE.match(f)
        if m:
            steps.append(int(m.group(1)))
    return sorted(steps)


def latest_checkpoint(model_dir: str) -> Optional[str]:
    """Honours the `checkpoint` state file like `tf.train.latest_checkpoint`; falls back to the highest step."""
    state = os.path.join(model_dir, "checkpoint")
    if os.path.exists(state):
        for line in open(state):
            m = re.match(r'\s*model_checkpoint_path:\s*"(.*)"', line)
            if m:
                p = m.group(1)
                p = p if os.path.isabs(p) else os.path.join(model_dir, p)
                if os.path.exists(p + ".index"):
                    return p
    for step in reversed(all_checkpoints(model_dir)):  # no usable state file: newest checkpoint with both files present
        p = os.path.join(model_dir, "model.ckpt-%d" % step)
        if glob.glob(glob.escape(p) + ".data-*"):  # one data file, or several behind a sharded (PS-style) index
            return p
    return None


def save_checkpoint(model_dir: str, state_dict: Dict[str, torch.Tensor], opt, keep: int = 5) -> str:
    os.makedirs(model_dir, exist_ok=True)
    step = int(state_dict["global_step"].reshape(-1)[0])
    prefix = os.path.join(model_dir, "model.ckpt-%d" % step)
    saved_step, opt.step = opt.step, step
    try:
        tf_sd = native_to_tf(state_dict, opt.name, opt.scalar_state())
    finally:
        opt.step = saved_step
    # atomic publication, like TF's Saver: a crash while writing never leaves a half-written "latest" checkpoint
    tmp = prefix + ".tmp-%d" % os.getpid()
    write_bundle(tmp, tf_sd)
    for ext in (".data-00000-of-00001", ".index"):  # the index last: its presence marks the checkpoint complete
        os.replace(tmp + ext, prefix + ext)
    steps = all_checkpoints(model_dir)
    for old in steps[:-keep] if keep > 0 else []:
        stem = os.path.join(model_dir, "model.ckpt-%d" % old)
        for path in [stem + ".index"] + glob.glob(glob.escape(stem) + ".data-*"):
            try:
                os.remove(path)
            except OSError:
                pass
    steps = all_checkpoints(model_dir)
    state = os.path.join(model_dir, "checkpoint")
    with open(state + ".tmp", "w") as f:
        f.write('model_checkpoint_path: "model.ckpt-%d"\n' % step)
        for s in steps:
            f.write('all_model_checkpoint_paths: "model.ckpt-%d"\n' % s)
    os.replace(state + ".tmp", state)
    return prefix


def load_checkpoint(prefix: str, native_names: List[str], opt_name: str) -> Dict[str, torch.Tensor]:
    return tf_to_native(read_bundle(prefix), native_names, opt_name)
