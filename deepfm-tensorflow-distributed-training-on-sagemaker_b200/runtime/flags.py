"""Flag system: the union of the two reference trainers' flag surfaces (PS:35-71 — 33 flags; HVD:40-68 — 28 flags),
parsed the way SageMaker feeds them (`hyperparameters` dict -> `--key value` argv, booleans as True/False
strings, NB-PS:89-94), plus the knobs the B200 framework adds.

Differences from the reference on purpose:
  * `SM_HOSTS` / `SM_CURRENT_HOST` / `SM_NUM_GPUS` / `SM_NUM_CPUS` are OPTIONAL (the reference reads them at import and
    crashes without them, PS:64-67 / HVD:61-62); single-box defaults are used instead.
  * unknown flags are tolerated and reported (SageMaker passes e.g. `--model_dir`, NB-HVD), never fatal.
  * `--optimizer GD` works (advertised but missing in the reference, PS:51 vs PS:260-267).
  * `log_steps` is honoured (declared but never read in the reference, PS:47).
"""
from __future__ import annotations

import json
import os
import sys
from types import SimpleNamespace
from typing import Any, Dict, List, Optional, Tuple


def _env_hosts(env) -> List[str]:
    try:
        h = json.loads(env.get("SM_HOSTS", "") or "null")
        if isinstance(h, list) and h:
            return [str(x) for x in h]
    except ValueError:
        pass
    return ["localhost"]


def _env_int(env, key, default):
    try:
        return int(env.get(key, default))
    except (TypeError, ValueError):
        return default


def flag_definitions(env=os.environ) -> List[Tuple[str, type, Any, str]]:
    hosts = _env_hosts(env)
    return [
        # ---- model (PS:38-56 / HVD:41-55)
        ("feature_size", int, 0, "Number of features"),
        ("field_size", int, 0, "Number of fields"),
        ("embedding_size", int, 32, "Embedding size"),
        ("deep_layers", str, "256,128,64", "deep layers"),
        ("dropout", str, "0.5,0.5,0.5", "dropout KEEP probabilities per deep layer (tf.nn.dropout keep_prob)"),
        ("batch_norm", bool, False, "perform batch normalization after ReLU"),
        ("batch_norm_decay", float, 0.9, "decay for the BN moving averages"),
        ("l2_reg", float, 0.0001, "L2 regularization on fm_w and fm_v"),
        ("loss_type", str, "log_loss", "loss type {log_loss} (accepted for parity; only log_loss exists, PS:50)"),
        ("optimizer", str, "Adam", "optimizer type {Adam, Adagrad, GD, Momentum, ftrl}"),
        ("learning_rate", float, 0.0005, "learning rate"),
        # ---- run
        ("num_epochs", int, 10, "Number of epochs"),
        ("batch_size", int, 64, "batch size per worker"),
        ("log_steps", int, 1000, "log loss / throughput every N steps"),
        ("task_type", str, "train", "task type {train, infer, eval, export}"),
        ("training_data_dir", str, "", "training data dir (tr*.tfrecords, searched recursively)"),
        ("val_data_dir", str, "", "validation data dir (va*.tfrecords, te*.tfrecords)"),
        ("model_dir", str, "", "checkpoint dir (PS trainer name)"),
        ("checkpoint_dir", str, "", "checkpoint dir (HVD trainer name; alias of model_dir)"),
        ("servable_model_dir", str, "", "export servable model here"),
        ("clear_existing_model", bool, False, "clear existing checkpoints before training (HVD:334-340)"),
        # ---- cluster (PS:57-67 / HVD:61-64)
        ("dist_mode", int, 0, "PS trainer: 0 single, 1 local test, 2 multi-worker (kept for parity; see --strategy)"),
        ("ps_hosts", str, "", "comma-separated ps hosts (manual cluster spec)"),
        ("worker_hosts", str, "", "comma-separated worker hosts (manual cluster spec)"),
        ("chief_hosts", str, "", "comma-separated chief hosts (the reference uses this flag without defining it, PS:297)"),
        ("job_name", str, "", "one of ps / worker / chief / evaluator"),
        ("task_index", int, 0, "index of the task within the job"),
        ("num_threads", int, 16, "loader / host threads"),
        ("hosts", list, hosts, "all cluster instance names (default: SM_HOSTS or ['localhost'])"),
        ("current_host", str, env.get("SM_CURRENT_HOST") or hosts[0], "this instance's name"),
        ("num_GPUs", int, _env_int(env, "SM_NUM_GPUS", 0), "GPUs on this instance"),
        ("num_CPUs", int, _env_int(env, "SM_NUM_CPUS", os.cpu_count() or 1), "vCPUs on this instance"),
        ("worker_per_host", int, 1, "worker processes per training instance"),
        # ---- data (PS:68-71 / HVD:63-68)
        ("pipe_mode", int, 0, "read TFRecords from per-channel FIFOs instead of files"),
        ("training_channel_name", str, "", "training channel name"),
        ("evaluation_channel_name", str, "", "evaluation channel name"),
        ("enable_s3_shard", bool, False, "data is already sharded per instance (ShardedByS3Key)"),
        ("enable_data_multi_path", bool, False, "pipe mode: every local worker has its own channel"),
        ("perform_shuffle", bool, False, "file mode: visit the records in a fresh pseudo-random order every epoch "
                                         "(declared but never read by the reference)"),
        # ---- deepfm_b200 additions
        ("strategy", str, "dp", "parallel strategy {dp, sharded, ps}: fused-allreduce data parallel, row-sharded "
                                "embeddings with fused all-to-all, async parameter server"),
        ("backend", str, "auto", "{auto, fused, torch}: fused = sm_100a kernels, torch = plain PyTorch"),
        ("l2_mode", str, "", "{exact, touched}; default exact for dp, touched for sharded/ps (SURVEY 7.3.3)"),
        ("comm", str, "auto", "{auto, p2p, nvls, nccl} gradient all-reduce transport for --strategy dp"),
        ("seed", int, 0, "initialisation / dropout seed"),
        ("max_steps", int, 0, "stop after this many optimizer steps (0 = by epochs)"),
        ("save_checkpoints_secs", int, 600, "checkpoint period in seconds (Estimator default)"),
        ("save_checkpoints_steps", int, 0, "checkpoint period in steps (overrides secs when > 0)"),
        ("keep_checkpoint_max", int, 5, "checkpoints to keep"),
        ("loader_threads", int, 4, "parser threads of the native loader (0 = as many as this rank has CPUs, 2..8)"),
        ("gpu_parse", bool, False, "decode the Example protos on the GPU (host only walks/CRC-checks/copies records)"),
        ("scale_lr_by_world", bool, True, "multiply learning_rate by the number of workers (HVD:149)"),
        ("data_base_dir", str, env.get("SM_INPUT_DIR", "/opt/ml/input") + "/data", "where pipe-mode FIFOs live"),
        ("metrics_file", str, "", "JSONL metrics output (default <model_dir>/metrics.jsonl)"),
        ("ps_staleness", int, 4, "strategy=ps: max steps a worker may run ahead of its last pull"),
        ("eval_start_delay_secs", int, 1000, "strategy=ps (train_and_evaluate): first evaluation not before this many "
                                             "seconds (EvalSpec start_delay_secs, PS:441)"),
        ("eval_throttle_secs", int, 1200, "strategy=ps (train_and_evaluate): at most one evaluation per this many seconds "
                                          "(EvalSpec throttle_secs, PS:441)"),
    ]


_TRUE = {"true", "1", "yes", "y", "t"}
_FALSE = {"false", "0", "no", "n", "f"}


def _to_bool(s: str) -> bool:
    v = str(s).strip().lower()
    if v in _TRUE:
        return True
    if v in _FALSE:
        return False
    raise ValueError("not a boolean: %r" % s)


def parse_flags(argv: Optional[List[str]] = None, env=os.environ) -> SimpleNamespace:
    argv = list(sys.argv[1:] if argv is None else argv)
    defs = {n: (t, d) for n, t, d, _ in flag_definitions(env)}
    vals: Dict[str, Any] = {n: d for n, (t, d) in defs.items()}
    unknown: List[str] = []
    i = 0
    while i < len(argv):
        tok = argv[i]
        i += 1
        if not tok.startswith("-"):
            unknown.append(tok)
            continue
        key = tok.lstrip("-")
        val: Optional[str] = None
        if "=" in key:
            key, val = key.split("=", 1)
        if key not in defs:
            if key.startswith("no") and key[2:] in defs and defs[key[2:]][0] is bool and val is None:
                vals[key[2:]] = False
                continue
            unknown.append(tok)
            if val is None and i < len(argv) and not argv[i].startswith("-"):
                unknown.append(argv[i])
                i += 1
            continue
        typ = defs[key][0]
        if typ is bool:
            if val is None and i < len(argv) and argv[i].strip().lower() in (_TRUE | _FALSE):
                val = argv[i]
                i += 1
            vals[key] = True if val is None else _to_bool(val)
            continue
        if val is None:
            if i >= len(argv):
                raise SystemExit("flag --%s needs a value" % key)
            val = argv[i]
            i += 1
        if typ is list:
            v = val.strip()
            vals[key] = json.loads(v) if v.startswith("[") else [x for x in v.split(",") if x]
        else:
            try:
                vals[key] = typ(val)
            except ValueError:
                raise SystemExit("flag --%s: cannot parse %r as %s" % (key, val, typ.__name__))
    f = SimpleNamespace(**vals)
    f.unknown = unknown
    # the two trainers name the checkpoint directory differently (PS:43 model_dir, HVD:57 checkpoint_dir)
    if not f.model_dir and f.checkpoint_dir:
        f.model_dir = f.checkpoint_dir
    if not f.checkpoint_dir and f.model_dir:
        f.checkpoint_dir = f.model_dir
    if not f.l2_mode:
        f.l2_mode = "exact" if f.strategy == "dp" else "touched"
    return f


def flags_help(env=os.environ) -> str:
    return "\n".join("  --%-26s %-6s default=%r  %s" % (n, t.__name__, d, h) for n, t, d, h in flag_definitions(env))


def print_flags(f: SimpleNamespace, out=sys.stdout) -> None:
    """The reference echoes its arguments for eyeball verification (PS:343-369, HVD:291-314)."""
    for k in sorted(vars(f)):
        if k != "unknown":
            print("%s %s" % (k, getattr(f, k)), file=out)
    if f.unknown:
        print("ignored unknown flags: %s" % f.unknown, file=out)
