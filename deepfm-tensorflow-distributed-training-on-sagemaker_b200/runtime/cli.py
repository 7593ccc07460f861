"""Task driver: `python -m deepfm_b200 --task_type {train,eval,infer,export} ...`

Reference: `main()` of both trainers (PS:341-467, HVD:289-431):
  * echo arguments, discover `tr*/va*/te*.tfrecords` recursively, shuffle the training file order
  * optional `--clear_existing_model`
  * train: file mode = per-epoch train -> chief-only evaluate loop (HVD:390-394); pipe mode = one train call with
    `num_epochs` baked into the stream, channel per local rank (HVD:403-405); then export on the chief
  * eval -> {auc, loss}; infer -> "%f\\n" per example into <val_data_dir>/pred.txt (PS:447-449); export
Strategy is chosen with --strategy {dp, sharded, ps} instead of by picking one of two scripts.
"""
from __future__ import annotations

import json
import os
import shutil
import sys
from typing import List, Optional

from ..models import DeepFMConfig
from ..optim.tf1 import TF1Optimizer
from ..parallel import shard as shard_rules
from ..parallel.dist import barrier, init_from_env, shutdown
from ..data.pipeline import InputPipeline, channel_fifo, discover_files
from .estimator import Estimator
from .flags import flags_help, parse_flags, print_flags


def build_config(F) -> DeepFMConfig:
    return DeepFMConfig(field_size=F.field_size, feature_size=F.feature_size, embedding_size=F.embedding_size,
                        deep_layers=F.deep_layers, dropout=F.dropout, l2_reg=F.l2_reg, batch_norm=F.batch_norm,
                        batch_norm_decay=F.batch_norm_decay, l2_mode=F.l2_mode)


def channel_names(env=os.environ) -> List[str]:
    """SM_CHANNELS, sorted the way SageMaker presents them (alphabetical: 'evaluation' first, HVD:375-380)."""
    try:
        names = json.loads(env.get("SM_CHANNELS", "") or "null")
        if isinstance(names, list) and names:
            return sorted(str(n) for n in names)
    except ValueError:
        pass
    return ["evaluation", "training"]


def configure_host_threads(F, env) -> None:
    """CPU threading knobs of the PS trainer (PS:405-411,428-430; README-EN:18-32): it disables MKL thread binding
    (`KMP_AFFINITY=verbose,disabled`), sets `OMP_NUM_THREADS` and TF's intra/inter-op pools to the vCPU count. Same idea
    here for the torch CPU backend and the host side of the loaders: split the vCPUs over the local workers."""
    import torch
    per_worker = max(1, int(F.num_CPUs) // max(1, int(F.worker_per_host) if F.worker_per_host > 1 else env.world_size))
    os.environ.setdefault("KMP_AFFINITY", "verbose,disabled" if os.environ.get("DFM_VERBOSE_THREADS") else "disabled")
    os.environ.setdefault("KMP_SETTINGS", "0")
    os.environ.setdefault("OMP_NUM_THREADS", str(per_worker))
    try:
        torch.set_num_threads(per_worker)
        torch.set_num_interop_threads(max(1, min(per_worker, F.num_threads)))
    except RuntimeError:
        pass  # inter-op pool already started


def main(argv: Optional[List[str]] = None) -> int:
    try:
        return _main(argv)
    except BaseException:
        # a device-side flag wait that timed out has trapped the kernel: say WHICH rank went missing before the CUDA
        # error ("unspecified launch failure") propagates (the analogue of Horovod's stall report, PDF p21)
        try:
            from ..ops import lib as _nlib
            wd = _nlib.watchdog_message()
            if wd:
                print("ERROR:" + wd, file=sys.stderr, flush=True)
        except Exception:
            pass
        raise


def _main(argv: Optional[List[str]] = None) -> int:
    argv = list(sys.argv[1:] if argv is None else argv)
    if "--help" in argv or "-h" in argv:
        print(__doc__)
        print(flags_help())
        return 0
    F = parse_flags(argv)
    use_cuda = None if F.backend == "auto" else (F.backend == "fused")
    if F.backend == "torch" and F.num_GPUs == 0:
        use_cuda = False
    env = init_from_env(use_cuda)
    configure_host_threads(F, env)
    if env.is_chief:
        print(sys.argv)
        print_flags(F)
        print("strategy %s on %d worker(s), backend %s" % (F.strategy, env.world_size, env.backend))
    if F.field_size <= 0 or F.feature_size <= 0:
        raise SystemExit("--field_size and --feature_size are required")

    # ------ init envs: file discovery (PS:373-381 / HVD:320-328)
    pipe = F.pipe_mode != 0
    if not pipe:
        tr_files = discover_files(F.training_data_dir, "tr", shuffle=True, seed=F.seed)
        va_files = discover_files(F.val_data_dir, "va")
        te_files = discover_files(F.val_data_dir, "te")
    else:
        tr_files = va_files = te_files = []
    if env.is_chief:
        print("tr_files:", tr_files)
        print("va_files:", va_files)
        print("te_files:", te_files)

    if F.clear_existing_model and env.is_chief and F.model_dir:
        try:
            shutil.rmtree(F.model_dir)
        except Exception as e:  # same tolerance as HVD:334-340
            print(e, "at clear_existing_model")
        else:
            print("existing model cleaned at %s" % F.model_dir)
    barrier(env)

    cfg = build_config(F)
    lr = F.learning_rate * (env.world_size if (F.scale_lr_by_world and F.strategy != "ps") else 1)  # HVD:149
    opt = TF1Optimizer(F.optimizer, lr)
    est = Estimator(cfg, F.batch_size, opt, F.model_dir or None, env, backend=F.backend, seed=F.seed,
                    log_steps=F.log_steps, save_checkpoints_secs=F.save_checkpoints_secs,
                    save_checkpoints_steps=F.save_checkpoints_steps, keep_checkpoint_max=F.keep_checkpoint_max,
                    metrics_file=F.metrics_file or None, comm=F.comm, strategy=F.strategy, ps_staleness=F.ps_staleness)

    num_hosts = len(F.hosts)
    if F.strategy == "ps":
        spec = shard_rules.ps_shard(F.enable_s3_shard, F.hosts, F.current_host)
        if spec is None or num_hosts == 1:  # several worker processes on one box still split the data
            spec = (env.world_size, env.rank)
    else:
        spec = shard_rules.hvd_shard(pipe, F.enable_s3_shard, F.enable_data_multi_path, env.world_size, env.rank,
                                     env.local_rank, F.worker_per_host, num_hosts)
    train_shard = shard_rules.normalise(spec)
    pin = env.device is not None

    def input_fn(files, num_epochs=1, shard=(1, 0), drop_remainder=True, shuffle_seed=0):
        gpu_parse = bool(F.gpu_parse) and drop_remainder and env.device is not None and F.strategy != "ps" \
            and F.backend != "torch"
        nthr = F.loader_threads
        if nthr <= 0:  # auto: the CPUs this rank can count on, between 2 and 8 (the deck's num_parallel_calls advice, PDF p13)
            local_world = max(1, int(os.environ.get("LOCAL_WORLD_SIZE", F.worker_per_host or 1)))
            nthr = min(8, max(2, (os.cpu_count() or 4) // local_world))
        return InputPipeline(files, F.field_size, F.batch_size, num_epochs, shard, drop_remainder, nthr, pin,
                             gpu_parse=gpu_parse, shuffle_seed=shuffle_seed)

    names = channel_names()
    metrics = None
    evaluator = None
    if F.task_type == "train" and F.strategy == "ps" and env.is_chief and F.model_dir and not pipe and va_files:
        # train_and_evaluate (PS:439-442): the chief evaluates every new checkpoint, throttled, while the asynchronous
        # workers (this process included) keep training; a final evaluation follows the last checkpoint
        from .train_and_evaluate import ThrottledCheckpointEvaluator
        evaluator = ThrottledCheckpointEvaluator(
            cfg, F.model_dir, F.optimizer, lambda: input_fn(va_files, 1, (1, 0), False), est.log.log_eval,
            F.eval_start_delay_secs, F.eval_throttle_secs, backend="auto" if F.backend != "torch" else "torch-cpu")
        evaluator.start()
    if F.task_type == "train":
        if not pipe:
            hint = None
            if env.world_size > 1 and F.strategy != "ps":
                # synchronous strategies: every rank must run the same number of steps or one of them ends up waiting
                # in a collective (PDF p22-23); agree on min over ranks of the local full-batch count
                from ..data.pipeline import local_steps
                hint = local_steps(tr_files, F.batch_size, train_shard, 1)
            for epoch in range(F.num_epochs):  # HVD:391-394
                # --perform_shuffle: a fresh record order every epoch (the reference only shuffles the FILE order once,
                # PS:373-381; its flag of this name is never read)
                sseed = (int(F.seed) * 1000003 + epoch + 1) if F.perform_shuffle else 0
                est.train(lambda: input_fn(tr_files, 1, train_shard, True, sseed), max_steps=F.max_steps,
                          steps_hint=hint)
                if env.is_chief and va_files and evaluator is None:  # HVD:393-394: chief evaluates after every epoch
                    metrics = est.evaluate(lambda: input_fn(va_files, 1, (1, 0), False))
                barrier(env)
                if F.max_steps and est.trainer.global_step >= F.max_steps:
                    break
        else:
            # one train() call; epochs are separate FIFOs <channel>_<epoch> consumed in order (HVD:396-405)
            ch = F.training_channel_name or shard_rules.pipe_train_channel(names, env.local_rank) \
                if (F.enable_data_multi_path or len(names) > 2) else (F.training_channel_name or names[-1])
            fifos = [channel_fifo(F.data_base_dir, ch, e) for e in range(F.num_epochs)]
            est.train(lambda: input_fn(fifos, 1, train_shard), max_steps=F.max_steps)
            if env.is_chief:
                ev = F.evaluation_channel_name or shard_rules.pipe_eval_channel(names)
                ev_fifo = channel_fifo(F.data_base_dir, ev, 0)
                if os.path.exists(ev_fifo):
                    metrics = est.evaluate(lambda: input_fn([ev_fifo], 1, (1, 0), False))
            barrier(env)
        if evaluator is not None:  # every worker's pushes are in the last checkpoint: evaluate it, like train_and_evaluate
            res = evaluator.finish(final_eval=True)
            metrics = res[-1] if res else None
    elif F.task_type == "eval":
        # the reference runs this on every rank although only rank 0 has a checkpoint (HVD:407-408); gate it
        if env.is_chief:
            metrics = est.evaluate(lambda: input_fn(va_files, 1, (1, 0), False))
    elif F.task_type == "infer":
        if env.is_chief:
            out_path = os.path.join(F.val_data_dir, "pred.txt")
            with open(out_path, "w") as fo:
                for p in est.predict(lambda: input_fn(te_files, 1, (1, 0), False)):
                    fo.write("%f\n" % p["prob"])
            print("predictions written to %s" % out_path)
    elif F.task_type != "export":
        raise SystemExit("unknown --task_type %r" % F.task_type)

    if F.task_type in ("export", "train") and F.servable_model_dir:  # PS:451-467 / HVD:415-431
        est.export_savedmodel(F.servable_model_dir)
    if metrics is not None and env.is_chief:
        print(json.dumps({"eval": metrics}))
    if hasattr(est.trainer, "close"):
        est.trainer.close()
    barrier(env)
    shutdown()
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
