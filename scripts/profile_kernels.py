#!/usr/bin/env python
"""Launch every kernel of the SURVEY section 8 worklist that is not part of the headline dp step, once, at production
shapes, inside a cudaProfilerStart/Stop window, so ONE single-GPU `ncu --profile-from-start off` call captures them:

    python scripts/profile_kernels.py                                   # plain run first (must exit 0)
    ncu --set full --clock-control none --import-source on --profile-from-start off \
        -o gpurun_out/prof_kernels python scripts/profile_kernels.py

Sections (each: 3 warm-up steps outside the window, 1 step inside, eager launches so each kernel is a separate result):
  sharded-pull   emb_fm_fwd_a2a, emb_bwd_a2a, sparse_row_opt        (4 Mi x 64 table, world 1 = self-mapped symmetric memory)
  sharded-push   a2a_dispatch / serve / combine / grad_push / grad_apply
  batch-norm     bn_fwd / bn_bwd (+ fp32-output GEMM epilogue)
  eval           auc_hist
  parse          example_parse on 1024 serialized Examples
  ps             ps_gather (zero-copy rows from the mapped host store)
The multi-GPU kernels (allreduce_opt, bcast) cannot run under ncu (one GPU only): their numbers are device-timed in
scripts/ar_bench.py / ar_sweep.py.
"""
from __future__ import annotations

import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from deepfm_b200.data.synthetic import SyntheticCTR
from deepfm_b200.models import DeepFMConfig
from deepfm_b200.optim import TF1Optimizer
from deepfm_b200.runtime.trainer import Trainer

B, F = 1024, 39
dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
prof = torch.cuda.profiler


def window(fn):
    torch.cuda.synchronize()
    prof.start()
    fn()
    torch.cuda.synchronize()
    prof.stop()


def train_section(name, cfg, strategy, env=None):
    for k, v in (env or {}).items():
        os.environ[k] = v
    tr = Trainer(cfg, B, TF1Optimizer("Adam", 5e-4), seed=1, device=dev, backend="fused", world_size=1, rank=0,
                 use_graph=False, strategy=strategy)
    gen = SyntheticCTR(F, cfg.feature_size, seed=0)
    batches = [tuple(t.to(dev) for t in gen.batch(B, i)) for i in range(4)]
    for i in range(3):
        tr.train_batch_device(*batches[i])
    window(lambda: tr.train_batch_device(*batches[3]))
    print("%-14s ok  loss=%.4f  kernels/step=%d  comm=%s" % (name, float(tr.last_loss()), tr.kernels_per_step, tr.comm_name),
          flush=True)
    for k in (env or {}):
        os.environ.pop(k, None)
    return tr, gen


def main():
    big = DeepFMConfig(field_size=F, feature_size=4 * 1024 * 1024, embedding_size=64, deep_layers="400,400,400",
                       dropout=[0.5] * 3, l2_reg=1e-4, l2_mode="touched")
    tr, _ = train_section("sharded-pull", big, "sharded", {"DFM_A2A": "pull"})
    del tr
    torch.cuda.empty_cache()
    tr, _ = train_section("sharded-push", big, "sharded", {"DFM_A2A": "push"})
    del tr
    torch.cuda.empty_cache()

    bn = DeepFMConfig(field_size=F, feature_size=117581, embedding_size=32, deep_layers="400,400,400",
                      dropout=[0.5] * 3, l2_reg=1e-4, batch_norm=True)
    tr, gen = train_section("batch-norm", bn, "dp")
    ev = [gen.batch(B, 100 + i, ids_dtype=torch.int64) for i in range(2)]
    tr.evaluate(ev[:1])
    window(lambda: tr.evaluate(ev[1:]))
    print("eval           ok", flush=True)
    del tr

    # ---- example_parse on real loader slots
    from deepfm_b200.data import native
    from deepfm_b200.ops import kernels as K
    from deepfm_b200.utils import tfrecord as T
    rng = np.random.default_rng(0)
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "tr.tfrecords")
        with native.NativeRecordWriter(path) as w:
            for i in range(B):
                w.write(T.make_ctr_example(float(i % 2), rng.integers(0, 117581, F).tolist(),
                                           rng.random(F).astype(np.float32).tolist()))
        ld = native.NativeRawLoader([path], B, pin_memory=True)
        slot, rows, raw = ld.next()
        draw = raw.cuda()
        ld.release(slot)
        ld.close()
    off = ((B + 1) * 4 + 255) & ~255
    labels = torch.empty(B, device=dev)
    ids = torch.empty(B, F, device=dev, dtype=torch.int32)
    vals = torch.empty(B, F, device=dev)
    err = torch.zeros(1, device=dev, dtype=torch.int32)
    parse = lambda: K.example_parse(draw[off:], draw[: (B + 1) * 4].view(torch.int32), B, F, labels, ids, vals, err)
    parse()
    window(parse)
    assert int(err) == 0
    print("parse          ok  (%d record bytes)" % (draw.numel() - off), flush=True)

    # ---- ps_gather from a mapped host store
    from deepfm_b200.parallel.dist import DistEnv
    from deepfm_b200.parallel.ps import AsyncPSTrainer
    cfg = DeepFMConfig(field_size=F, feature_size=117581, embedding_size=32, deep_layers="400,400,400",
                       dropout=[0.5] * 3, l2_reg=1e-4, l2_mode="touched")
    ps = AsyncPSTrainer(cfg, B, TF1Optimizer("Adam", 5e-4), 0, DistEnv(0, 1, 0, "none", dev), backend="fused", staleness=2)
    gen = SyntheticCTR(F, 117581, seed=0)
    for i in range(3):
        ps.train_batch(*gen.batch(B, i))
    ps.drain()
    window(lambda: (ps.train_batch(*gen.batch(B, 3)), ps.drain()))
    print("ps             ok  kernels/step=%d" % ps.kernels_per_step, flush=True)
    ps.close()
    if torch.distributed.is_initialized():
        torch.distributed.destroy_process_group()
    print("PROFILE_KERNELS DONE", flush=True)


if __name__ == "__main__":
    main()
