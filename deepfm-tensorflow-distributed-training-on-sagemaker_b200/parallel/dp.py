"""Synchronous data parallelism (reference: Horovod, HVD:149,262,372 — one process per GPU, full replica, gradients
averaged by `hvd.DistributedOptimizer`, LR x world size, state broadcast from rank 0, chief-only checkpoints).

Two implementations behind `Trainer`:

  * FusedDataParallel — the product. The flat fp32 gradient buffer lives in symmetric memory
    (`torch.distributed._symmetric_memory`: one VMM allocation per rank mapped into every peer, plus an NVLS
    multicast mapping when the fabric offers one). ONE kernel per step (csrc/allreduce_opt.cu) then does what
    Horovod + TF need an NCCL all-reduce, a cast/scale pass and 11 ApplyAdam launches for:
        rank r reduces ITS 1/W slice of the gradients across all peers (P2P loads over NVSwitch, or
        `multimem.ld_reduce` in-switch), applies L2 decay + the TF-1 optimizer update on that slice only
        (optimizer state is therefore sharded W ways — ZeRO-1 — cutting the HBM sweep W-fold), and stores the
        updated fp32 weights and bf16 GEMM shadows straight into every peer's replica (P2P stores / `multimem.st`).
    NCCL is not on this path; `comm='nccl'` selects the baseline (NCCL all-reduce + local flat optimizer).
  * install_torch_allreduce — plain `torch.distributed.all_reduce` hook for the torch backend (gloo on CPU).

Because the loss is the mean over the GLOBAL batch and dropout masks are keyed by global row, DP-N on N shards of
a batch equals DP-1 on the concatenated batch (tests/test_dp_cpu.py, tests/test_multi_gpu.py).
"""
from __future__ import annotations

from typing import Dict, Optional

import torch
import torch.distributed as dist


def install_torch_allreduce(model, group=None) -> None:
    """Sum the (already 1/global-batch-scaled) data gradients over ranks with one flat all-reduce."""

    def hook(grads: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        keys = sorted(grads)
        flat = torch.cat([grads[k].reshape(-1) for k in keys])
        dist.all_reduce(flat, group=group)
        out, o = {}, 0
        for k in keys:
            n = grads[k].numel()
            out[k] = flat[o:o + n].view_as(grads[k])
            o += n
        return out

    model.grad_hook = hook


def broadcast_model(model, src: int = 0, group=None) -> None:
    """`BroadcastGlobalVariablesHook(0)` (HVD:372): parameters, optimizer slots and the step from rank `src`."""
    sd = model.state_dict()
    for k in sorted(sd):
        t = sd[k].contiguous()
        dist.broadcast(t, src, group=group)
        sd[k] = t
    model.load_state_dict(sd, strict=False)
