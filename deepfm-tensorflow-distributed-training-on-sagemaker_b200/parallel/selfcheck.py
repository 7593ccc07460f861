"""Multi-GPU self-checks that run in any job (bench.py prints them, tests/test_multi_gpu.py and scripts/dp_check.py assert
them): SURVEY section 4 "allreduce kernel == torch.distributed.all_reduce" and "replicas stay identical".

Nothing here is on a timed path. Every function is collective: all ranks must call it with the same arguments.
"""
from __future__ import annotations

from typing import Dict

import torch
import torch.distributed as dist


def replicas_equal(model, group=None) -> Dict[str, object]:
    """Bit-equality of the parameter buffer, the bf16 GEMM shadows and (where replicated) the optimizer slots across
    ranks: every rank compares its buffers with rank 0's. Returns {"equal": bool, "first_bad_rank": int | None, ...}."""
    bufs = {"P": model.P, "shadow": model.shadow}
    if getattr(model, "slots_replicated", False):  # sparse dp: slots are replicated too (ZeRO-1 dp shards them)
        for i, s in enumerate((model.S0, model.S1)):
            if s is not None:
                bufs["S%d" % i] = s
    bad = torch.zeros(1, device=model.dev, dtype=torch.int32)
    detail = {}
    for name, t in bufs.items():
        ref = t.clone()
        dist.broadcast(ref, 0, group=group)
        if t.dtype == torch.bfloat16:  # compare bit patterns (NaN-safe, and -0 != +0)
            same = bool(torch.equal(ref.view(torch.int16), t.view(torch.int16)))
        else:
            same = bool(torch.equal(ref.view(torch.int32), t.view(torch.int32)))
        detail[name] = same
        if not same:
            bad.fill_(1)
    flags = [torch.zeros_like(bad) for _ in range(dist.get_world_size(group))]
    dist.all_gather(flags, bad, group=group)
    bad_ranks = [r for r, f in enumerate(flags) if int(f) != 0]
    return {"equal": not bad_ranks, "first_bad_rank": bad_ranks[0] if bad_ranks else None, "local": detail}


def exchange_equals_allreduce(cfg, batch_size: int, device, world: int, rank: int, comm: str = "auto",
                              exchange: str = "sparse", seed: int = 123, group=None) -> Dict[str, object]:
    """Fused exchange kernel vs `dist.all_reduce` on a copy of the same gradients.

    A throw-away replica with GD, lr = 1, no L2 and ZERO parameters takes one optimizer pass over hand-made gradients
    (random, different per rank, non-zero only where a real backward pass writes: the dense blocks and the rows this
    rank's batch looks up). With w = 0 the update is w' = -sum_r g_r exactly, so -w' IS the kernel's reduction and can be
    compared element by element with NCCL's sum of the same buffers."""
    from ..optim.tf1 import TF1Optimizer
    from dataclasses import replace
    from ..data.synthetic import SyntheticCTR
    cfg0 = replace(cfg, l2_reg=0.0)
    if exchange == "sparse":
        from .dp_sparse import SparseDataParallel as DP
    else:
        from .dp import FusedDataParallel as DP
    m = DP(cfg0, batch_size, TF1Optimizer("GD", 1.0), seed, device, world, rank, comm=comm, use_graph=False,
           process_group=group)
    L = m.layout
    lt = L.trainable_numel
    ids, vals, labels = SyntheticCTR(cfg.field_size, cfg.feature_size, seed=1000 + rank).batch(batch_size, 7)
    m.load_inputs(ids.to(device), vals.to(device), labels.to(device), 0)
    gen = torch.Generator(device=device).manual_seed(seed * 100 + rank)
    g = torch.zeros(lt, device=device)
    dense = torch.ones(lt, dtype=torch.bool, device=device)
    if L.include_embeddings:
        sw, sv = L.segments["fm_w"], L.segments["fm_v"]
        dense[sw.offset:sv.offset + sv.numel] = False
        rows = torch.unique(ids.to(device).long().reshape(-1))
        K_ = cfg.embedding_size
        g[sw.offset + rows] = torch.randn(rows.numel(), device=device, generator=gen)
        gv = g[sv.offset:sv.offset + sv.numel].view(-1, K_)
        gv[rows] = torch.randn(rows.numel(), K_, device=device, generator=gen)
    nd = int(dense.sum())
    g[dense] = torch.randn(nd, device=device, generator=gen)
    for name, seg in L.segments.items():  # keep the alignment padding at zero like a real gradient buffer
        if seg.trainable:
            end = seg.offset + seg.numel
            nxt = min([s.offset for s in L.segments.values() if s.offset > seg.offset] + [lt])
            g[end:nxt] = 0
    ref = g.clone()
    torch.cuda.synchronize(device)
    dist.barrier(group=group)
    dist.all_reduce(ref, group=group)
    m.P.zero_()
    m.G.zero_()
    m.G[:lt].copy_(g)
    if getattr(m, "_routed", False):
        # owner routing: in a real step emb_bwd has already summed the table gradients into their owner's buffer and left
        # nothing elsewhere — put the buffer into that state (the routed emb_bwd itself is checked against the oracle by
        # scripts/dp_check.py); the dense ranges still go through the kernel's own reduction
        a, b = m._local_range
        lo, hi = m.slice
        m.G[a:b].zero_()
        oa, ob = max(a, lo), min(b, hi)
        if ob > oa:
            m.G[oa:ob].copy_(ref[oa:ob])
    torch.cuda.synchronize(device)
    dist.barrier(group=group)
    cur = torch.cuda.current_stream()
    if getattr(m, "split", None) is not None:
        m._split_side_kernels(m.in_slots[0])
    elif hasattr(m, "_pre_step"):
        pass  # dense path: nothing to prepare (gradients are in place)
    getattr(m, "_exchange_all", m._optimizer_kernels)()  # every bucket of the exchange
    cur.synchronize()
    torch.cuda.synchronize(device)
    dist.barrier(group=group)
    got = -m.P[:lt]
    scale = float(ref.abs().max()) + 1e-12
    err = float((got - ref).abs().max()) / scale
    zeroed = bool(int(m.G[:lt].count_nonzero()) == 0) if exchange == "sparse" else None
    t = torch.tensor([err], device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    out = {"exchange": exchange, "comm": m.comm_name, "max_rel_err_vs_nccl_allreduce": float(t), "elements": lt,
           "grads_zeroed": zeroed}
    del m
    torch.cuda.synchronize(device)
    dist.barrier(group=group)
    return out
