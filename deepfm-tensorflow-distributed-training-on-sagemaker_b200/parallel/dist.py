"""Process-group plumbing: one process per GPU, rendezvous from the torchrun-style environment
(RANK / LOCAL_RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT). Replaces `hvd.init()` + mpirun (HVD:295,
NB-HVD:87-92) and SageMaker's TF_CONFIG (README-EN:16). NCCL on GPU, gloo on CPU."""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Optional

import torch
import torch.distributed as dist


@dataclass
class DistEnv:
    rank: int = 0
    world_size: int = 1
    local_rank: int = 0
    backend: str = "none"
    device: Optional[torch.device] = None

    @property
    def is_chief(self) -> bool:  # rank 0 alone checkpoints / evaluates / exports (HVD:365-368,393,429)
        return self.rank == 0


def init_from_env(use_cuda: Optional[bool] = None, env=os.environ) -> DistEnv:
    rank = int(env.get("RANK", env.get("OMPI_COMM_WORLD_RANK", "0")))
    world = int(env.get("WORLD_SIZE", env.get("OMPI_COMM_WORLD_SIZE", "1")))
    local = int(env.get("LOCAL_RANK", env.get("OMPI_COMM_WORLD_LOCAL_RANK", "0")))
    if use_cuda is None:
        use_cuda = torch.cuda.is_available()
    dev = None
    if use_cuda:
        torch.cuda.set_device(local % max(torch.cuda.device_count(), 1))  # pin GPU by local rank (HVD:355-357)
        dev = torch.device("cuda", torch.cuda.current_device())
    backend = "none"
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        backend = "nccl" if use_cuda else "gloo"
        if not dist.is_initialized():
            if backend == "nccl":
                dist.init_process_group(backend, rank=rank, world_size=world, device_id=dev)
            else:
                dist.init_process_group(backend, rank=rank, world_size=world)
    return DistEnv(rank, world, local, backend, dev)


def barrier(env: DistEnv):
    if env.world_size > 1 and dist.is_initialized():
        dist.barrier()


def agree_min(env: DistEnv, value: int) -> int:
    """Per-rank step-count agreement: every rank runs the minimum number of steps any rank has data for, which
    removes the uneven-data deadlock Horovod suffers from (PDF p22-23)."""
    if env.world_size <= 1 or not dist.is_initialized():
        return value
    t = torch.tensor([value], dtype=torch.int64, device=env.device if env.backend == "nccl" else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return int(t.item())


def shutdown():
    if dist.is_initialized():
        dist.destroy_process_group()
