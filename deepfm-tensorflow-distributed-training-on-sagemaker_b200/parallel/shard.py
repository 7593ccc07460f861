"""The dataset shard decision, as a pure function ('shard decision table as a pure function').

Horovod trainer (HVD:92-120, table at README-EN:86-91) — returns (num_shards, index) or None for "no shard":

  file mode:  enable_s3_shard  -> shard(worker_per_host, local_rank)     data already split per instance
              otherwise        -> shard(world_size, rank)                every worker sees every file
  pipe mode:  enable_data_multi_path (each local worker has its own channel)
                  enable_s3_shard   -> no shard
                  otherwise         -> shard(num_hosts, rank // worker_per_host) if num_hosts > 1 else no shard
              single shared channel
                  enable_s3_shard   -> shard(worker_per_host, local_rank)
                  otherwise         -> shard(world_size, rank)

Parameter-server trainer (PS:114-117): shard by HOST unless S3 already sharded the data.
"""
from __future__ import annotations

from typing import List, Optional, Tuple


def hvd_shard(pipe_mode: bool, enable_s3_shard: bool, enable_data_multi_path: bool, world_size: int, rank: int,
              local_rank: int, worker_per_host: int, num_hosts: int) -> Optional[Tuple[int, int]]:
    if not pipe_mode:
        if enable_s3_shard:
            return (worker_per_host, local_rank)
        return (world_size, rank)
    if enable_data_multi_path:
        if not enable_s3_shard and num_hosts > 1:
            return (num_hosts, rank // max(worker_per_host, 1))
        return None
    if enable_s3_shard:
        return (worker_per_host, local_rank)
    return (world_size, rank)


def ps_shard(enable_s3_shard: bool, hosts: List[str], current_host: str) -> Optional[Tuple[int, int]]:
    if enable_s3_shard:
        return None
    return (len(hosts), hosts.index(current_host))


def normalise(spec: Optional[Tuple[int, int]]) -> Tuple[int, int]:
    if spec is None or spec[0] <= 1:
        return (1, 0)
    n, i = spec
    if not 0 <= i < n:
        raise ValueError("bad shard spec %r" % (spec,))
    return (n, i)


def pipe_train_channel(channel_names: List[str], local_rank: int) -> str:
    """HVD:376-380,403: SM_CHANNELS is sorted alphabetically, so 'evaluation' comes first and the training channels
    follow; local worker r reads channel_names[1 + r]."""
    return channel_names[1 + local_rank]


def pipe_eval_channel(channel_names: List[str]) -> str:
    return channel_names[0]
