"""Row sharding across GPUs/ranks.  Articles are independent (decode.py:135-148 decodes one batch = one article
at a time), so the only multi-GPU structure is a partition of the row stream -- the reference's operator
parallelism worker_num (param/HasClusterConfig.java:20-24) -- with NO collective on the data path.
torch.distributed is used only to agree on timings and counts (NCCL on GPUs, gloo in the CPU tests)."""
from __future__ import annotations

from typing import List, Sequence, TypeVar

T = TypeVar("T")


def owner(i: int, world: int) -> int:
    """Subtask (rank, GPU) that decodes the i-th row of the stream: round-robin, like a rebalanced Flink stream.  Rank r
    returns its rows in submission order, so reading the ranks in the same round-robin order restores the stream's."""
    return i % world


def shard_rows(rows: Sequence[T], rank: int, world: int) -> List[T]:
    """Round-robin partition, the way a rebalanced Flink stream deals rows to subtasks."""
    if not 0 <= rank < world:
        raise ValueError("rank %d outside world of %d" % (rank, world))
    return list(rows[rank::world])


def unshard_rows(shards: Sequence[Sequence[T]]) -> List[T]:
    """Inverse of shard_rows over all ranks: restores the original row order."""
    world = len(shards)
    n = sum(len(s) for s in shards)
    out: List[T] = []
    for i in range(n):
        out.append(shards[i % world][i // world])
    return out


def job_throughput(n_local: int, seconds_local: float, dist=None, device=None) -> float:
    """Whole-job units/s: all ranks' units divided by the slowest rank's time (max over ranks)."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return n_local / seconds_local
    import torch
    t = torch.tensor([seconds_local], dtype=torch.float64, device=device)
    n = torch.tensor([float(n_local)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.all_reduce(n, op=dist.ReduceOp.SUM)
    return float(n.item() / t.item())


def wait_for_rank0(dist, key: str, done: bool = False, timeout_s: float = 1800.0) -> None:
    """Ranks other than 0 SLEEP until rank 0 has called this with done=True (a key in the process group's store).

    bench.py's CPU-baseline leg runs on rank 0's host cores only; a NCCL barrier would have the other ranks spin on
    their stream (one busy core each, plus their OpenMP pools) and starve it: on an 8-GPU box the 0.5 s leg took more
    than a minute that way.  No device work and no collective is involved."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return
    import datetime
    from torch.distributed import distributed_c10d
    store = distributed_c10d._get_default_store()
    if dist.get_rank() == 0:
        if done:
            store.set(key, "1")
        return
    store.wait([key], datetime.timedelta(seconds=timeout_s))

