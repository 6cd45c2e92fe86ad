"""Multi-GPU sharding and the commit barrier.

The ingest path shards by Kafka partition: rank r of W owns the partitions
``{p : p % W == r}`` -- what the reference obtains from the broker's consumer-group
assignment with one consumer per worker (``src/kafka_dataset.py:219-231``,
``README.md:122-126``).  Each rank has its own consumer, staging slots, streams and kernels;
there is no collective on the data path.  The only exchange is the *commit barrier*: every
rank contributes its per-partition commit offsets (``-1`` for partitions it does not own) and
all ranks obtain the element-wise maximum, i.e. the group's global committed-offset vector.
That is P int64 values per rank (4 KiB for 8 ranks x 64 partitions): latency-bound, done with
``torch.distributed.all_gather`` on host tensors (gloo) or on a tiny device tensor (NCCL).
"""
from __future__ import annotations

from typing import Dict, Iterable, List, Optional

import torch
import torch.distributed as dist


def owned_partitions(n_partitions: int, rank: int, world_size: int) -> List[int]:
    """Partitions of rank ``rank``: p % world_size == rank."""
    return [p for p in range(n_partitions) if p % world_size == rank]


def owns(partition: int, rank: int, world_size: int) -> bool:
    return partition % world_size == rank


def offsets_vector(offsets: Dict[int, int], n_partitions: int, device=None) -> torch.Tensor:
    """{partition: offset} -> int64[n_partitions], -1 where the rank has nothing to report."""
    vec = torch.full((n_partitions,), -1, dtype=torch.int64)
    for p, o in offsets.items():
        vec[int(p)] = int(o)
    return vec.to(device) if device is not None else vec


def commit_barrier(local: torch.Tensor, group: Optional[dist.ProcessGroup] = None) -> torch.Tensor:
    """All ranks exchange their commit vectors; returns the element-wise maximum.

    ``local`` is int64[P] (CPU tensor with the gloo backend, CUDA tensor with NCCL).  Without an
    initialised process group (single GPU) the input is returned unchanged.
    """
    if not (dist.is_available() and dist.is_initialized()):
        return local.clone()
    world = dist.get_world_size(group)
    if world == 1:
        return local.clone()
    gathered = [torch.empty_like(local) for _ in range(world)]
    dist.all_gather(gathered, local, group=group)
    return torch.stack(gathered, 0).max(dim=0).values


def commit_barrier_min_max(local: torch.Tensor, n_max: int, group: Optional[dist.ProcessGroup] = None) -> torch.Tensor:
    """``commit_barrier`` for a vector whose first ``n_max`` entries are commit offsets (maximum
    over the ranks) and whose remaining entries are flags that only count when *every* rank set
    them (minimum over the ranks): what ``auto_commit(..., commit_barrier=...)`` exchanges per batch
    (offsets + "my stream has ended")."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return local.clone()
    world = dist.get_world_size(group)
    gathered = [torch.empty_like(local) for _ in range(world)]
    dist.all_gather(gathered, local, group=group)
    stacked = torch.stack(gathered, 0)
    return torch.cat([stacked[:, :n_max].max(dim=0).values, stacked[:, n_max:].min(dim=0).values])


def global_commit(offsets: Dict[int, int], n_partitions: int, device=None,
                  group: Optional[dist.ProcessGroup] = None) -> Dict[int, int]:
    """Convenience: dict in, dict out (partitions never seen by any rank are omitted)."""
    vec = commit_barrier(offsets_vector(offsets, n_partitions, device), group).cpu()
    return {p: int(o) for p, o in enumerate(vec.tolist()) if o >= 0}
