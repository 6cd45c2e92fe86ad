"""``auto_commit``: iterate a DataLoader and commit Kafka offsets after each batch.

Behaviour follows the reference generator (``/root/reference/src/auto_commit.py:22-72``,
facts in SURVEY.md §4.1-6/9):

* not a ``DataLoader``            -> ``TypeError("A DataLoader must be provided.")``
* dataset is not a KafkaDataset   -> batches are passed through untouched
* ``num_workers == 0``            -> yield the batch, then ``dataset.commit()``
  (so commit *k* runs when batch *k+1* is requested; breaking out of the loop
  leaves the last batch uncommitted: at-least-once)
* ``num_workers > 0``             -> the worker that produced a batch is signalled after it
  (``commit_worker``).  Two fixes over the reference: the worker is
  the one the DataLoader says the batch came from (the reference assumes round-robin, which
  breaks once a worker's stream ended), and it commits the offsets of the batches the main
  process has received, not its read-ahead position (``commit_delivered_only``)

New, for datasets that declare a ``payload_schema``: the DataLoader's per-sample
``next()`` loop (``torch/utils/data/_utils/fetch.py:27-45``) is bypassed; the
loader only supplies ``batch_size`` and ``drop_last`` and whole device batches
come from ``KafkaDataset.device_batches``.  The commit after each batch carries
explicit offsets computed on the GPU (K4), identical to what the reference
commits in single-process mode.

Also new, for several ranks that each ingest their own partitions (one process per GPU,
``torchkafka_b200.distributed``): ``auto_commit(dataloader, commit_barrier=True)`` (or a
``torch.distributed`` process group) makes the commit after each batch a *commit barrier*.  In
place of the reference's per-worker ``os.kill(pid, SIGUSR1)`` fan-out (``auto_commit.py:63-72``)
every rank contributes its per-partition commit offsets, all ranks obtain the group's vector
(element-wise maximum, ``dataset.group_offsets``) and then commit their own partitions, so the
committed offsets of the whole group always describe the same training step.  A rank whose
stream ended keeps answering the barrier until every rank has finished.
"""
from __future__ import annotations

import collections
import ctypes
import itertools
import logging
import multiprocessing
import time

from torch.utils.data import DataLoader

from . import kafka_dataset as _kd
from .kafka_dataset import KafkaDataset

#: "no offsets pending": an empty map that is never written to (``_note_offsets`` replaces an empty
#: pending map instead of updating it)
_NOTHING_PENDING: dict = {}


class _CommitBarrier:
    """Per-batch exchange of the ranks' commit vectors (see ``distributed.commit_barrier``)."""

    def __init__(self, dataset, group):
        import torch
        import torch.distributed as dist
        from . import distributed as tkd
        if not (dist.is_available() and dist.is_initialized()):
            raise RuntimeError("commit_barrier needs an initialised torch.distributed process group")
        self.tkd = tkd
        self.group = None if group is True else group
        backend = str(dist.get_backend(self.group))
        # host tensors when the group can move them (gloo), else a small device tensor (NCCL)
        self.device = None if "gloo" in backend or "mpi" in backend else torch.device(
            "cuda", dataset.device_index if dataset.device_index is not None else torch.cuda.current_device())
        self.n = int(dataset.max_partitions)
        # [0, n): commit offset per partition id (-1: nothing to report), [n]: 1 = this rank's stream ended
        self.vec = torch.full((self.n + 1,), -1, dtype=torch.int64)
        self.vec[self.n] = 0
        self.dataset = dataset
        dataset.group_offsets = None
        dataset.commit_barrier_us = collections.deque(maxlen=1 << 16)

    def note(self, offsets):
        """Merge the {TopicPartition: OffsetAndMetadata} of the batch just consumed."""
        for tp, om in offsets.items():
            p = int(tp.partition)
            if not 0 <= p < self.n:
                raise RuntimeError(f"partition {p} outside [0, max_partitions={self.n})")
            self.vec[p] = int(om.offset)

    def exchange(self, done: bool) -> bool:
        """One barrier; returns True once every rank has reported the end of its stream."""
        self.vec[self.n] = 1 if done else 0
        t0 = time.perf_counter()
        local = self.vec.to(self.device) if self.device is not None else self.vec
        merged = self.tkd.commit_barrier_min_max(local, self.n, self.group)
        merged = merged.cpu() if self.device is not None else merged
        self.dataset.commit_barrier_us.append((time.perf_counter() - t0) * 1e6)
        self.dataset.group_offsets = merged[:self.n]
        return bool(merged[self.n] == 1)


def _device_path(dataloader: DataLoader, commit_barrier=None):
    dataset = dataloader.dataset
    if dataloader.num_workers != 0:
        raise RuntimeError(
            "A dataset with a payload_schema runs on the GPU in the main process; "
            "use num_workers=0."
        )
    if dataloader.batch_size is None:
        raise ValueError("auto_commit needs a DataLoader with a batch_size for schema datasets.")
    if commit_barrier is None or commit_barrier is False:
        # Per slot, not per batch: the pipeline hands over the lists of a slot's batches and commit
        # maps; per batch this loop does one yield and one commit (with batch_size=4 the depth of
        # the generator chain is what the host spends its time on).
        # When commit() is the stock method and nothing is pending from an earlier batch, the commit
        # of a batch is issued right here (same steps as KafkaDataset.commit: the consumer lock, the
        # explicit offsets of the batch, failures logged not raised).
        stock = type(dataset).commit is KafkaDataset.commit and not _kd._logger.isEnabledFor(logging.DEBUG)
        failed = _kd._commit_failed_types()
        for views, maps, pipe in dataset._device_slot_batches(dataloader.batch_size, dataloader.drop_last):  # pylint: disable=protected-access
            lock = pipe.consumer_lock
            commit = dataset._consumer.commit                       # pylint: disable=protected-access
            for j, batch in enumerate(views):
                pipe.delivered_index = j
                offsets = maps[j]
                if stock and not dataset._explicit_offsets:         # pylint: disable=protected-access
                    dataset._explicit_offsets = offsets             # (what leaving the loop here leaves pending)
                    yield batch
                    try:
                        with lock:
                            commit(offsets)
                        dataset._explicit_offsets = _NOTHING_PENDING
                    except failed:
                        _kd._logger.error("Commit failed.")
                else:
                    dataset._note_offsets(offsets)                  # pylint: disable=protected-access
                    yield batch
                    dataset.commit()
        return
    barrier = _CommitBarrier(dataset, commit_barrier)
    for batch in dataset.device_batches(dataloader.batch_size, dataloader.drop_last):
        yield batch
        barrier.note(dataset._explicit_offsets)         # pylint: disable=protected-access
        barrier.exchange(False)                         # every rank has consumed this step's batch ...
        dataset.commit()                                # ... then each commits its own partitions
    while not barrier.exchange(True):                   # finished: answer the others' barriers
        pass


def auto_commit(dataloader: DataLoader, commit_barrier=None):
    """Generator over ``dataloader`` that commits each batch once it was consumed.

    ``commit_barrier`` (additive; schema datasets, one rank per GPU): ``True`` for the default
    ``torch.distributed`` group or a process group -- see the module docstring."""
    if not isinstance(dataloader, DataLoader):
        raise TypeError("A DataLoader must be provided.")
    dataset = dataloader.dataset
    if not isinstance(dataset, KafkaDataset):
        yield from dataloader
        return
    if dataset.payload_schema is not None:
        yield from _device_path(dataloader, commit_barrier)
        return
    if commit_barrier:
        raise ValueError("commit_barrier needs a dataset with a payload_schema (one rank per GPU)")
    if dataloader.num_workers == 0:
        for batch in dataloader:
            yield batch
            dataset.commit()
        return
    # Multi-process: the iterator owns the worker processes.  Every worker keeps the positions at
    # the end of each batch it produced; the count of batches received from it lives in shared
    # memory (set up before the workers are forked), the signal only says "look at it".
    delivered = None
    if dataset.commit_delivered_only:
        delivered = getattr(dataset, "_delivered", None)
        if delivered is None or len(delivered) != dataloader.num_workers:
            delivered = multiprocessing.Array(ctypes.c_longlong, dataloader.num_workers, lock=False)
        for w in range(dataloader.num_workers):                     # persistent workers keep the array they
            delivered[w] = 0                                        # were forked with: reuse it, restart the counts
        dataset._delivered = delivered                              # pylint: disable=protected-access
        dataset._delivered_batch_size = dataloader.batch_size or 1  # pylint: disable=protected-access
    batches = iter(dataloader)
    workers = batches._workers  # pylint: disable=protected-access
    # which worker a batch came from: the iterator passes it to _process_data (torch >= 1.8);
    # without that hook fall back to the reference's round-robin assumption
    source = {}
    inner = getattr(batches, "_process_data", None)
    if inner is not None:
        def spy(data, *args, **kwargs):
            source["worker"] = args[0] if args else kwargs.get("worker_idx")
            return inner(data, *args, **kwargs)
        batches._process_data = spy  # pylint: disable=protected-access
    fallback = itertools.cycle(range(len(workers)))
    for batch in batches:
        w = source.pop("worker", None)
        if w is None:
            w = next(fallback)
        yield batch
        if delivered is not None:
            delivered[w] += 1
        dataset.commit_worker(workers[w])
