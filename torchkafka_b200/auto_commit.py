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
"""
from __future__ import annotations

import ctypes
import itertools
import multiprocessing

from torch.utils.data import DataLoader

from .kafka_dataset import KafkaDataset


def _device_path(dataloader: DataLoader):
    dataset = dataloader.dataset
    if dataloader.num_workers != 0:
        raise RuntimeError(
            "A dataset with a payload_schema runs on the GPU in the main process; "
            "use num_workers=0."
        )
    if dataloader.batch_size is None:
        raise ValueError("auto_commit needs a DataLoader with a batch_size for schema datasets.")
    for batch in dataset.device_batches(dataloader.batch_size, dataloader.drop_last):
        yield batch
        dataset.commit()


def auto_commit(dataloader: DataLoader):
    """Generator over ``dataloader`` that commits each batch once it was consumed."""
    if not isinstance(dataloader, DataLoader):
        raise TypeError("A DataLoader must be provided.")
    dataset = dataloader.dataset
    if not isinstance(dataset, KafkaDataset):
        yield from dataloader
        return
    if dataset.payload_schema is not None:
        yield from _device_path(dataloader)
        return
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
