"""``auto_commit``: iterate a DataLoader and commit Kafka offsets after each batch.

Behaviour follows the reference generator (``/root/reference/src/auto_commit.py:22-72``,
facts in SURVEY.md §4.1-6/9):

* not a ``DataLoader``            -> ``TypeError("A DataLoader must be provided.")``
* dataset is not a KafkaDataset   -> batches are passed through untouched
* ``num_workers == 0``            -> yield the batch, then ``dataset.commit()``
  (so commit *k* runs when batch *k+1* is requested; breaking out of the loop
  leaves the last batch uncommitted: at-least-once)
* ``num_workers > 0``             -> workers are signalled round-robin after
  each batch (``commit_worker``)

New, for datasets that declare a ``payload_schema``: the DataLoader's per-sample
``next()`` loop (``torch/utils/data/_utils/fetch.py:27-45``) is bypassed; the
loader only supplies ``batch_size`` and ``drop_last`` and whole device batches
come from ``KafkaDataset.device_batches``.  The commit after each batch carries
explicit offsets computed on the GPU (K4), identical to what the reference
commits in single-process mode.
"""
from __future__ import annotations

import itertools

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
    # Multi-process: the iterator owns the worker processes; tasks are handed
    # out round-robin, so batch i came from worker i % num_workers (holds while
    # all workers are alive and in_order=True).
    batches = iter(dataloader)
    owners = itertools.cycle(batches._workers)  # pylint: disable=protected-access
    for worker, batch in zip(owners, batches):
        yield batch
        dataset.commit_worker(worker)
