"""ORACLE -- TEST / BASELINE INFRASTRUCTURE ONLY.  Never imported by torchkafka_b200/.

Restatement of the reference's *multi-process* path for the timed CPU baseline
(`bench.py --impl reference`): an IterableDataset whose workers each own a
consumer, run `_process` per record, drop `None`, and commit when the main
process signals them after each delivered batch.

    dataset loop + in-loop commit check   /root/reference/src/kafka_dataset.py:147-171
    signal handler sets a flag            /root/reference/src/kafka_dataset.py:107-114
    commit, failures logged               /root/reference/src/kafka_dataset.py:120-145
    worker init (one consumer per worker) /root/reference/src/kafka_dataset.py:208-233
    round-robin SIGUSR1 after each batch  /root/reference/src/auto_commit.py:63-72

Protocol notes from SURVEY.md §4.1-13: a finite stream makes the reference kill
its own workers (handler reset to SIG_DFL on exhaustion, then SIGUSR1), so the
stub consumers here *block* after their last record and the harness stops after a
fixed number of batches -- which is also how a real, never-ending consumer behaves.
"""
from __future__ import annotations

import itertools
import os
import signal

from torch.utils.data import DataLoader, IterableDataset, get_worker_info


class OracleWorkerDataset(IterableDataset):
    def __init__(self, process, consumer_factory):
        self.process = process
        self.consumer_factory = consumer_factory   # worker_id, num_workers -> consumer
        self.consumer = None
        self.worker_id = None
        self.commit_flag = False

    def on_signal(self, signum, frame):
        self.commit_flag = True

    def maybe_commit(self):
        if self.commit_flag:
            try:
                self.consumer.commit()
            finally:
                self.commit_flag = False

    def __iter__(self):
        info = get_worker_info()
        if info is None:
            self.consumer = self.consumer_factory(0, 1)
        else:
            self.worker_id = info.id
            self.consumer = self.consumer_factory(info.id, info.num_workers)
            signal.signal(signal.SIGUSR1, self.on_signal)
        for record in self.consumer:
            sample = self.process(record)
            if sample is not None:
                yield sample
            if self.worker_id is not None:
                self.maybe_commit()


def iterate_with_commits(loader: DataLoader, n_batches: int):
    """auto_commit.py:49-72 for both modes; stops after n_batches (see module docstring)."""
    ds = loader.dataset
    if loader.num_workers == 0:
        it = iter(loader)
        for _ in range(n_batches):
            batch = next(it)
            yield batch
            ds.consumer.commit()
        return
    batches = iter(loader)
    owners = itertools.cycle(batches._workers)  # pylint: disable=protected-access
    for k, (worker, batch) in enumerate(zip(owners, batches)):
        yield batch
        os.kill(worker.pid, signal.SIGUSR1)
        if k + 1 >= n_batches:
            break
    # the workers are blocked inside their consumers: let the DataLoader tear them down
    del batches
