"""IngestPipeline: Kafka consumer -> pinned CSR slots -> GPU kernels -> batches.

This is the schema path of ``KafkaDataset.__iter__`` (reference
``src/kafka_dataset.py:147-171``) re-shaped for a GPU: instead of one Python
``_process`` call per record, records are polled in bulk, packed into a pinned
staging slot, copied asynchronously (K0) and handled by the kernel sequence of
``SlotEngine.process``.  Slots are software-pipelined: while the GPU works on
slot *s*, the host polls and packs slot *s+1* and reads back the (tiny) result
block of slot *s-1*.

Iteration yields ``(batch, offsets)`` where ``offsets`` is the
``{TopicPartition: OffsetAndMetadata}`` map to commit once the batch has been
consumed -- exactly the consumer positions the reference commits after the same
batch in single-process mode (``src/auto_commit.py:55-58``).
"""
from __future__ import annotations

import collections
import sys
from typing import Dict, Optional

import numpy as np

from .engine import SlotEngine
from .schema import PayloadError


def _offset_struct():
    """OffsetAndMetadata constructor of kafka-python if present, else of the stub."""
    mod = sys.modules.get("kafka.structs")
    cls = getattr(mod, "OffsetAndMetadata", None) if mod is not None else None
    if cls is None:
        try:
            from kafka.structs import OffsetAndMetadata as cls  # type: ignore
        except Exception:
            from .testing.stub_kafka import OffsetAndMetadata as cls
    nfields = len(getattr(cls, "_fields", ("offset", "metadata")))
    if nfields >= 3:        # kafka-python >= 2.1 added leader_epoch
        return lambda off: cls(off, "", -1)
    return lambda off: cls(off, "")


def _tp_struct():
    mod = sys.modules.get("kafka.structs")
    cls = getattr(mod, "TopicPartition", None) if mod is not None else None
    if cls is None:
        try:
            from kafka.structs import TopicPartition as cls  # type: ignore
        except Exception:
            from .testing.stub_kafka import TopicPartition as cls
    return cls


class _Pending:
    __slots__ = ("slot", "n", "out", "partitions", "offsets")

    def __init__(self, slot, n, out, partitions, offsets):
        self.slot = slot
        self.n = n
        self.out = out
        self.partitions = partitions
        self.offsets = offsets


class IngestPipeline:
    def __init__(self, consumer, schema, batch_size: int, drop_last: bool = False,
                 slot_bytes: int = 64 << 20, n_slots: int = 3, device: Optional[int] = None,
                 max_partitions: int = 64, poll_timeout_ms: int = 100, max_empty_polls: int = 1):
        self.consumer = consumer
        self.schema = schema
        self.batch_size = int(batch_size)
        self.drop_last = bool(drop_last)
        self.engine = SlotEngine(schema, batch_size, slot_bytes=slot_bytes, n_slots=n_slots,
                                 device=device, n_partitions=max_partitions)
        self.poll_timeout_ms = poll_timeout_ms
        self.max_empty_polls = max_empty_polls
        self._mk_offset = _offset_struct()
        self._tp_cls = _tp_struct()
        self._tp_of: Dict[int, object] = {}        # dense partition index -> TopicPartition
        self._dense_of: Dict[object, int] = {}     # TopicPartition -> dense index
        self._packed = hasattr(consumer, "poll_packed")
        self._leftover = None                      # records polled but not fitting the last slot
        self._closed = False
        self.stats = collections.Counter()

    # ------------------------------------------------------------------ lifecycle
    def close(self):
        if not self._closed:
            self._closed = True
            self.engine.close()

    # ------------------------------------------------------------------ host side: poll + pack
    def _dense(self, tp) -> int:
        d = self._dense_of.get(tp)
        if d is None:
            d = len(self._dense_of)
            if d >= self.engine.n_partitions:
                raise RuntimeError(
                    f"more than max_partitions={self.engine.n_partitions} partitions assigned; "
                    "raise KafkaDataset.max_partitions")
            self._dense_of[tp] = d
            self._tp_of[d] = tp
        return d

    def _fill_from_packed(self, slot: int):
        eng = self.engine
        run = self.consumer.poll_packed(eng.slot_records, eng.slot_bytes)
        if run is None or len(run) == 0:
            return 0, 0, None, None
        topic = getattr(run, "topic", None) or "topic"
        parts = np.asarray(run.partitions)
        # dense index == raw partition id on this path (single topic)
        for p in np.unique(parts):
            p = int(p)
            if p not in self._tp_of:
                if not 0 <= p < eng.n_partitions:
                    raise RuntimeError(f"partition {p} outside [0, max_partitions={eng.n_partitions})")
                tp = self._tp_cls(topic, p)
                self._tp_of[p] = tp
                self._dense_of[tp] = p
        n, nbytes = eng.fill_packed(slot, run.values, run.starts, parts, run.offsets)
        return n, nbytes, parts, np.asarray(run.offsets)

    def _fill_from_poll(self, slot: int):
        """Flatten ``consumer.poll()`` results (dict order, list order == iterator order)."""
        eng = self.engine
        hb, hr = eng.host_views(slot)
        n = 0
        nbytes = 0
        empty = 0
        parts_acc, offs_acc = [], []
        full = False
        while not full and n < eng.slot_records:
            if self._leftover:
                items, self._leftover = self._leftover, []
            else:
                room = eng.slot_records - n
                polled = self.consumer.poll(timeout_ms=self.poll_timeout_ms,
                                            max_records=min(room, 1 << 16))
                if not polled:
                    empty += 1
                    if n > 0 or empty >= self.max_empty_polls:
                        break
                    continue
                items = list(polled.items())
            for k, (tp, recs) in enumerate(items):
                d = self._dense(tp)
                vals = [r.value for r in recs]
                lens = np.fromiter(map(len, vals), dtype=np.int64, count=len(vals))
                take = len(vals)
                tot = int(lens.sum())
                if n + take > eng.slot_records or nbytes + tot > eng.slot_bytes:
                    # the slot is full inside this partition's run: split it, keep the rest
                    # (records cannot be un-polled) in consumption order for the next slot
                    cum = np.cumsum(lens)
                    take = int(min(np.searchsorted(cum, eng.slot_bytes - nbytes, side="right"),
                                   eng.slot_records - n))
                    if take == 0 and n == 0:
                        raise ValueError(
                            f"a single record of {int(lens[0])} bytes exceeds the slot size")
                    self._leftover = [(tp, recs[take:])] + items[k + 1:]
                    vals, lens, recs = vals[:take], lens[:take], recs[:take]
                    tot = int(lens.sum())
                    full = True
                if take:
                    hb[nbytes:nbytes + tot] = np.frombuffer(b"".join(vals), dtype=np.uint8)
                    starts = np.empty(take, np.int64)
                    starts[0] = nbytes
                    if take > 1:
                        np.cumsum(lens[:-1], out=starts[1:])
                        starts[1:] += nbytes
                    hr["start"][n:n + take] = starts
                    hr["partition"][n:n + take] = d
                    offs = np.fromiter((r.offset for r in recs), dtype=np.int64, count=take)
                    hr["offset"][n:n + take] = offs
                    parts_acc.append(np.full(take, d, np.int32))
                    offs_acc.append(offs)
                    n += take
                    nbytes += tot
                if full:
                    break
        if n == 0:
            return 0, 0, None, None
        hr["start"][n] = nbytes
        return n, nbytes, np.concatenate(parts_acc), np.concatenate(offs_acc)

    # ------------------------------------------------------------------ iteration
    def _error(self, item, res) -> PayloadError:
        err = self.engine.error_for(item.slot, res, item.partitions, item.offsets)
        tp = self._tp_of.get(err.partition)
        if tp is not None:                      # dense index -> the consumer's partition id
            err.partition = getattr(tp, "partition", err.partition)
        return err

    def _offsets_map(self, column: np.ndarray):
        mk = self._mk_offset
        tp_of = self._tp_of
        return {tp_of[int(p)]: mk(int(column[p])) for p in np.flatnonzero(column >= 0)}

    def __iter__(self):
        eng = self.engine
        B = self.batch_size
        fill = self._fill_from_packed if self._packed else self._fill_from_poll
        self._leftover = []
        pending = collections.deque()
        depth = max(1, eng.n_slots - 1)
        prev_slot, prev_out = None, None
        open_lmax = 0                  # longest record among the rows of the open batch
        last = None                    # (_Pending, SlotResult) of the most recent drained slot
        cursor = 0

        def drain(item):
            nonlocal open_lmax, last
            res = eng.result(item.slot)
            self.stats["records"] += res.n_records
            self.stats["kept"] += res.n_kept
            err_rec = res.first_error_rec if res.n_errors else None
            first = B - res.carry_in
            for j in range(res.n_closed):
                if err_rec is not None and res.batch_end[j] >= err_rec:
                    raise self._error(item, res)
                lmax = max(int(res.batch_lmax[j]), open_lmax) if j == 0 else int(res.batch_lmax[j])
                batch = eng.batch_view(item.out, first + j * B, B, lmax)
                self.stats["batches"] += 1
                yield batch, self._offsets_map(res.commit[:, j])
            if err_rec is not None:
                # bad record sits in the still-open batch: it surfaces with that batch
                raise self._error(item, res)
            tail_lmax = int(res.batch_lmax[res.n_closed])
            open_lmax = max(open_lmax, tail_lmax) if res.n_closed == 0 else tail_lmax
            last = (item, res)

        while True:
            slot = cursor % eng.n_slots
            n, nbytes, parts, offs = fill(slot)
            if n == 0:
                break
            eng.submit(slot, n, nbytes)
            out = eng.process(slot, n, prev_slot, prev_out)
            pending.append(_Pending(slot, n, out, parts, offs))
            prev_slot, prev_out = slot, out
            cursor += 1
            self.stats["slots"] += 1
            self.stats["bytes"] += nbytes
            if len(pending) >= depth:
                yield from drain(pending.popleft())
        while pending:
            yield from drain(pending.popleft())
        # stream exhausted: the open batch becomes the short tail batch (fetch.py:39-42)
        if last is not None and not self.drop_last:
            item, res = last
            if res.carry_out:
                first = B + res.n_kept - res.carry_out
                batch = eng.batch_view(item.out, first, res.carry_out, open_lmax)
                self.stats["batches"] += 1
                yield batch, self._offsets_map(res.commit[:, res.n_closed])
