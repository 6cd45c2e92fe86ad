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
import contextlib
import queue
import sys
import threading
from typing import Dict, Optional

import numpy as np

from . import _tkpack      # CPython helper built next to libtorchkafka_b200.so (csrc/tk_pymod.cpp)
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
                 max_partitions: int = 64, poll_timeout_ms: int = 100, max_empty_polls: int = 1,
                 threaded: bool = True):
        self.consumer = consumer
        self.schema = schema
        self.batch_size = int(batch_size)
        self.drop_last = bool(drop_last)
        self.engine = SlotEngine(schema, batch_size, slot_bytes=slot_bytes, n_slots=max(n_slots, 2),
                                 device=device, n_partitions=max_partitions)
        self.poll_timeout_ms = poll_timeout_ms
        self.max_empty_polls = max_empty_polls
        self.threaded = bool(threaded)
        #: serialises consumer calls: the producer thread polls, the user's thread commits,
        #: and kafka-python consumers are not safe for concurrent calls
        self.consumer_lock = threading.Lock()
        self._mk_offset = _offset_struct()
        self._tp_cls = _tp_struct()
        self._tp_of: Dict[int, object] = {}        # dense partition index -> TopicPartition
        self._dense_of: Dict[object, int] = {}     # TopicPartition -> dense index
        self._packed = hasattr(consumer, "poll_packed")
        self._leftover = []                        # (tp, records) polled but not fitting the last slot
        self._field_idx = {}                       # record type -> (value index, offset index)
        cap = self.engine.slot_records
        self._ptrs = np.zeros(cap, np.uint64)
        self._lens = np.zeros(cap, np.uint32)
        self._offs = np.zeros(cap, np.int64)
        self._parts = np.zeros(cap, np.int32)
        self._closed = False
        self._stop = threading.Event()
        self._thread = None
        self._free_q = None
        self.stats = collections.Counter()

    # ------------------------------------------------------------------ lifecycle
    def close(self):
        if self._closed:
            return
        self._closed = True
        self._stop.set()
        if self._thread is not None:
            try:
                self._free_q.put_nowait(None)
            except Exception:
                pass
            self._thread.join(timeout=30)
            self._thread = None
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
        with self.consumer_lock:
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

    def _indices(self, rec):
        """(value index, offset index) for tuple-like records, (-1, -1) for attribute access."""
        t = type(rec)
        idx = self._field_idx.get(t)
        if idx is None:
            fields = getattr(t, "_fields", None)
            if isinstance(rec, tuple) and fields and "value" in fields and "offset" in fields:
                idx = (fields.index("value"), fields.index("offset"))
            else:
                idx = (-1, -1)
            self._field_idx[t] = idx
        return idx

    def _fill_from_poll(self, slot: int):
        """Flatten ``consumer.poll()`` results (dict order, list order == iterator order).

        Per record only ~25 ns of C run under the GIL (``_tkpack.harvest``: address, length
        and offset of each value); the payload bytes are then gathered into pinned memory by
        ``tk_slot_pack`` on several threads with the GIL released.
        """
        eng = self.engine
        ptrs, lens, offs, parts = self._ptrs, self._lens, self._offs, self._parts
        n = 0
        nbytes = 0
        empty = 0
        hold = []                       # keeps the record objects (and their bytes) alive until packed
        full = False
        while not full and n < eng.slot_records:
            if self._leftover:
                items, self._leftover = self._leftover, []
            else:
                with self.consumer_lock:
                    polled = self.consumer.poll(timeout_ms=self.poll_timeout_ms,
                                                max_records=min(eng.slot_records - n, 1 << 16))
                if not polled:
                    empty += 1
                    if n > 0 or empty >= self.max_empty_polls or self._stop.is_set():
                        break
                    continue
                items = list(polled.items())
            for k, (tp, recs) in enumerate(items):
                if not recs:
                    continue
                d = self._dense(tp)
                vi, oi = self._indices(recs[0])
                taken, tot, keep = _tkpack.harvest(
                    recs, vi, oi, 0, eng.slot_records - n, eng.slot_bytes - nbytes,
                    ptrs.ctypes.data + 8 * n, lens.ctypes.data + 4 * n, offs.ctypes.data + 8 * n)
                hold.append(recs)
                if keep is not None:
                    hold.append(keep)
                parts[n:n + taken] = d
                n += taken
                nbytes += tot
                if taken < len(recs):
                    # the slot is full inside this partition's run: records cannot be un-polled,
                    # so the rest waits, in consumption order, for the next slot
                    if taken == 0 and n == 0:
                        raise ValueError("a single record exceeds the slot size "
                                         f"({eng.slot_bytes} bytes); raise device_slot_bytes")
                    self._leftover = [(tp, recs[taken:])] + items[k + 1:]
                    full = True
                    break
        if n == 0:
            return 0, 0, None, None
        after = eng.pack_pointers(slot, ptrs, lens, parts, offs, n)
        assert after == nbytes
        del hold
        return n, nbytes, parts[:n].copy(), offs[:n].copy()

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

    def _produce(self, free_q, filled_q):
        """Producer thread: poll + pack + submit (H2D) slot after slot."""
        fill = self._fill_from_packed if self._packed else self._fill_from_poll
        try:
            while not self._stop.is_set():
                slot = free_q.get()
                if slot is None or self._stop.is_set():
                    break
                n, nbytes, parts, offs = fill(slot)
                if n == 0:
                    break
                self.engine.submit(slot, n, nbytes)
                filled_q.put((slot, n, nbytes, parts, offs))
            filled_q.put(None)
        except BaseException as exc:  # forwarded to the consuming thread
            filled_q.put(exc)

    def _filled_slots(self):
        """Generator of (slot, n, nbytes, partitions, offsets) in consumption order, plus the
        callable that hands a drained slot back."""
        eng = self.engine
        if not self.threaded:
            fill = self._fill_from_packed if self._packed else self._fill_from_poll
            free = collections.deque(range(eng.n_slots))

            def gen():
                while True:
                    slot = free.popleft()       # never empty: pending depth < n_slots
                    n, nbytes, parts, offs = fill(slot)
                    if n == 0:
                        return
                    eng.submit(slot, n, nbytes)
                    yield slot, n, nbytes, parts, offs

            return gen(), free.append
        free_q = queue.Queue()
        filled_q = queue.Queue()
        for s in range(eng.n_slots):
            free_q.put(s)
        self._free_q = free_q
        self._thread = threading.Thread(target=self._produce, args=(free_q, filled_q),
                                        name="torchkafka-producer", daemon=True)
        self._thread.start()

        def gen():
            while True:
                item = filled_q.get()
                if item is None:
                    return
                if isinstance(item, BaseException):
                    raise item
                yield item

        return gen(), free_q.put

    def __iter__(self):
        eng = self.engine
        B = self.batch_size
        self._leftover = []
        pending = collections.deque()
        depth = max(1, eng.n_slots - 2) if self.threaded else max(1, eng.n_slots - 1)
        prev_slot, prev_out = None, None
        open_lmax = 0                  # longest record among the rows of the open batch
        last = None                    # (_Pending, SlotResult) of the most recent drained slot
        slots, release = self._filled_slots()

        def drain(item):
            nonlocal open_lmax, last
            res = eng.result(item.slot)
            release(item.slot)         # pinned memory is reusable: H2D and kernels are done
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

        for slot, n, nbytes, parts, offs in slots:
            out = eng.process(slot, n, prev_slot, prev_out)
            pending.append(_Pending(slot, n, out, parts, offs))
            prev_slot, prev_out = slot, out
            self.stats["slots"] += 1
            self.stats["bytes"] += nbytes
            if len(pending) > depth:
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
