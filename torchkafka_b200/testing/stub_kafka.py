"""In-memory stand-in for the parts of kafka-python that torch-kafka touches.

There is no broker (and no kafka-python wheel) on the build or GPU machines, so
every config in BASELINE.json injects a stub consumer through the documented
``new_consumer`` override (reference ``src/kafka_dataset.py:188-206``,
``README.md:46-57``).  This module provides that stub.  It follows kafka-python's
*documented* behaviour for the calls the dataset makes:

* iterating the consumer yields ``ConsumerRecord`` objects and advances the
  partition position to ``record.offset + 1``;
* ``poll()`` returns ``{TopicPartition: [ConsumerRecord, ...]}`` and advances
  positions the same way;
* ``commit()`` with no argument commits the current position of every partition
  that has one; ``commit(offsets)`` commits the given
  ``{TopicPartition: OffsetAndMetadata}`` map;
* ``close(autocommit=...)``.

Besides that it offers one *additive* fast path, ``poll_packed()``, that hands a
whole run of records out as four flat arrays (values, starts, partitions,
offsets) without creating per-record Python objects.  That is the shape a
C-level fetch-response parser would produce.

``install()`` registers the stub as the ``kafka`` / ``kafka.errors`` modules so
that the unmodified reference (which does ``from kafka import KafkaConsumer`` at
import time, ``src/kafka_dataset.py:21-22``) can be imported by the golden
fixture generator.
"""
from __future__ import annotations

import collections
import sys
import threading
import types
from typing import Dict, Iterable, List, Optional, Sequence

import numpy as np

ConsumerRecord = collections.namedtuple(
    "ConsumerRecord",
    [
        "topic", "partition", "offset", "timestamp", "timestamp_type", "key",
        "value", "headers", "checksum", "serialized_key_size",
        "serialized_value_size", "serialized_header_size",
    ],
)
TopicPartition = collections.namedtuple("TopicPartition", ["topic", "partition"])
OffsetAndMetadata = collections.namedtuple("OffsetAndMetadata", ["offset", "metadata"])

# dtype of one packed record header: 16 bytes, matches tk_rec_t in
# include/torchkafka_b200.h
REC_DTYPE = np.dtype([("start", "<u4"), ("partition", "<i4"), ("offset", "<i8")])


class CommitFailedError(Exception):
    """Same name as kafka.errors.CommitFailedError (reference kafka_dataset.py:22)."""


class PackedRecords:
    """A run of records as flat arrays.

    ``values`` is a uint8 array holding the record values back to back,
    ``starts`` (uint32, n+1 entries) the CSR boundaries inside ``values``,
    ``partitions`` (int32) and ``offsets`` (int64) the Kafka coordinates.
    """

    __slots__ = ("values", "starts", "partitions", "offsets", "topic")

    def __init__(self, values, starts, partitions, offsets, topic="topic"):
        self.topic = topic
        self.values = values
        self.starts = starts
        self.partitions = partitions
        self.offsets = offsets

    def __len__(self):
        return len(self.partitions)

    @property
    def nbytes(self):
        return int(self.starts[-1]) - int(self.starts[0]) if len(self.starts) else 0


class RawFetch:
    """A fetch response as the wire delivers it: per partition one blob of RecordBatch v2 bytes.

    ``blobs`` is a list of ``(partition, uint8 array, fetch_offset, offset_delta)``; the array
    is a view of the broker's log segment, ``fetch_offset`` the consumer's position in that
    partition (records below it are to be skipped), ``offset_delta`` is added to every offset
    (0 on a real broker; the stub uses it to serve the same bytes again at higher offsets).
    """

    __slots__ = ("blobs", "topic")

    def __init__(self, blobs, topic="topic"):
        self.blobs = blobs
        self.topic = topic


class RecordLog:
    """An immutable, ordered log of records shared by stub consumers.

    The order of the log *is* the consumption order.  Records are stored packed
    (one byte buffer + arrays); ``ConsumerRecord`` objects are created lazily,
    once, so that both the reference CPU path and the GPU path read the very
    same bytes.
    """

    def __init__(self, values: Sequence[bytes], partitions: Sequence[int],
                 offsets: Sequence[int], topic: str = "topic"):
        n = len(values)
        assert len(partitions) == n and len(offsets) == n
        self.topic = topic
        lens = np.fromiter((len(v) for v in values), dtype=np.int64, count=n)
        starts = np.zeros(n + 1, dtype=np.int64)
        np.cumsum(lens, out=starts[1:])
        self.starts = starts
        self.values = np.frombuffer(b"".join(values), dtype=np.uint8) if n else np.zeros(0, np.uint8)
        self.partitions = np.asarray(partitions, dtype=np.int32)
        self.offsets = np.asarray(offsets, dtype=np.int64)
        self._byte_values = list(values)
        self._records: Optional[List[ConsumerRecord]] = None
        self._bounds = None

    def __len__(self):
        return len(self.partitions)

    @property
    def nbytes(self):
        return int(self.starts[-1])

    def records(self) -> List[ConsumerRecord]:
        if self._records is None:
            t = self.topic
            self._records = [
                ConsumerRecord(t, int(p), int(o), 0, 0, None, v, [], None, -1, len(v), -1)
                for v, p, o in zip(self._byte_values, self.partitions, self.offsets)
            ]
        return self._records

    def per_partition_count(self) -> np.ndarray:
        if getattr(self, "_ppc", None) is None:
            self._ppc = np.bincount(self.partitions, minlength=int(self.partitions.max()) + 1 if len(self) else 1)
        return self._ppc

    def rep_offsets(self, rep: int) -> np.ndarray:
        """Offsets of the records in repetition ``rep`` of the log (a consumer created with
        ``repeat=R`` serves the log R times; each partition's offsets keep increasing)."""
        if rep == 0:
            return self.offsets
        cache = self.__dict__.setdefault("_rep_offsets", {})
        if rep not in cache:
            cache[rep] = self.offsets + rep * self.per_partition_count()[self.partitions]
        return cache[rep]

    def rep_records(self, rep: int) -> List[ConsumerRecord]:
        if rep == 0:
            return self.records()
        cache = self.__dict__.setdefault("_rep_records", {})
        if rep not in cache:
            t = self.topic
            offs = self.rep_offsets(rep).tolist()
            parts = self.partitions.tolist()
            cache[rep] = [ConsumerRecord(t, p, o, 0, 0, None, v, [], None, -1, len(v), -1)
                          for v, p, o in zip(self._byte_values, parts, offs)]
        return cache[rep]

    def run_bounds(self) -> np.ndarray:
        """Indices i > 0 where partitions[i] != partitions[i-1] (starts of partition runs)."""
        if getattr(self, "_bounds", None) is None:
            p = self.partitions
            self._bounds = (np.flatnonzero(p[1:] != p[:-1]) + 1) if len(p) > 1 else np.zeros(0, np.int64)
        return self._bounds

    def encoded(self, batch_bytes: int = 16384, codec: int = 0):
        """The log as RecordBatch v2 bytes (``testing.record_batch.EncodedLog``), cached.
        ``codec`` 3 = LZ4-compressed batches."""
        cache = self.__dict__.setdefault("_encoded", {})
        if (batch_bytes, codec) not in cache:
            from .record_batch import EncodedLog
            cache[(batch_bytes, codec)] = EncodedLog(self, batch_bytes, codec)
        return cache[(batch_bytes, codec)]

    def select_partitions(self, keep) -> "RecordLog":
        """Sub-log holding only partitions for which ``keep(p)`` is true."""
        idx = [i for i, p in enumerate(self.partitions) if keep(int(p))]
        return RecordLog([self._byte_values[i] for i in idx],
                         self.partitions[idx], self.offsets[idx], self.topic)


class KafkaConsumer:
    """Stub consumer over a :class:`RecordLog`.

    Parameters mirror kafka-python: positional topics, keyword configuration.
    Stub-only keywords: ``log`` (the RecordLog to serve), ``block_at_end`` (after
    the last record, block forever instead of raising StopIteration — what a
    real consumer with ``consumer_timeout_ms=inf`` does),
    ``fail_commits`` (raise CommitFailedError from commit()).
    """

    def __init__(self, *topics, log: Optional[RecordLog] = None, block_at_end: bool = False,
                 fail_commits: bool = False, repeat: int = 1, raw_batch_bytes: int = 16384,
                 raw_codec=0, keep_commit_log: bool = True, **configs):
        self.topics = topics
        self.config = dict(configs)
        # the end of the in-memory log is the end of the stream, unless the stub is told to
        # behave like a live consumer (block_at_end)
        self.consumer_timeout_ms = configs.get("consumer_timeout_ms", float("inf") if block_at_end else 0)
        self._log = log if log is not None else RecordLog([], [], [])
        self._cursor = 0
        self._rep = 0                  # current repetition of the log (see ``repeat``)
        self._repeat = max(1, int(repeat))
        self._block_at_end = block_at_end
        self._raw_batch_bytes = int(raw_batch_bytes)
        self._raw_codec = raw_codec
        self._fail_commits = fail_commits
        self._positions: Dict[int, int] = {}
        self._seek_pending = False
        #: False: only the current table (``committed_table``) and the number of commits are kept
        #: (long benchmark runs commit millions of times)
        self._keep_commit_log = bool(keep_commit_log)
        self.committed_table: Dict[int, int] = {}
        self.n_commits = 0
        self.committed_log: List[Dict[int, int]] = []
        self.commit_calls: List[Optional[int]] = []
        self.closed = False
        self.close_autocommit = None

    # -- iteration (what the reference uses, kafka_dataset.py:156) -----------
    def __iter__(self):
        return self

    def _roll(self) -> bool:
        """Move to the next repetition when the current one is exhausted; False at the very end."""
        if self._seek_pending:
            self._apply_seek()
        if self._cursor < len(self._log):
            return True
        if self._rep + 1 < self._repeat and len(self._log):
            self._rep += 1
            self._cursor = 0
            return True
        return False

    @property
    def consumed(self) -> int:
        """Records handed out so far (all repetitions)."""
        return self._rep * len(self._log) + self._cursor

    def __next__(self):
        if not self._roll():
            if self._block_at_end:
                threading.Event().wait()
            raise StopIteration
        rec = self._log.rep_records(self._rep)[self._cursor]
        self._cursor += 1
        self._positions[rec.partition] = rec.offset + 1
        return rec

    # -- poll (kafka-python API; used by the GPU path) -----------------------
    def poll(self, timeout_ms=0, max_records=None, update_offsets=True):
        """Return the next records grouped per TopicPartition.

        The flattening of the result (dict order, then list order) is always a
        prefix of the log order, so a poll-driven reader sees records in exactly
        the order the iterator would have produced them.
        """
        n = len(self._log)
        if not self._roll():
            return {}
        limit = n - self._cursor if max_records is None else min(max_records, n - self._cursor)
        recs = self._log.rep_records(self._rep)
        bounds = self._log.run_bounds()          # indices where the partition changes
        out: Dict[TopicPartition, List[ConsumerRecord]] = {}
        i = self._cursor
        end = self._cursor + limit
        parts = self._log.partitions
        k = int(np.searchsorted(bounds, i, side="right"))
        while i < end:
            stop = min(int(bounds[k]) if k < len(bounds) else n, end)
            tp = TopicPartition(self._log.topic, int(parts[i]))
            if tp in out:
                break  # would re-open a partition: stop, keep order exact
            out[tp] = recs[i:stop]
            i = stop
            k += 1
        if update_offsets:
            for tp, lst in out.items():
                self._positions[tp.partition] = lst[-1].offset + 1
            self._cursor = i
        return out

    # -- packed fast path (additive) -----------------------------------------
    def poll_packed(self, max_records: int, max_bytes: int) -> Optional[PackedRecords]:
        """Next run of records as flat array views (no per-record objects).

        Returns ``None`` when the log is exhausted.  Positions advance as in
        ``poll``.  At least one record is returned even if it exceeds
        ``max_bytes`` on its own (the caller sizes its slots for that).
        """
        n = len(self._log)
        if not self._roll():
            return None
        a = self._cursor
        starts = self._log.starts
        b = min(n, a + max_records)
        # largest b with starts[b] - starts[a] <= max_bytes
        hi = int(np.searchsorted(starts, starts[a] + max_bytes, side="right")) - 1
        b = max(a + 1, min(b, hi))
        parts = self._log.partitions[a:b]
        offs = self._log.rep_offsets(self._rep)[a:b]
        self._cursor = b
        # positions: last offset per partition in the run
        rev_parts = parts[::-1]
        uniq, first_rev = np.unique(rev_parts, return_index=True)
        for p, j in zip(uniq, first_rev):
            self._positions[int(p)] = int(offs[len(offs) - 1 - j]) + 1
        base = starts[a]
        return PackedRecords(self._log.values[base:starts[b]],
                             (starts[a:b + 1] - base).astype(np.uint32),
                             parts, offs, self._log.topic)

    # -- raw fast path (additive) --------------------------------------------
    def poll_raw(self, max_bytes: int, max_records: int) -> Optional[RawFetch]:
        """Next fetch response as RecordBatch v2 blobs (no parsing on the host).

        Whole batches only, every partition at most once per response, blobs in consumption
        order.  If the cursor stands inside a batch (earlier records of the batch were already
        handed out through ``poll()`` / iteration), the batch is returned whole with
        ``fetch_offset`` = the position, as a broker does.  Returns ``None`` at the end of the log.
        """
        if not self._roll():
            return None
        enc = self._log.encoded(self._raw_batch_bytes, self._raw_codec)
        first, start, part = enc.first_rec, enc.start, enc.partition
        run_end = enc.run_end                    # run_end[b] = first batch after b of another partition
        b = int(np.searchsorted(first, self._cursor, side="right")) - 1      # batch holding the cursor
        offs = self._log.rep_offsets(self._rep)
        delta_of = (self._rep * self._log.per_partition_count()) if self._rep else None
        blobs = []
        seen = set()
        nbytes = nrec = 0
        cursor = self._cursor
        nb = enc.n_batches
        while b < nb:
            p = int(part[b])
            if p in seen:
                break                                    # would re-open a partition
            # whole batches of this partition's run, as many as the limits allow (at least one if
            # the response is still empty)
            e = int(run_end[b])
            byte_room = max_bytes - nbytes
            rec_room = max_records - nrec
            if int(start[e] - start[b]) > byte_room or int(first[e] - first[b]) > rec_room:
                e = min(int(np.searchsorted(start, start[b] + byte_room, side="right")) - 1,
                        int(np.searchsorted(first, first[b] + rec_room, side="right")) - 1, e)
                if e <= b:
                    if blobs:
                        break
                    e = b + 1
            data = enc.data[int(start[b]):int(start[e])]
            delta = int(delta_of[p]) if delta_of is not None else 0
            blobs.append((p, data, int(offs[cursor]), delta))
            seen.add(p)
            nbytes += int(start[e] - start[b])
            nrec += int(first[e] - first[b])
            cursor = int(first[e])
            self._positions[p] = int(offs[cursor - 1]) + 1
            partial = e < int(run_end[b])
            b = e
            if partial or nbytes >= max_bytes or nrec >= max_records:
                break
        self._cursor = cursor
        return RawFetch(blobs, self._log.topic)

    def poll_raw_into(self, buffer, max_records: int):
        """``poll_raw`` that *receives into the caller's buffer* (additive protocol): the next
        fetch response's RecordBatch v2 bytes are written back to back into ``buffer`` (a writable
        uint8 array, e.g. a page-locked staging slot), at most ``len(buffer)`` bytes.  Returns a
        ``RawFetch`` whose blobs are ``(partition, start, length, fetch_offset, offset_delta)`` with
        ``start`` an offset into ``buffer``, or ``None`` at the end of the log.  For a network
        client this is ``recv_into``; the stub copies from its in-memory log segment."""
        fetch = self.poll_raw(len(buffer), max_records)
        if fetch is None:
            return None
        out = np.asarray(buffer)
        pos = 0
        blobs = []
        for p, data, fetch_offset, delta in fetch.blobs:
            n = len(data)
            out[pos:pos + n] = data
            blobs.append((p, pos, n, fetch_offset, delta))
            pos += n
        return RawFetch(blobs, fetch.topic)

    def fetch_arenas(self):
        """Long-lived buffers that every blob returned by ``poll_raw`` is a view of (additive
        protocol): a caller may page-lock them once and DMA out of them.  The stub's arena is the
        encoded log segment."""
        return [self._log.encoded(self._raw_batch_bytes, self._raw_codec).data]

    # -- seek (kafka-python API) ---------------------------------------------
    def seek(self, partition, offset):
        """Next fetch of ``partition`` starts at ``offset``.  The stub serves one global log order,
        so after seeking it resumes at the earliest log entry that some partition still has to
        deliver (exact whenever the sought positions describe a prefix of the log order, which
        is what a pipeline rewinding to its last delivered batch produces)."""
        self._positions[partition.partition] = int(offset)
        self._seek_pending = True

    def _apply_seek(self):
        self._seek_pending = False
        log = self._log
        n = len(log)
        if not n:
            return
        counts = log.per_partition_count()
        by_part = log.__dict__.setdefault("_idx_by_part", {})
        best = None
        for p in np.unique(log.partitions).tolist():
            idx = by_part.get(p)
            if idx is None:
                idx = by_part[p] = np.flatnonzero(log.partitions == p)
            offs = log.offsets[idx]                      # ascending within a partition
            pos = self._positions.get(p, int(offs[0]))
            cnt = int(counts[p])
            rep = max(0, (pos - int(offs[0])) // cnt) if cnt else 0
            k = int(np.searchsorted(offs, pos - rep * cnt, side="left"))
            if k >= len(idx):
                rep, k = rep + 1, 0
            if rep >= self._repeat:
                continue                                 # nothing left in this partition
            g = rep * n + int(idx[k])
            best = g if best is None else min(best, g)
        if best is None:
            self._rep, self._cursor = self._repeat - 1, n
        else:
            self._rep, self._cursor = divmod(best, n)

    # -- offsets -------------------------------------------------------------
    def position(self, tp):
        return self._positions.get(tp.partition)

    def commit(self, offsets=None):
        if self._fail_commits:
            raise CommitFailedError("stub commit failure")
        self.n_commits += 1
        if not self._keep_commit_log:
            if offsets is None:
                self.committed_table.update(self._positions)
            else:
                for tp, om in offsets.items():
                    self.committed_table[tp.partition] = om.offset
            return
        # committed_log[k] = the group's committed-offset table right after commit k
        snap = dict(self.committed_log[-1]) if self.committed_log else {}
        if offsets is None:
            snap.update(self._positions)
        else:
            for tp, om in offsets.items():
                snap[tp.partition] = om.offset
        self.committed_log.append(snap)
        self.committed_table = snap
        self.commit_calls.append(None if offsets is None else len(offsets))

    def committed(self, tp):
        for snap in reversed(self.committed_log):
            if tp.partition in snap:
                return snap[tp.partition]
        return None

    def close(self, autocommit=True):
        self.closed = True
        self.close_autocommit = autocommit


class PollOnlyConsumer(KafkaConsumer):
    """Stub without the packed fast path: only what kafka-python itself offers."""

    @property
    def poll_packed(self):  # hasattr(consumer, "poll_packed") is False
        raise AttributeError("poll_packed")


def install() -> types.ModuleType:
    """Register this stub as ``kafka`` and ``kafka.errors`` in ``sys.modules``.

    Needed only to import the unmodified reference package; the new package
    never requires it.
    """
    mod = types.ModuleType("kafka")
    mod.KafkaConsumer = KafkaConsumer
    mod.TopicPartition = TopicPartition
    mod.OffsetAndMetadata = OffsetAndMetadata
    mod.ConsumerRecord = ConsumerRecord
    err = types.ModuleType("kafka.errors")
    err.CommitFailedError = CommitFailedError
    structs = types.ModuleType("kafka.structs")
    structs.TopicPartition = TopicPartition
    structs.OffsetAndMetadata = OffsetAndMetadata
    mod.errors = err
    mod.structs = structs
    sys.modules["kafka"] = mod
    sys.modules["kafka.errors"] = err
    sys.modules["kafka.structs"] = structs
    return mod
