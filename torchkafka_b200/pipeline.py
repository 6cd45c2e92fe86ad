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
import ctypes as _ctypes
import queue
import sys
import threading
import time
from typing import Dict, Optional

import numpy as np

from . import _tkpack      # CPython helper built next to libtorchkafka_b200.so (csrc/tk_pymod.cpp)
from . import affinity as _affinity
from . import engine as _engine
from .schema import PayloadError


from .structs import _offset_make, _offset_struct, _tp_struct  # noqa: E402  (kafka-python's structs, or the stub's)


try:  # NVTX ranges around the per-slot stages (visible in nsys / ncu timelines; ~0.1 us each otherwise)
    from torch.cuda import nvtx as _nvtx
    _nvtx.range_push, _nvtx.range_pop        # noqa: B018  (torch builds without NVTX lack them)
except Exception:  # pragma: no cover
    _nvtx = None


@contextlib.contextmanager
def _range(name: str):
    if _nvtx is None:
        yield
        return
    _nvtx.range_push(name)
    try:
        yield
    finally:
        _nvtx.range_pop()


class _Pending:
    __slots__ = ("slot", "n", "out", "partitions", "offsets")

    def __init__(self, slot, n, out, partitions, offsets):
        self.slot = slot
        self.n = n
        self.out = out
        self.partitions = partitions     # host copy of the record coordinates; raw slots: the
        self.offsets = offsets           # fetch's {dense partition: fetch offset} and None


#: staging slots of a pipeline fed with record batches (see IngestPipeline.__init__)
RAW_MIN_SLOTS = 6


class IngestPipeline:
    def __init__(self, consumer, schema, batch_size: int, drop_last: bool = False,
                 slot_bytes: int = 64 << 20, n_slots: int = 3, device: Optional[int] = None,
                 max_partitions: int = 64, poll_timeout_ms: int = 100,
                 threaded: bool = True, commit_all_partitions: bool = False,
                 n_producers: int = 2, input_mode: str = "auto", numa_affinity: bool = True,
                 max_latency_ms: Optional[float] = None, output_bytes: Optional[int] = None,
                 json_single_pass: Optional[bool] = None):
        self.consumer = consumer
        self.schema = schema
        self.batch_size = int(batch_size)
        self.drop_last = bool(drop_last)
        # Record-batch input may be compressed, and a decompression unit (an LZ4 block, the DEFLATE
        # stream or the zstd frames of one batch) is decoded by one warp from end to end: a slot of
        # 64 KB batches occupies 5 warps per SM for milliseconds.  What fills the GPU then is slots
        # in flight (each decompresses on its own stream): at least RAW_MIN_SLOTS of them.
        raw_input = input_mode in ("raw", "raw_into", "raw_dma")
        self.engine = _engine.acquire(schema, batch_size, slot_bytes=slot_bytes,
                                      n_slots=max(n_slots, 2 + (int(n_producers) if threaded else 0),
                                                  RAW_MIN_SLOTS if raw_input else 0),
                                      device=device, n_partitions=max_partitions, output_bytes=output_bytes)
        if json_single_pass is not None:
            self.engine.set_json_fused(bool(json_single_pass))
        self.poll_timeout_ms = poll_timeout_ms
        # kafka-python ends `for record in consumer` after consumer_timeout_ms without a record
        # (default: never).  The GPU path follows the same setting; the in-memory stub
        # advertises 0 ("end of log is end of stream").
        timeout = getattr(consumer, "consumer_timeout_ms", None)
        if timeout is None:
            cfg = getattr(consumer, "config", None)
            timeout = cfg.get("consumer_timeout_ms", float("inf")) if isinstance(cfg, dict) else float("inf")
        self.idle_timeout_s = float(timeout) / 1e3
        self.threaded = bool(threaded)
        #: serialises consumer calls: the producer thread polls, the user's thread commits,
        #: and kafka-python consumers are not safe for concurrent calls
        self.consumer_lock = threading.Lock()
        #: producers take turns gathering (poll + harvest) so slots follow consumption order
        self._order_lock = threading.Lock()
        self._mk_offset = _offset_struct()
        self._om_make, self._om_rest = _offset_make()
        self._tp_cls = _tp_struct()
        self._tp_of: Dict[int, object] = {}        # dense partition index -> TopicPartition
        self._dense_of: Dict[object, int] = {}     # TopicPartition -> dense index
        # how records reach the slot:
        #   "records"  consumer.poll() -> ConsumerRecord objects (kafka-python as it is)
        #   "packed"   consumer.poll_packed() -> flat arrays (additive protocol)
        #   "raw"      consumer.poll_raw() -> RecordBatch v2 blobs of the fetch response, split on
        #              the GPU (additive protocol)
        #   "auto"     packed if the consumer offers it, else records
        #   "raw_into" consumer.poll_raw_into(buffer): the consumer receives the fetch response
        #              straight into the slot's page-locked buffer -- no gather copy on the host
        #   "raw_dma"  consumer.poll_raw() blobs are copied to the device straight out of the
        #              consumer's own buffers (page-locked once if it offers fetch_arenas())
        if input_mode not in ("auto", "records", "packed", "raw", "raw_into", "raw_dma"):
            raise ValueError(f"unknown input_mode {input_mode!r}")
        if input_mode == "auto":
            input_mode = "packed" if hasattr(consumer, "poll_packed") else "records"
        if input_mode == "packed" and not hasattr(consumer, "poll_packed"):
            raise TypeError("input_mode='packed' needs a consumer with poll_packed()")
        if input_mode in ("raw", "raw_dma") and not hasattr(consumer, "poll_raw"):
            raise TypeError(f"input_mode={input_mode!r} needs a consumer with poll_raw()")
        if input_mode == "raw_into" and not hasattr(consumer, "poll_raw_into"):
            raise TypeError("input_mode='raw_into' needs a consumer with poll_raw_into()")
        self.input_mode = input_mode
        self._raw = input_mode in ("raw", "raw_into", "raw_dma")
        if input_mode == "raw_dma" and hasattr(consumer, "fetch_arenas"):
            for arena in consumer.fetch_arenas():
                _engine.register_arena(arena)
        self._raw_hold = {}                        # slot -> fetch whose memory the H2D copies read
        self._packed = input_mode == "packed"
        self._leftover = []                        # (tp, records) polled but not fitting the last slot
        self._mean_len = 0.0                       # running mean of the value size (poll sizing)
        # raw input: decompressed bytes per fetched byte.  A slot must hold the records of its fetch
        # *after* decompression, so fetches of compressed topics ask for less than a slot; until
        # the first slot came back the codec is unknown and the budget is conservative.
        self._raw_expansion = 4.0
        self._raw_seen = False
        self._raw_bytes_of = {}
        self._field_idx = {}                       # record type -> (value index, offset index)
        self.n_producers = max(1, int(n_producers))
        # host threads of this rank stay on the cores of the GPU's NUMA node / on the rank's share
        # of the cores (affinity.py); None = leave them alone
        self._cpus = _affinity.rank_cpus(self.engine.device) if numa_affinity else None
        if self._cpus:
            arr = (_ctypes.c_int * len(self._cpus))(*self._cpus)
            self.engine.lib.tk_pack_set_affinity(arr, len(self._cpus))
            self.engine.pack_threads = max(1, min(self.engine.pack_threads, len(self._cpus)))
        self._threads = []
        self._seq = 0
        self._eof = False
        self._last_column = None
        self.commit_all_partitions = bool(commit_all_partitions)
        self._closed = False
        #: a partly filled slot is shipped once its first record has waited this long (records
        #: mode; the other protocols ship one fetch per slot anyway).  None = only when the
        #: stream pauses or the slot is full.
        self.max_latency_s = None if max_latency_ms is None else float(max_latency_ms) / 1e3
        # what an early close() needs to seek the consumer back (see _rewind)
        self._iterating = False
        self._base_delivered = None                # (commit columns of a slot, index of its last yielded batch)
        self._base_tails = []                      # [(_Pending, first record not in a yielded batch)], in order
        self._cur = None                           # (_Pending, result, commit columns) of the slot being handed out
        self.delivered_index = -1                  # position of the batch being yielded within that slot
        self._pending = collections.deque()
        self._ready = {}
        self._stop = threading.Event()
        self._free_q = None
        self._saved_switch = None
        self.stats = collections.Counter()
        self.errors_by_status = collections.Counter()   # status code -> slots whose first bad record had it
        self._t_start = self._t_end = None

    # ------------------------------------------------------------------ lifecycle
    def close(self):
        if self._closed:
            return
        self._closed = True
        if self._t_start is not None and self._t_end is None:
            self._t_end = time.perf_counter()
        self._stop.set()
        alive = []
        if self._saved_switch is not None:
            sys.setswitchinterval(self._saved_switch)
            self._saved_switch = None
        if self._threads:
            try:
                self._free_q.put_nowait(None)
            except Exception:
                pass
            for t in self._threads:
                t.join(timeout=30)
            alive = list(self._threads)
            self._threads = []
        if self._iterating and not any(t.is_alive() for t in alive):
            self._iterating = False
            try:
                self._rewind()
            except Exception as exc:  # pragma: no cover - depends on the consumer
                import logging
                logging.getLogger(__name__).error("Could not seek the consumer back to the last delivered batch: %s", exc)
        # (a producer that did not stop in time may still touch its slot: such an engine is not reused)
        if any(t.is_alive() for t in alive):
            self.engine.close()
        else:
            _engine.release(self.engine)

    def snapshot_stats(self) -> dict:
        """Counters so far plus the derived rates (``KafkaDataset.ingest_stats``)."""
        out = dict(self.stats)
        t0 = getattr(self, "_t_start", None)
        if t0 is not None:
            dt = max((self._t_end or time.perf_counter()) - t0, 1e-9)
            out["seconds"] = dt
            out["records_per_s"] = out.get("records", 0) / dt
            out["h2d_GBps"] = out.get("bytes", 0) / dt / 1e9
        out["errors_by_status"] = dict(self.errors_by_status)
        return out

    def _rewind(self):
        """Early close (``break`` out of the loop, an exception): the producers have polled ahead of
        what was delivered.  Seek every partition back to the position right after the last
        yielded batch, so that iterating the dataset again continues there -- the reference,
        which never reads ahead, behaves like that by construction (``src/kafka_dataset.py:156``).
        Needs ``consumer.seek(TopicPartition, offset)`` (kafka-python has it)."""
        seek = getattr(self.consumer, "seek", None)
        if seek is None:
            return
        target = {}
        delivered, tails = self._base_delivered, list(self._base_tails)
        if self._cur is not None:
            item, res, cols = self._cur
            if self.delivered_index >= 0:
                delivered = (cols, self.delivered_index)
                tails = [(item, int(res.batch_end[self.delivered_index]) + 1)]
            else:
                tails.append((item, 0))
        if delivered is not None:
            cols, j = delivered
            col = cols[j]
            for p in np.flatnonzero(col >= 0).tolist():
                target[p] = int(col[p])

        def first_seen(parts, offs, start=0):
            """First offset of every partition not delivered yet, in consumption order."""
            if isinstance(parts, dict):                      # raw slot: {partition: fetch offset}
                for p, off in parts.items():
                    target.setdefault(int(p), int(off))
                return
            if parts is None or start >= len(parts):
                return
            sub = np.asarray(parts[start:])
            uniq, idx = np.unique(sub, return_index=True)
            for p, i in zip(uniq.tolist(), idx.tolist()):
                target.setdefault(int(p), int(offs[start + i]))

        for item, k0 in tails:
            first_seen(item.partitions, item.offsets, k0)
        for item in self._pending:
            first_seen(item.partitions, item.offsets)
        for seq in sorted(self._ready):
            it = self._ready[seq]
            if it is not None:
                first_seen(it[3], it[4])
        for tp, recs in self._leftover:
            if recs:
                r = recs[0]
                target.setdefault(self._dense(tp), int(getattr(r, "offset", None) if not isinstance(r, tuple)
                                                       else r[self._indices(r)[1]]))
        with self.consumer_lock:
            for p, off in target.items():
                tp = self._tp_of.get(p)
                if tp is not None:
                    seek(tp, off)
        self._pending = collections.deque()
        self._ready = {}
        self._leftover = []

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

    # Both consumer protocols are split in two phases so that several producer threads can
    # work on different slots: _gather_* talks to the consumer (serialised, in consumption
    # order), _pack_* moves the payload into pinned memory (parallel, GIL released).
    def _gather_packed(self, slot: int, scratch):
        eng = self.engine
        idle_since = None
        while True:
            t0 = time.perf_counter()
            with self.consumer_lock:
                run = self.consumer.poll_packed(eng.slot_records, eng.slot_bytes)
            self.stats["t_poll"] += time.perf_counter() - t0
            if run is not None and len(run):
                break
            if idle_since is None:
                idle_since = time.monotonic()
            if self._stop.is_set() or time.monotonic() - idle_since >= self.idle_timeout_s:
                return 0, 0, None
            time.sleep(min(self.poll_timeout_ms, 50) / 1e3)
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
        return len(parts), int(run.starts[len(parts)]) - int(run.starts[0]), run

    def _pack_packed(self, slot: int, scratch, n: int, run):
        t0 = time.perf_counter()
        parts = np.asarray(run.partitions)
        n, nbytes = self.engine.fill_packed(slot, run.values, run.starts, parts, run.offsets)
        self.stats["t_pack"] += time.perf_counter() - t0
        return parts, np.asarray(run.offsets)

    def _gather_raw(self, slot: int, scratch):
        """Next fetch as RecordBatch v2 blobs: no per-record work on the host at all."""
        eng = self.engine
        idle_since = None
        while True:
            t0 = time.perf_counter()
            budget = eng.slot_bytes
            if self._raw_expansion > 1.0:
                budget = max(int(eng.slot_bytes / (self._raw_expansion * 1.15)), min(eng.slot_bytes, 4096))
            with self.consumer_lock:
                if self.input_mode == "raw_into":
                    fetch = self.consumer.poll_raw_into(eng.host_views(slot)[0][:budget], eng.slot_records)
                else:
                    fetch = self.consumer.poll_raw(budget, eng.slot_records)
            self.stats["t_poll"] += time.perf_counter() - t0
            if fetch is not None and len(fetch.blobs):
                break
            if idle_since is None:
                idle_since = time.monotonic()
            if self._stop.is_set() or time.monotonic() - idle_since >= self.idle_timeout_s:
                return 0, 0, None
            time.sleep(min(self.poll_timeout_ms, 50) / 1e3)
        topic = getattr(fetch, "topic", None) or "topic"
        nbytes = 0
        into = self.input_mode == "raw_into"
        for blob in fetch.blobs:
            p = int(blob[0])
            if p not in self._tp_of:                       # dense index == partition id (single topic)
                if not 0 <= p < eng.n_partitions:
                    raise RuntimeError(f"partition {p} outside [0, max_partitions={eng.n_partitions})")
                tp = self._tp_cls(topic, p)
                self._tp_of[p] = tp
                self._dense_of[tp] = p
            nbytes += int(blob[2]) if into else len(blob[1])
        if nbytes > eng.slot_bytes:
            raise ValueError(f"a fetch of {nbytes} bytes exceeds the slot size ({eng.slot_bytes} bytes)")
        # n is not known before the device has split the batches: report blobs, fix up in _pack_raw
        return len(fetch.blobs), nbytes, fetch

    def _pack_raw(self, slot: int, scratch, n_blobs: int, fetch):
        """Blobs -> pinned buffer (multi-threaded gather); the batch headers give the record count;
        H2D and the split kernels are queued.  Returns (record entries, raw bytes)."""
        eng = self.engine
        t0 = time.perf_counter()
        k = len(fetch.blobs)
        if self.input_mode == "raw_into":
            # the bytes are already in the slot's pinned buffer: (partition, start, length, fetch offset, delta)
            total = sum(int(b[2]) for b in fetch.blobs)
            n = eng.submit_recordbatches(slot, total, [(int(b[1]), int(b[2]), int(b[0]), int(b[3]), int(b[4]))
                                                       for b in fetch.blobs])
            self.stats["t_pack"] += time.perf_counter() - t0
            self._raw_bytes_of[slot] = total
            return n, total
        if self.input_mode == "raw_dma":
            lens = [len(b[1]) for b in fetch.blobs]
            starts = [0]
            for ln in lens:
                starts.append(starts[-1] + ln)
            self._raw_hold[slot] = fetch           # the copies read the consumer's memory: keep it alive
            n = eng.submit_recordbatches(
                slot, starts[k], [(starts[i], lens[i], int(b[0]), int(b[2]), int(b[3]))
                                  for i, b in enumerate(fetch.blobs)],
                host_ptrs=[np.asarray(b[1]).ctypes.data for b in fetch.blobs])
            self.stats["t_pack"] += time.perf_counter() - t0
            self._raw_bytes_of[slot] = starts[k]
            return n, starts[k]
        ptrs = np.fromiter((b[1].ctypes.data for b in fetch.blobs), dtype=np.uint64, count=k)
        lens = np.fromiter((len(b[1]) for b in fetch.blobs), dtype=np.uint32, count=k)
        parts = np.fromiter((int(b[0]) for b in fetch.blobs), dtype=np.int32, count=k)
        fetch_offsets = np.fromiter((int(b[2]) for b in fetch.blobs), dtype=np.int64, count=k)
        eng.pack_pointers(slot, ptrs, lens, parts, fetch_offsets, k)      # bytes back to back
        starts = np.zeros(k + 1, dtype=np.int64)
        np.cumsum(lens, out=starts[1:])
        n = eng.submit_recordbatches(slot, int(starts[k]),
                                     [(int(starts[i]), int(lens[i]), int(parts[i]), int(fetch_offsets[i]),
                                       int(fetch.blobs[i][3])) for i in range(k)])
        self.stats["t_pack"] += time.perf_counter() - t0
        self._raw_bytes_of[slot] = int(starts[k])
        return n, int(starts[k])

    @staticmethod
    def _fetch_offsets(fetch, into: bool):
        """{dense partition: fetch offset} of a raw fetch (first blob of every partition)."""
        out = {}
        for b in fetch.blobs:
            out.setdefault(int(b[0]), int(b[3] if into else b[2]))
        return out

    def _note_raw_expansion(self, slot: int, n_records: int):
        """Feedback for the fetch budget (the split of this slot is done: no waiting here)."""
        raw = self._raw_bytes_of.get(slot, 0)
        if not raw:
            return
        _, value_bytes, _ = self.engine.recordbatch_result(slot)
        decompressed = value_bytes + 12 * n_records          # values + record framing, roughly
        obs = decompressed / raw if decompressed > raw * 1.02 else 1.0
        if not self._raw_seen:
            self._raw_expansion, self._raw_seen = obs, True
        else:
            self._raw_expansion = max(obs, 0.9 * self._raw_expansion + 0.1 * obs)

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

    def _gather_poll(self, slot: int, scratch):
        """Flatten ``consumer.poll()`` results (dict order, list order == iterator order).

        Per record only ~25 ns of C run under the GIL (``_tkpack.harvest``: address, length
        and offset of each value); the payload bytes are gathered afterwards by ``_pack_poll``.
        """
        eng = self.engine
        ptrs, lens, offs, parts, addr = scratch
        n = 0
        nbytes = 0
        idle_since = None
        hold = []                       # keeps the record objects (and their bytes) alive until packed
        full = False
        t_first = None                  # when the slot received its first record
        while not full and n < eng.slot_records:
            if t_first is not None and self.max_latency_s is not None and \
                    time.monotonic() - t_first >= self.max_latency_s:
                break                   # a trickling topic: ship what is there
            if self._leftover:
                items, self._leftover = self._leftover, []
            else:
                t0 = time.perf_counter()
                # ask for about as many records as still fit (mean value size seen so far): records
                # polled beyond the slot's byte budget would be harvested twice
                want = eng.slot_records - n
                if self._mean_len:
                    want = min(want, int((eng.slot_bytes - nbytes) / self._mean_len * 1.02) + 16)
                timeout_ms = self.poll_timeout_ms
                if t_first is not None and self.max_latency_s is not None:
                    left = self.max_latency_s - (time.monotonic() - t_first)
                    timeout_ms = max(0, min(timeout_ms, int(left * 1e3)))
                with self.consumer_lock:
                    polled = self.consumer.poll(timeout_ms=timeout_ms,
                                                max_records=max(1, min(want, 1 << 16)))
                self.stats["t_poll"] += time.perf_counter() - t0
                if not polled:
                    # a partly filled slot is shipped as soon as the stream pauses (bounded
                    # latency); an empty one keeps polling until the consumer's idle timeout
                    if n > 0 or self._stop.is_set():
                        break
                    if idle_since is None:
                        idle_since = time.monotonic()
                    if time.monotonic() - idle_since >= self.idle_timeout_s:
                        break
                    continue
                idle_since = None
                items = list(polled.items())
            items = [it for it in items if it[1]]
            if not items:
                continue
            dense = [self._dense(tp) for tp, _ in items]
            vi, oi = self._indices(items[0][1][0])
            t0 = time.perf_counter()
            taken, tot, g_done, partial, keep = _tkpack.harvest_groups(
                items, dense, vi, oi, eng.slot_records - n, eng.slot_bytes - nbytes,
                addr[0] + 8 * n, addr[1] + 4 * n, addr[2] + 8 * n, addr[3] + 4 * n)
            self.stats["t_harvest"] += time.perf_counter() - t0
            hold.append(items)
            if keep is not None:
                hold.append(keep)
            n += taken
            nbytes += tot
            if taken and t_first is None:
                t_first = time.monotonic()
            if taken:
                m = tot / taken
                self._mean_len = m if not self._mean_len else 0.75 * self._mean_len + 0.25 * m
            if g_done < len(items):
                # the slot is full inside group g_done: records cannot be un-polled, so the
                # rest waits, in consumption order, for the next slot
                if taken == 0 and n == 0:
                    raise ValueError("a single record exceeds the slot size "
                                     f"({eng.slot_bytes} bytes); raise device_slot_bytes")
                tp, recs = items[g_done]
                self._leftover = [(tp, recs[partial:])] + items[g_done + 1:]
                full = True
        return n, nbytes, hold

    def _pack_poll(self, slot: int, scratch, n: int, hold):
        ptrs, lens, offs, parts, _ = scratch
        t0 = time.perf_counter()
        self.engine.pack_pointers(slot, ptrs, lens, parts, offs, n)
        self.stats["t_pack"] += time.perf_counter() - t0
        del hold
        return parts[:n].copy(), offs[:n].copy()

    def _new_scratch(self):
        cap = self.engine.slot_records
        arrs = (np.zeros(cap, np.uint64), np.zeros(cap, np.uint32), np.zeros(cap, np.int64),
                np.zeros(cap, np.int32))
        return arrs + (tuple(a.ctypes.data for a in arrs),)

    # ------------------------------------------------------------------ iteration
    def _error(self, item, res) -> PayloadError:
        coords = (item.partitions, item.offsets) if item.offsets is not None else (None, None)
        err = self.engine.error_for(item.slot, res, *coords)
        tp = self._tp_of.get(err.partition)
        if tp is not None:                      # dense index -> the consumer's partition id
            err.partition = getattr(tp, "partition", err.partition)
        return err

    def _offsets_map(self, column: np.ndarray):
        """{TopicPartition: OffsetAndMetadata} to commit after a batch.

        Only partitions whose position moved since the previous batch's map are listed
        (``commit_all_partitions=False``): the broker's committed-offset table after the commit
        is the same as with the full map the reference sends, at a fraction of the host cost.
        """
        mk = self._mk_offset
        tp_of = self._tp_of
        prev = self._last_column
        if prev is None or self.commit_all_partitions:
            idx = np.flatnonzero(column >= 0)
        else:
            idx = np.flatnonzero(column != prev)
        self._last_column = column
        vals = column[idx].tolist()
        return {tp_of[p]: mk(v) for p, v in zip(idx.tolist(), vals)}

    def _offsets_maps(self, cols: np.ndarray):
        """``_offsets_map`` for every row of ``cols`` (one row per closed batch of a slot) at once:
        one vectorised comparison per slot instead of a handful of numpy calls per batch, which is
        what the host spends its time on when batches are small."""
        nc = cols.shape[0]
        if nc == 0:
            return []
        prev = self._last_column
        if prev is None or self.commit_all_partitions:
            head = cols[:1] >= 0
            changed = np.concatenate([head, (cols[1:] >= 0) if self.commit_all_partitions
                                      else (cols[1:] != cols[:-1])], axis=0)
        else:
            changed = np.concatenate([cols[:1] != prev[None, :], cols[1:] != cols[:-1]], axis=0)
        self._last_column = cols[nc - 1]
        rows, parts = np.nonzero(changed)
        vals = cols[rows, parts].tolist()
        cut = np.searchsorted(rows, np.arange(nc + 1)).tolist()
        # Everything per entry runs in C: the OffsetAndMetadata tuples come from the class's _make
        # over zipped columns, the TopicPartitions from a table lookup; with small batches (one or
        # two partitions move per batch) building these maps is what the host spends its time on.
        tps = list(map(self._tp_of.__getitem__, parts.tolist()))
        oms = list(map(self._om_make, zip(vals, *self._om_rest)))
        return [{tps[a]: oms[a]} if b == a + 1 else dict(zip(tps[a:b], oms[a:b]))
                for a, b in zip(cut, cut[1:])]

    def _produce(self, free_q, deliver):
        """Producer thread.  Gather (consumer calls, serialised, defines the slot order) then
        pack + submit (parallel across producer threads)."""
        raw = self._raw
        gather = self._gather_raw if raw else self._gather_packed if self._packed else self._gather_poll
        pack = self._pack_raw if raw else self._pack_packed if self._packed else self._pack_poll
        scratch = self._new_scratch() if self.input_mode == "records" else None
        _affinity.pin_current_thread(self._cpus)
        try:
            while not self._stop.is_set():
                slot = free_q.get()
                if slot is None or self._stop.is_set():
                    free_q.put(None)            # let the sibling producers see it too
                    return
                with self._order_lock:
                    if self._eof:
                        return
                    with _range("torchkafka.gather"):
                        n, nbytes, token = gather(slot, scratch)
                    seq = self._seq
                    self._seq += 1
                    if n == 0:
                        self._eof = True
                if n == 0:
                    deliver(seq, None)
                    return
                if raw:
                    # the device splits the batches: record coordinates stay there
                    fo = self._fetch_offsets(token, self.input_mode == "raw_into")
                    with _range("torchkafka.pack+submit"):
                        n, nbytes = pack(slot, scratch, n, token)
                    deliver(seq, (slot, n, nbytes, fo, None))
                    continue
                with _range("torchkafka.pack"):
                    parts, offs = pack(slot, scratch, n, token)
                with _range("torchkafka.submit"):
                    self.engine.submit(slot, n, nbytes)
                deliver(seq, (slot, n, nbytes, parts, offs))
        except BaseException as exc:  # forwarded to the consuming thread
            with self._cv:
                self._failure = exc
                self._cv.notify_all()

    def _filled_slots(self):
        """Generator of (slot, n, nbytes, partitions, offsets) in consumption order, plus the
        callable that hands a drained slot back."""
        eng = self.engine
        self._seq = 0
        self._eof = False
        if not self.threaded:
            raw = self._raw
            gather = self._gather_raw if raw else self._gather_packed if self._packed else self._gather_poll
            pack = self._pack_raw if raw else self._pack_packed if self._packed else self._pack_poll
            scratch = self._new_scratch() if self.input_mode == "records" else None
            free = collections.deque(range(eng.n_slots))

            def gen():
                while True:
                    slot = free.popleft()       # never empty: pending depth < n_slots
                    n, nbytes, token = gather(slot, scratch)
                    if n == 0:
                        return
                    if raw:
                        fo = self._fetch_offsets(token, self.input_mode == "raw_into")
                        n, nbytes = pack(slot, scratch, n, token)
                        yield slot, n, nbytes, fo, None
                        continue
                    parts, offs = pack(slot, scratch, n, token)
                    eng.submit(slot, n, nbytes)
                    yield slot, n, nbytes, parts, offs

            return gen(), free.append
        free_q = queue.Queue()
        for s in range(eng.n_slots):
            free_q.put(s)
        self._free_q = free_q
        self._cv = threading.Condition()
        self._failure = None
        ready = self._ready = {}

        def deliver(seq, item):
            with self._cv:
                ready[seq] = item
                self._cv.notify_all()

        self._threads = [threading.Thread(target=self._produce, args=(free_q, deliver),
                                          name=f"torchkafka-producer-{k}", daemon=True)
                         for k in range(self.n_producers)]
        for t in self._threads:
            t.start()

        def gen():
            want = 0
            while True:
                with self._cv:
                    while want not in ready and self._failure is None:
                        self._cv.wait(timeout=1.0)
                    if self._failure is not None:
                        raise self._failure
                    item = ready.pop(want)
                want += 1
                if item is None:
                    return
                yield item

        return gen(), free_q.put

    def __iter__(self):
        """(batch, offsets) pairs; see ``slot_batches`` for the per-slot form."""
        for views, maps in self.slot_batches():
            for j in range(len(views)):
                self.delivered_index = j
                yield views[j], maps[j]

    def slot_batches(self):
        """Per drained slot, the list of batches it closed and the list of their commit maps
        (same length), in order.  A caller that hands the batches out one by one sets
        ``delivered_index`` to the position of the batch it is about to yield, so that an early
        ``close()`` knows where to seek the consumer back to (``__iter__`` does exactly that;
        ``auto_commit`` uses this form directly: with small batches the cost per batch is the
        depth of the generator chain)."""
        eng = self.engine
        B = self.batch_size
        self._leftover = []
        pending = self._pending = collections.deque()
        self._base_delivered = None
        self._base_tails = []
        self._cur = None
        self.delivered_index = -1
        self._iterating = True
        self._t_start, self._t_end = time.perf_counter(), None
        depth = 1 if self.threaded else max(1, eng.n_slots - 1)
        prev_slot, prev_out = None, None
        open_lmax = 0                  # longest record among the rows of the open batch
        last = None                    # (_Pending, SlotResult) of the most recent drained slot
        # The producers hold the GIL in short C stretches (poll glue, harvest); with CPython's
        # default 5 ms switch interval the consuming thread would wait that long to get the GIL
        # back after every blocking call.  Ask for prompt hand-offs while the pipeline runs.
        if self.threaded and self._saved_switch is None:
            self._saved_switch = sys.getswitchinterval()
            sys.setswitchinterval(min(self._saved_switch, 5e-5))
        slots, release = self._filled_slots()

        def drain(item):
            nonlocal open_lmax, last
            t0 = time.perf_counter()
            with _range("torchkafka.drain"):
                res = eng.result(item.slot)
            self.stats["t_wait_result"] += time.perf_counter() - t0
            if self._raw:
                self._note_raw_expansion(item.slot, res.n_records)
                self._raw_hold.pop(item.slot, None)
            # the error report reads the slot's device buffers: build it before the slot goes back
            # to the producers, whose next submit overwrites them
            error = self._error(item, res) if res.n_errors else None
            release(item.slot)         # pinned memory is reusable: H2D and kernels are done
            eng.publish(item.out)      # caller's stream waits for this slot only
            self.stats["records"] += res.n_records
            self.stats["kept"] += res.n_kept
            if res.n_errors:
                self.stats["decode_errors"] += res.n_errors
                self.errors_by_status[int(error.status)] += 1
            err_rec = res.first_error_rec if res.n_errors else None
            first = B - res.carry_in
            nc = res.n_closed
            if nc:
                lmax = res.batch_lmax[:nc].copy()
                lmax[0] = max(int(lmax[0]), open_lmax)
                views = eng.batch_views(item.out, first, nc, lmax)
                commit_cols = np.ascontiguousarray(res.commit.T[:nc])
                # batches in front of the first bad record are delivered, the one holding it raises
                good = nc if err_rec is None else int(np.searchsorted(res.batch_end[:nc], err_rec, side="left"))
                maps = self._offsets_maps(commit_cols[:good])
                self.stats["batches"] += good
                # (what an early close() seeks back to: the positions after the last yielded batch)
                self._cur = (item, res, commit_cols)
                self.delivered_index = -1
                if good:
                    yield views[:good] if good < nc else views, maps
                    self._base_delivered = (commit_cols, good - 1)
                    self._base_tails = [(item, int(res.batch_end[good - 1]) + 1)]
                else:
                    self._base_tails.append((item, 0))
                self._cur = None
                if good < nc:
                    raise error
            if err_rec is not None:
                # bad record sits in the still-open batch: it surfaces with that batch
                raise error
            if not nc:
                self._base_tails.append((item, 0))      # all of it sits in the still-open batch
            tail_lmax = int(res.batch_lmax[res.n_closed])
            open_lmax = max(open_lmax, tail_lmax) if res.n_closed == 0 else tail_lmax
            last = (item, res)

        t_get = time.perf_counter()
        for slot, n, nbytes, parts, offs in slots:
            t0 = time.perf_counter()
            self.stats["t_wait_slot"] += t0 - t_get
            with _range("torchkafka.process"):
                out = eng.process(slot, n, prev_slot, prev_out)
            self.stats["t_launch"] += time.perf_counter() - t0
            pending.append(_Pending(slot, n, out, parts, offs))
            prev_slot, prev_out = slot, out
            self.stats["slots"] += 1
            self.stats["bytes"] += nbytes
            if len(pending) > depth:
                yield from drain(pending.popleft())
            t_get = time.perf_counter()
        while pending:
            yield from drain(pending.popleft())
        # stream exhausted: the open batch becomes the short tail batch (fetch.py:39-42)
        if last is not None and not self.drop_last:
            item, res = last
            if res.carry_out:
                first = B + res.n_kept - res.carry_out
                batch = eng.batch_view(item.out, first, res.carry_out, open_lmax)
                self.stats["batches"] += 1
                self._iterating = False         # everything the consumer handed out is delivered
                yield [batch], [self._offsets_map(np.ascontiguousarray(res.commit[:, res.n_closed]))]
        self._iterating = False
