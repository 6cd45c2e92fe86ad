"""``KafkaDataset``: the drop-in dataset class.

Public surface and behaviour follow the reference class
(``/root/reference/src/kafka_dataset.py:31-247``; contract frozen in SURVEY.md
§4.1 and §8b): same constructor, ``_process``, ``new_consumer``,
``init_worker``, ``placeholder``, ``commit``, ``commit_worker``, ``close``, same
exception types and messages, same log lines.

What is new is *additive*: a subclass may set ``payload_schema`` (see
``torchkafka_b200.schema``).  Such a dataset never calls ``_process``; records
are polled in bulk, packed CSR-style into pinned memory, copied to the GPU on a
side stream and decoded / filtered / collated / offset-reduced by the CUDA
kernels behind the C-ABI in ``include/torchkafka_b200.h``.  There is no CPU
fallback for a schema dataset: if the CUDA library or a GPU is missing, using it
raises.

Two things differ from the reference on purpose, both on the schema path only:

* offsets are committed *explicitly* (``consumer.commit({tp: offset})``), never
  as "the consumer's current position" (reference ``kafka_dataset.py:130``),
  because the GPU path reads ahead of the batch it has delivered; the committed
  values are exactly those the reference commits in single-process mode
  (SURVEY.md §4.1-3/5);
* a record outside its schema raises ``PayloadError`` (with partition/offset)
  when the batch containing it is requested, instead of whatever
  ``json.loads`` / ``torch.tensor`` would have raised inside ``_process``.
"""
from __future__ import annotations

import collections
import contextlib
import logging
import os
import signal
import sys

import torch.multiprocessing as mp
from torch.utils.data import IterableDataset, get_worker_info

try:  # kafka-python is optional at import time: tests and benches inject a stub consumer
    from kafka import KafkaConsumer as _KafkaConsumer
    from kafka.errors import CommitFailedError as _CommitFailedError
except Exception:  # pragma: no cover - depends on the environment
    _KafkaConsumer = None

    class _CommitFailedError(Exception):
        """Placeholder used when kafka-python is not installed."""


_logger = logging.getLogger(__name__)


def _commit_signal():
    # same platform table as the reference (kafka_dataset.py:50-55)
    if sys.platform in {"linux", "linux2"}:
        return signal.SIGUSR1
    if sys.platform in {"darwin", "win32", "win64"}:
        return signal.SIGINT
    raise RuntimeError(f"Unsupported platform '{sys.platform}'.")


def _commit_failed_types():
    """Exception classes treated as 'commit failed' (logged, not raised)."""
    kinds = [_CommitFailedError]
    mod = sys.modules.get("kafka.errors")
    cls = getattr(mod, "CommitFailedError", None) if mod is not None else None
    if isinstance(cls, type) and cls not in kinds:
        kinds.append(cls)
    try:
        from .testing.stub_kafka import CommitFailedError as stub_cls
        if stub_cls not in kinds:
            kinds.append(stub_cls)
    except Exception:  # pragma: no cover
        pass
    return tuple(kinds)


class KafkaDataset(IterableDataset):
    """Stream samples from Kafka into a PyTorch ``DataLoader``.

    Subclass it and override ``_process`` (reference behaviour), or declare a
    ``payload_schema`` to take the GPU path.  Every positional / keyword
    argument goes to the Kafka consumer; auto-commit is always disabled.
    """

    _COMMIT_SIGNAL = _commit_signal()

    #: optional schema (``JsonArray`` / ``TokenList`` / ``RawBytes`` / ``StructBytes`` / ``ImageU8``)
    payload_schema = None
    #: GPU-path tuning knobs (bytes per staging slot, slots in flight, device index)
    device_slot_bytes = 64 << 20
    device_slots = 3
    device_index = None
    max_partitions = 64
    #: when set, plain iteration (``for b in DataLoader(ds, batch_size=None)``) yields whole
    #: device batches of this size instead of single samples
    device_batch_size = None
    #: poll + pack + H2D on a producer thread (overlaps the user's loop); False = same thread
    device_producer_thread = True
    device_producers = 2
    #: send every assigned partition with each commit (reference behaviour) instead of only the
    #: partitions whose position moved since the previous commit (same broker state, cheaper)
    commit_all_partitions = False
    #: how records reach the GPU: "auto" (``poll_packed()`` if the consumer has it, else
    #: ``poll()``), "records", "packed", or "raw" (``poll_raw()``: RecordBatch v2 blobs of the
    #: fetch response, CRC-checked and split on the device), "raw_into" (``poll_raw_into(buffer)``:
    #: the consumer receives the response straight into the page-locked staging slot), "raw_dma"
    #: (``poll_raw()`` blobs go to the device straight out of the consumer's own buffers,
    #: page-locked once via ``fetch_arenas()``)
    device_input = "auto"
    #: pin the producer threads and the pack helper's threads to the cores of the GPU's NUMA node,
    #: and give each rank of a multi-GPU job its own share of the cores (affinity.py)
    device_numa_affinity = True
    #: ship a partly filled staging slot once its first record has waited this many milliseconds
    #: (``poll()`` input): bounds the delivery latency on a slow topic, where a slot would otherwise
    #: go out only when full or when ``poll()`` comes back empty.  None = no time bound.
    device_max_latency_ms = None
    #: budget for the output tensors of one staging slot (default 2 GiB).  Rows are as wide as the
    #: schema allows (``max_len``) and there is one per record, so a slot of very short records
    #: under a wide schema is cut by this budget rather than by ``device_slot_bytes``.
    device_output_bytes = None
    #: JSON arrays through the single-pass kernel (``tk_json_ingest``: payload read from HBM once)
    #: instead of the three-kernel chain; None = the library's default (the chain, or
    #: ``TK_JSON_FUSED=1``).  Same results; the chain is the faster one today.
    device_json_single_pass = None
    #: multi-worker CPU path: a worker commits the offsets of the batches the main process has
    #: *received* from it (explicit offsets, same values as single-process iteration), not its
    #: consumer's position, which runs ahead by the DataLoader's prefetched batches and would lose
    #: them on a crash.  False = the reference's behaviour.
    commit_delivered_only = True

    def __init__(self, *args, **kwargs):
        self._worker_id = None          # set by init_worker inside a DataLoader worker
        self._commit_required = False   # raised by the signal handler, cleared after each commit
        self._delivered = None          # shared per-worker count of delivered batches (auto_commit)
        self._delivered_batch_size = None
        self._snapshots = None          # worker: positions at the end of every batch it produced
        self._snapshot_base = 0         # worker: batches whose snapshots were already dropped
        self._explicit_offsets = None   # {TopicPartition: OffsetAndMetadata} of the last device batch
        self._pipeline = None
        #: counters of the last GPU-path iteration (``ingest_stats()`` reads the running one)
        self.last_pipeline_stats = {}
        #: auto_commit(..., commit_barrier=...): the group's commit offsets after the last barrier
        #: (int64[max_partitions], -1 = no rank has consumed that partition) and the barrier's
        #: latencies in microseconds
        self.group_offsets = None
        self.commit_barrier_us = None
        if kwargs.get("_is_placeholder", False):
            self._consumer = None
            return
        if not args:
            raise ValueError(
                "No topic was provided. "
                "Please use the placeholder() method "
                "to create a dataset without consumer."
            )
        self._consumer = self.new_consumer(*args, **kwargs)

    # ------------------------------------------------------------------ lifecycle
    def __del__(self):
        try:
            self.close()
        except Exception:  # interpreter teardown
            pass

    def close(self):
        """Close the consumer; offsets are NOT committed (reference :85-91)."""
        pipe = getattr(self, "_pipeline", None)
        if pipe is not None:
            pipe.close()
            self._pipeline = None
        consumer = getattr(self, "_consumer", None)
        if consumer is not None:
            consumer.close(autocommit=False)
        self._commit_required = False

    # ------------------------------------------------------------------ commits
    def commit(self, signum=None, stack=None):  # pylint: disable=unused-argument
        """Commit offsets now (main process) or flag a commit (worker, via signal).

        Reference ``kafka_dataset.py:93-118``: three-way behaviour kept as is.
        """
        if self._consumer is None:
            raise RuntimeError("Consumer is not initialized.")
        if self._worker_id is None:
            pending = self._explicit_offsets
            pipe = self._pipeline
            if pending is not None and pipe is not None and not _logger.isEnabledFor(logging.DEBUG):
                # GPU path, once per batch: the same steps as _commit_if_required(force=True) without
                # the (disabled) log calls -- with batches of a few records this is the host's hot loop
                try:
                    with pipe.consumer_lock:
                        self._consumer.commit(pending)
                    if self._explicit_offsets is pending:
                        self._explicit_offsets = {}
                except _commit_failed_types():
                    _logger.error("Commit failed.")
                self._commit_required = False
                return
            self._commit_if_required(force=True)
            return
        if signum is None:
            raise RuntimeError("Direct commit should not be used with multiprocessing.")
        if signum != self._COMMIT_SIGNAL:
            raise ValueError(f"Worker {self._worker_id} received a bad signal ({signum}).")
        self._commit_required = True

    def _commit_if_required(self, force: bool = False):
        if not (force or self._commit_required):
            return
        in_worker = self._worker_id is not None
        if in_worker:
            _logger.info("Committing offsets on worker %d.", self._worker_id)
        else:
            _logger.debug("Committing offsets.")
        try:
            if in_worker and self._snapshots is not None:
                self._commit_delivered()
            elif self._explicit_offsets is not None:
                # GPU path: exactly the records up to the delivered batch, never the read-ahead;
                # serialised with the producer thread's poll() (the consumer is not thread safe)
                pipe = self._pipeline
                lock = pipe.consumer_lock if pipe is not None else contextlib.nullcontext()
                pending = self._explicit_offsets
                with lock:
                    self._consumer.commit(pending)
                # committed: start a fresh accumulation (a failed commit keeps its entries)
                if self._explicit_offsets is pending:
                    self._explicit_offsets = {}
            else:
                self._consumer.commit()
        except _commit_failed_types():
            if in_worker:
                _logger.error("Commit failed on worker %d.", self._worker_id)
            else:
                _logger.error("Commit failed.")
        else:
            if in_worker:
                _logger.debug("Committed offsets on worker %d.", self._worker_id)
            else:
                _logger.debug("Committed offsets.")
        finally:
            self._commit_required = False

    def _commit_delivered(self):
        """Worker: commit the positions at the end of the last batch the main process received
        from this worker (its count sits in shared memory: signals may coalesce, counts do not)."""
        from .structs import _offset_struct, _tp_struct
        target = int(self._delivered[self._worker_id])
        snap = None
        while self._snapshots and self._snapshot_base < target:
            snap = self._snapshots.popleft()
            self._snapshot_base += 1
        if snap:
            tp, off = _tp_struct(), _offset_struct()
            self._consumer.commit({tp(t, p): off(o) for (t, p), o in snap.items()})

    # ------------------------------------------------------------------ iteration
    def __iter__(self):
        if self._consumer is None:
            raise RuntimeError("Consumer is not initialized.")
        if self.payload_schema is not None:
            yield from self._iter_device_samples()
            return
        in_worker = self._worker_id is not None
        if in_worker:
            signal.signal(self._COMMIT_SIGNAL, self.commit)
        process = self._process
        track = in_worker and self.commit_delivered_only and self._delivered is not None
        if track:
            # batch boundaries as the DataLoader's fetcher draws them (fetch.py:27-45): a batch
            # closes on its B-th kept sample; records skipped behind it belong to the next one
            per_batch = int(self._delivered_batch_size or 1)
            self._snapshots = collections.deque()
            self._snapshot_base = 0
            positions = {}
            kept = 0
        for record in self._consumer:
            sample = process(record)
            if track:
                positions[(getattr(record, "topic", None), record.partition)] = record.offset + 1
                if sample is not None:
                    kept += 1
                    if kept % per_batch == 0:
                        self._snapshots.append(dict(positions))
            if sample is not None:
                yield sample
            if in_worker:
                # committing from inside the loop (not from the handler) is the
                # reference's 1.1.0 deadlock fix (CHANGELOG.md:15-17)
                self._commit_if_required()
        if in_worker:
            # The reference restores SIG_DFL here (kafka_dataset.py:170-171); auto_commit keeps
            # signalling the worker after its stream ended, and the default action of SIGUSR1
            # kills it ("DataLoader worker is killed by signal").  A finished
            # worker instead honours a pending commit request once more and then ignores further
            # signals.
            if track and kept % per_batch:
                self._snapshots.append(dict(positions))     # the short tail batch
            self._commit_if_required()
            signal.signal(self._COMMIT_SIGNAL, signal.SIG_IGN)

    def _process(self, record):
        """Turn one Kafka record into a sample; return ``None`` to skip it."""
        raise NotImplementedError()

    # ------------------------------------------------------------------ GPU path
    def _open_pipeline(self, batch_size, drop_last):
        if self._worker_id is not None or get_worker_info() is not None:
            raise RuntimeError(
                "A dataset with a payload_schema runs on the GPU and cannot be used "
                "inside DataLoader worker processes (num_workers must be 0)."
            )
        from .pipeline import IngestPipeline  # imports the CUDA library; fails loudly if missing
        if self._pipeline is not None:
            self._pipeline.close()
        self._pipeline = IngestPipeline(
            self._consumer, self.payload_schema, batch_size=batch_size, drop_last=drop_last,
            slot_bytes=self.device_slot_bytes, n_slots=self.device_slots,
            device=self.device_index, max_partitions=self.max_partitions,
            threaded=self.device_producer_thread, n_producers=self.device_producers,
            commit_all_partitions=self.commit_all_partitions, input_mode=self.device_input,
            numa_affinity=self.device_numa_affinity, max_latency_ms=self.device_max_latency_ms,
            output_bytes=self.device_output_bytes, json_single_pass=self.device_json_single_pass,
        )
        return self._pipeline

    def device_batches(self, batch_size: int, drop_last: bool = False):
        """Yield batch-ready device tensors (schema datasets only).

        After each yielded batch, ``commit()`` commits exactly the offsets the
        reference would have committed after the same batch.
        """
        if self._consumer is None:
            raise RuntimeError("Consumer is not initialized.")
        if self.payload_schema is None:
            raise RuntimeError("device_batches() needs a payload_schema.")
        pipe = self._open_pipeline(int(batch_size), bool(drop_last))
        # (offsets of batches that were delivered but never committed -- the loop was left before
        # the commit -- stay pending: they go out with the next commit)
        if self._explicit_offsets is None:
            self._explicit_offsets = {}
        try:
            for batch, offsets in pipe:
                # offsets list the partitions that moved with this batch; merge them so that a
                # caller who commits every few batches still commits everything consumed so far
                self._note_offsets(offsets)
                yield batch
        finally:
            self.last_pipeline_stats = pipe.snapshot_stats()
            pipe.close()
            if self._pipeline is pipe:
                self._pipeline = None

    def ingest_stats(self):
        """Counters of the GPU path (running iteration, else the last one): ``records``, ``kept``,
        ``batches``, ``slots``, ``bytes`` (over the host-device link), ``decode_errors``,
        ``errors_by_status``, ``seconds``, ``records_per_s``, ``h2d_GBps`` and the host-side
        stage times ``t_poll`` / ``t_harvest`` / ``t_pack`` / ``t_launch`` / ``t_wait_slot`` /
        ``t_wait_result`` (seconds, summed over the producer threads)."""
        pipe = self._pipeline
        return pipe.snapshot_stats() if pipe is not None else dict(self.last_pipeline_stats)

    def _note_offsets(self, offsets):
        """Offsets of the batch about to be delivered; they go out with the next commit."""
        pending = self._explicit_offsets
        if pending:
            pending.update(offsets)
        else:
            self._explicit_offsets = offsets

    def _device_slot_batches(self, batch_size: int, drop_last: bool = False):
        """``device_batches`` one staging slot at a time: (list of batches, list of their commit
        maps, pipeline).  The caller sets ``pipeline.delivered_index`` and calls ``_note_offsets``
        per batch (``auto_commit`` does)."""
        if self._consumer is None:
            raise RuntimeError("Consumer is not initialized.")
        if self.payload_schema is None:
            raise RuntimeError("device_batches() needs a payload_schema.")
        pipe = self._open_pipeline(int(batch_size), bool(drop_last))
        if self._explicit_offsets is None:
            self._explicit_offsets = {}
        try:
            for views, maps in pipe.slot_batches():
                yield views, maps, pipe
        finally:
            self.last_pipeline_stats = pipe.snapshot_stats()
            pipe.close()
            if self._pipeline is pipe:
                self._pipeline = None

    def _iter_device_samples(self):
        """Per-sample view of the GPU path, for a plain ``DataLoader(ds, batch_size=B)``.

        Each sample is what the equivalent ``_process`` would have returned (moved
        to the GPU); ``default_collate`` then stacks them.  Correct but slow:
        ``auto_commit(dataloader)`` or ``device_batches()`` deliver whole batches.
        """
        from .schema import JsonArray, StructBytes
        schema = self.payload_schema
        if self.device_batch_size:
            yield from self.device_batches(int(self.device_batch_size), False)
            return
        for batch in self.device_batches(1, False):
            if isinstance(schema, JsonArray) and schema.fixed_len is None:
                n = int(batch.lengths[0])
                yield batch[0][0, :n]
            elif isinstance(schema, StructBytes):
                yield tuple(f[0] for f in batch)
            else:
                yield batch[0]

    # ------------------------------------------------------------------ factories
    @classmethod
    def new_consumer(cls, *args, **kwargs):
        """Build the Kafka consumer (auto-commit forced off).  Override to inject
        configuration or another consumer (reference :188-206, README.md:46-57)."""
        if not args:
            raise ValueError("Cannot create a consumer without topic.")
        kwargs["enable_auto_commit"] = False
        kwargs.pop("_is_placeholder", None)
        if _KafkaConsumer is None:
            consumer_cls = getattr(sys.modules.get("kafka"), "KafkaConsumer", None)
            if consumer_cls is None:
                raise ImportError(
                    "kafka-python is not installed; install it or override "
                    "new_consumer() to provide a consumer."
                )
            return consumer_cls(*args, **kwargs)
        return _KafkaConsumer(*args, **kwargs)

    @classmethod
    def init_worker(cls, *args, **kwargs):
        """``worker_init_fn`` factory: one consumer per DataLoader worker
        (reference :208-233)."""

        def func(worker_id):
            info = get_worker_info()
            if info is None:
                raise RuntimeError(
                    "Custom initialization should be used for multiprocessing only."
                )
            dataset = info.dataset
            # pylint: disable=protected-access
            dataset._consumer = cls.new_consumer(*args, **kwargs)
            dataset._worker_id = worker_id

        return func

    @classmethod
    def commit_worker(cls, worker: mp.Process):
        """Ask a worker process to commit (reference :235-239)."""
        os.kill(worker.pid, cls._COMMIT_SIGNAL)

    @classmethod
    def placeholder(cls):
        """Dataset without a consumer, for the multi-worker DataLoader
        (reference :241-247)."""
        return cls(_is_placeholder=True)
