"""SlotEngine: object wrapper over the C-ABI for one GPU.

It owns a native context (pinned staging slots + device twins), knows how to run
the kernel sequence of a schema on a slot and how to turn the fetched result
block into batch views.  ``IngestPipeline`` (pipeline.py) drives it from a Kafka
consumer; tests and ``bench.py`` drive it directly with packed records.

Kernel sequence per slot (all on the caller's current CUDA stream):

    tk_slot_wait        wait for the slot's H2D copy (K0)
    tk_carry_rows       rows of the open batch: previous tensor -> head of this one
    tk_json_scan | tk_raw_scan      element count / size check, keep flag (K1 phase 1)
    tk_compact          filter compaction, batch boundaries (K2)
    tk_commit_offsets   per batch, per partition commit offsets (K4)
    tk_json_decode | tk_json_object_decode | tk_collate_raw | tk_collate_field |
    tk_collate_image_u8_bf16_nchw   (K1 phase 2 + K3)
    tk_slot_fetch_result            async D2H of the small result block
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional

import numpy as np
import torch

from . import _native
from .schema import ImageU8, JsonArray, JsonNested, JsonObject, PayloadError, RawBytes, StructBytes

# one packed record header: tk_rec_t in include/torchkafka_b200.h
REC_DTYPE = np.dtype([("start", "<u4"), ("partition", "<i4"), ("offset", "<i8")])

#: a slot never closes more batches than this (bounds the commit matrix)
MAX_BATCHES_CAP = 16384
#: default budget for the output tensors of one slot (rows of every plane)
DEFAULT_OUTPUT_BYTES = 2 << 30
#: kernels launched per slot by ``process``: carry rows, scan, plan (compaction + commit maxima,
#: one kernel), decode / collate (one kernel), commit running max (one block, beside the decode);
#: JSON arrays through the single-pass kernel: carry rows, tile table + resets, the single-pass
#: kernel, commit running max
LAUNCHES_PER_SLOT_JSON = 5
LAUNCHES_PER_SLOT_JSON_FUSED = 4
LAUNCHES_PER_SLOT_RAW = 5


def _np_view(ptr: int, nbytes: int, dtype) -> np.ndarray:
    buf = (C.c_uint8 * nbytes).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype)


class SlotOutput:
    """Tensors one slot wrote into (kept alive until the next slot took its carry)."""

    __slots__ = ("tensors", "planes", "n_records", "done")

    def __init__(self, tensors, planes, n_records):
        self.done = None
        self.tensors = tensors
        self.planes = planes
        self.n_records = n_records


class SlotResult:
    """Host copy of a slot's result block."""

    __slots__ = ("n_records", "n_kept", "carry_in", "carry_out", "n_closed", "n_errors",
                 "first_error_rec", "batch_end", "batch_lmax", "commit")

    def __init__(self, hdr, batch_end, batch_lmax, commit):
        self.n_records = hdr.n_records
        self.n_kept = hdr.n_kept
        self.carry_in = hdr.carry_in
        self.carry_out = hdr.carry_out
        self.n_closed = hdr.n_closed
        self.n_errors = hdr.n_errors
        self.first_error_rec = hdr.first_error_rec
        self.batch_end = batch_end
        self.batch_lmax = batch_lmax
        self.commit = commit          # [n_partitions, n_closed + 1] int64


class SlotEngine:
    #: key of the engine pool entry this engine belongs to (``acquire`` sets it)
    pool_key = None

    def __init__(self, schema, batch_size: int, slot_bytes: int = 64 << 20,
                 slot_records: Optional[int] = None, n_slots: int = 3,
                 device: Optional[int] = None, n_partitions: int = 64,
                 output_bytes: Optional[int] = None):
        if not torch.cuda.is_available():
            raise RuntimeError(
                "payload_schema datasets need a CUDA device (B200); there is no CPU fallback."
            )
        self.lib = _native.load()
        self.schema = schema
        self.batch_size = int(batch_size)
        if self.batch_size <= 0:
            raise ValueError("batch_size must be positive")
        self.device = torch.cuda.current_device() if device is None else int(device)
        #: kernels run on the engine's own stream so that a consumer synchronising on a
        #: delivered batch never waits for the slot that is being decoded behind it
        self.stream = torch.cuda.Stream(device=self.device)
        self.n_slots = int(n_slots)
        self.n_partitions = int(n_partitions)
        self.slot_bytes = int(slot_bytes)
        if isinstance(schema, (RawBytes, ImageU8, StructBytes)):
            rb = schema.record_bytes
            if self.slot_bytes < rb:
                self.slot_bytes = rb
            auto_records = max(1, self.slot_bytes // rb)
        else:
            auto_records = max(1, self.slot_bytes // 48)
        cap = MAX_BATCHES_CAP * self.batch_size
        # Output tensors hold one row per record of the slot (kept or not) plus batch_size head
        # rows, each as wide as the schema allows: short records under a wide schema
        # (TokenList(max_len=2048) over ten-token records) would otherwise ask for gigabytes per
        # slot.  The slot is therefore also cut by an output-byte budget.
        self.output_bytes = int(output_bytes) if output_bytes else DEFAULT_OUTPUT_BYTES
        by_output = max(1, self.output_bytes // max(self.row_bytes(schema), 1) - self.batch_size)
        self.slot_records = int(min(slot_records or auto_records, cap, 1 << 22, by_output))
        self.max_batches = (self.slot_records + self.batch_size - 1) // self.batch_size + 1
        ctx = C.c_void_p()
        _native.check(self.lib.tk_ctx_create(self.device, self.slot_bytes, self.slot_records,
                                              self.n_slots, self.max_batches, self.n_partitions,
                                              C.byref(ctx)))
        self.ctx = ctx
        self._host_bytes: List[np.ndarray] = []
        self._host_recs: List[np.ndarray] = []
        for s in range(self.n_slots):
            pb, pr = C.c_void_p(), C.c_void_p()
            _native.check(self.lib.tk_slot_host_ptrs(self.ctx, s, C.byref(pb), C.byref(pr)))
            self._host_bytes.append(_np_view(pb.value, self.slot_bytes, np.uint8))
            self._host_recs.append(_np_view(pr.value, (self.slot_records + 1) * 16, REC_DTYPE))
        if isinstance(schema, ImageU8):
            mean = (C.c_float * 4)(*schema.mean)
            std = (C.c_float * 4)(*schema.std)
            with torch.cuda.device(self.device):
                _native.check(self.lib.tk_image_set_norm(self.ctx, schema.channels, mean, std,
                                                         self._stream()))
        if isinstance(schema, JsonObject):
            names, offs = schema.name_table()
            blob = np.frombuffer(names, dtype=np.uint8).copy()
            off = np.asarray(offs, dtype=np.uint32)
            with torch.cuda.device(self.device):
                _native.check(self.lib.tk_ctx_set_object_fields(self.ctx, blob.ctypes.data, off.ctypes.data,
                                                                len(schema.fields), self._stream()))
        self.launches = 0
        #: JSON arrays through the single-pass kernel (tk_json_fused.cu) instead of the three-kernel
        #: chain (scan, plan, parse): an independent decomposition of the same contract, used as a
        #: cross-check and for comparison runs (``set_json_fused``)
        self.json_fused = bool(int(os.environ.get("TK_JSON_FUSED", "0") or 0))
        #: variant of the JSON parse kernel (csrc/tk_json_kernels.cu): None = chosen from a sample of
        #: the payload (``_choose_json_variant``), True / False = fixed (TK_JSON_WIDE=1 / 0)
        _w = os.environ.get("TK_JSON_WIDE")
        self.json_wide_fixed = None if _w in (None, "") else bool(int(_w))
        self.json_wide = bool(self.json_wide_fixed)
        self._variant_checks = 0
        self._slot_payload = {}              # slot -> payload bytes in its pinned buffer (not for raw submissions)
        try:
            cores = len(os.sched_getaffinity(0))
        except AttributeError:  # pragma: no cover
            cores = os.cpu_count() or 1
        # several ranks on one host (torchrun) share its cores
        cores = max(1, cores // max(1, int(os.environ.get("LOCAL_WORLD_SIZE", "1"))))
        #: threads used by the pinned-memory pack (payload memcpy).  Packs of different producer
        #: threads take turns on one process-wide pool, so each pack gets all of them; two cores
        #: are left for the producer (GIL) and the consuming thread.  Measured on the 16-core
        #: B200 host: 7 threads 61 GB/s, 14 threads 92 GB/s (benchmarks/pack_bandwidth.py).
        self.pack_threads = int(os.environ.get("TK_PACK_THREADS", 0)) or max(1, min(16, cores - 2))

    @staticmethod
    def row_bytes(sc) -> int:
        """Bytes one record's row takes in the output tensors (all planes)."""
        if isinstance(sc, JsonArray):
            width = sc.max_len * torch.empty((), dtype=sc.dtype).element_size()
            return width + (4 + sc.max_len if sc.fixed_len is None else 0)
        if isinstance(sc, JsonObject):
            return len(sc.fields) * torch.empty((), dtype=sc.dtype).element_size()
        if isinstance(sc, JsonNested):
            return sc.elements * torch.empty((), dtype=sc.dtype).element_size()
        if isinstance(sc, ImageU8):
            return sc.record_bytes * 2
        if isinstance(sc, (RawBytes, StructBytes)):
            return sc.record_bytes
        raise TypeError(f"unsupported schema {sc!r}")

    # ------------------------------------------------------------------ plumbing
    def _stream(self):
        st = getattr(self, "stream", None)
        if st is None:
            return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        return C.c_void_p(st.cuda_stream)

    def close(self):
        if getattr(self, "ctx", None) is not None and self.ctx.value:
            self._host_bytes = []
            self._host_recs = []
            self.lib.tk_ctx_destroy(self.ctx)
            self.ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def reset(self):
        _native.check(self.lib.tk_ctx_reset(self.ctx, self._stream()))

    def h2d_peak_gbs(self, nbytes: int = 256 << 20, iters: int = 10) -> float:
        out = C.c_double()
        _native.check(self.lib.tk_h2d_peak_gbs(self.device, nbytes, iters, C.byref(out)))
        return out.value

    # ------------------------------------------------------------------ filling
    def host_views(self, slot: int):
        """(uint8 bytes view, structured recs view) over the slot's pinned memory."""
        return self._host_bytes[slot], self._host_recs[slot]

    def fill_packed(self, slot: int, values: np.ndarray, starts: np.ndarray,
                    partitions: np.ndarray, offsets: np.ndarray):
        """Copy a packed run of records (CSR arrays) into the slot's pinned memory
        (tk_slot_fill: headers in C, payload memcpy on several threads, GIL released)."""
        n = len(partitions)
        nbytes = int(starts[n]) - int(starts[0])
        if n > self.slot_records or nbytes > self.slot_bytes:
            raise ValueError(f"run of {n} records / {nbytes} bytes exceeds the slot "
                             f"({self.slot_records} records / {self.slot_bytes} bytes)")
        values = np.ascontiguousarray(values, dtype=np.uint8)
        starts = np.ascontiguousarray(starts, dtype=np.uint32)
        partitions = np.ascontiguousarray(partitions, dtype=np.int32)
        offsets = np.ascontiguousarray(offsets, dtype=np.int64)
        if len(values) != nbytes and int(starts[0]) == 0:
            values = values[:nbytes]
        after = C.c_uint64()
        _native.check(self.lib.tk_slot_fill(self.ctx, slot, 0, 0, values.ctypes.data, starts.ctypes.data,
                                            partitions.ctypes.data, offsets.ctypes.data, n,
                                            self.pack_threads, C.byref(after)))
        return n, nbytes

    def pack_pointers(self, slot: int, ptrs: np.ndarray, lens: np.ndarray, partitions: np.ndarray,
                      offsets: np.ndarray, n: int) -> int:
        """Gather n values given by (address, length) into the slot (tk_slot_pack)."""
        after = C.c_uint64()
        _native.check(self.lib.tk_slot_pack(self.ctx, slot, 0, 0, ptrs.ctypes.data, lens.ctypes.data,
                                            partitions.ctypes.data, offsets.ctypes.data, n,
                                            self.pack_threads, C.byref(after)))
        return int(after.value)

    def submit(self, slot: int, n: int, nbytes: int):
        self._slot_payload[slot] = nbytes
        _native.check(self.lib.tk_slot_submit(self.ctx, slot, n, nbytes))

    def set_json_wide(self, enabled: bool):
        """Parse kernel with the ``0.00...`` literals inline (True) or the leanest element loop (False)."""
        self.json_wide = bool(enabled)
        _native.check(self.lib.tk_ctx_set_json_wide(self.ctx, 1 if enabled else 0))

    def _choose_json_variant(self, slot: int):
        """Every 32nd slot: how many numbers of a 64 KB sample of the payload start with ``0.0`` or
        carry an exponent?  The lean parse kernel leaves those to an out-of-line call (one element in a
        hundred of them costs 15 %); above one in 300 the wide kernel runs, below one in 3 000 the lean
        one.  Slots submitted as record batches hold no plain payload on the host: wide."""
        if self.json_wide_fixed is not None or not isinstance(self.schema, JsonArray):
            return
        self._variant_checks += 1
        if (self._variant_checks - 1) % 32:
            return
        nbytes = self._slot_payload.get(slot)
        if nbytes is None:
            want = True
        else:
            lo = max(0, nbytes // 2 - 32768)
            sample = self._host_bytes[slot][lo:lo + 65536].tobytes()
            elements = sample.count(b",") + 1
            odd = (sample.count(b" 0.0") + sample.count(b"-0.0") + sample.count(b",0.0") + sample.count(b"[0.0")
                   + sample.count(b"e-") + sample.count(b"e+"))
            want = True if odd * 300 > elements else False if odd * 3000 < elements else self.json_wide
        if want != self.json_wide:
            self.set_json_wide(want)

    # ------------------------------------------------------------------ kernels
    def _allocate(self, n: int) -> SlotOutput:
        """Output tensors for a slot of ``n`` records: batch_size head rows + n body rows."""
        rows = self.batch_size + n
        dev = torch.device("cuda", self.device)
        sc = self.schema
        if isinstance(sc, JsonArray):
            values = torch.empty((rows, sc.max_len), dtype=sc.dtype, device=dev)
            tensors = [values]
            planes = [(values.data_ptr(), sc.max_len * values.element_size())]
            if sc.fixed_len is None:
                lengths = torch.empty((rows,), dtype=torch.int32, device=dev)
                mask = torch.empty((rows, sc.max_len), dtype=torch.bool, device=dev)
                tensors += [lengths, mask]
                planes += [(lengths.data_ptr(), 4), (mask.data_ptr(), sc.max_len)]
        elif isinstance(sc, JsonObject):
            values = torch.empty((rows, len(sc.fields)), dtype=sc.dtype, device=dev)
            tensors = [values]
            planes = [(values.data_ptr(), len(sc.fields) * values.element_size())]
        elif isinstance(sc, JsonNested):
            values = torch.empty((rows,) + sc.shape, dtype=sc.dtype, device=dev)
            tensors = [values]
            planes = [(values.data_ptr(), sc.elements * values.element_size())]
        elif isinstance(sc, ImageU8):
            out = torch.empty((rows, sc.channels, sc.height, sc.width), dtype=torch.bfloat16, device=dev)
            tensors = [out]
            planes = [(out.data_ptr(), sc.record_bytes * 2)]
        elif isinstance(sc, RawBytes):
            out = torch.empty((rows,) + sc.shape, dtype=sc.dtype, device=dev)
            tensors = [out]
            planes = [(out.data_ptr(), sc.record_bytes)]
        elif isinstance(sc, StructBytes):
            tensors = [torch.empty((rows,) + sh, dtype=d, device=dev) for _, d, sh in sc.fields]
            planes = [(t.data_ptr(), nb) for t, nb in zip(tensors, sc.field_bytes)]
        else:
            raise TypeError(f"unsupported schema {sc!r}")
        arr = (_native.TkPlane * len(planes))()
        for k, (ptr, rb) in enumerate(planes):
            arr[k].base = ptr
            arr[k].row_bytes = rb
        return SlotOutput(tensors, arr, n)

    def _job(self, out: SlotOutput, prev_slot, prev_out, resident: bool) -> "_native.TkJob":
        sc = self.schema
        job = _native.TkJob()
        if isinstance(sc, JsonArray):
            job.schema, job.dtype = 0, sc.kind
            job.max_len, job.min_size, job.fixed_len = sc.max_len, sc.min_size, sc.fixed_len or 0
        elif isinstance(sc, JsonObject):
            job.schema, job.dtype = 4, sc.kind
        elif isinstance(sc, JsonNested):
            job.schema, job.dtype, job.max_len = 5, sc.kind, len(sc.shape)
            for k, extent in enumerate(sc.shape):
                job.field_bytes[k] = extent
        elif isinstance(sc, ImageU8):
            job.schema = 2
            job.height, job.width, job.channels = sc.height, sc.width, sc.channels
            job.record_bytes = sc.record_bytes
        elif isinstance(sc, StructBytes):
            job.schema = 3
            job.record_bytes = sc.record_bytes
            for k, (off, nb) in enumerate(zip(sc.field_offsets, sc.field_bytes)):
                job.field_offset[k], job.field_bytes[k] = off, nb
        else:
            job.schema = 1
            job.record_bytes = sc.record_bytes
        job.batch_size = self.batch_size
        job.prev_slot = -1 if prev_out is None else prev_slot
        job.n_planes = len(out.planes)
        job.resident = 1 if resident else 0
        for k in range(len(out.planes)):
            job.cur[k] = out.planes[k]
            if prev_out is not None:
                job.prev[k] = prev_out.planes[k]
        return job

    def process(self, slot: int, n: int, prev_slot: Optional[int], prev_out: Optional[SlotOutput],
                resident: bool = False, timers: Optional[list] = None) -> SlotOutput:
        """Enqueue the whole kernel sequence for ``slot`` on the engine's stream (asynchronous).

        Normally a single C call (``tk_process_slot``).  With ``timers`` the sequence is issued
        kernel by kernel and a (start, end) pair of CUDA events bracketing the decode / collate
        kernel is appended (roofline measurement in bench.py)."""
        lib, ctx, st = self.lib, self.ctx, self._stream()
        sc = self.schema
        is_json = isinstance(sc, JsonArray)
        if is_json:
            self._choose_json_variant(slot)
        with torch.cuda.device(self.device), torch.cuda.stream(self.stream):
            out = self._allocate(n)
            if timers is None:
                job = self._job(out, prev_slot, prev_out, resident)
                _native.check(lib.tk_process_slot(ctx, slot, C.byref(job), st))
            else:
                if not resident:
                    _native.check(lib.tk_slot_wait(ctx, slot, st))
                if prev_out is not None:
                    _native.check(lib.tk_carry_rows(ctx, prev_slot, len(out.planes), prev_out.planes,
                                                    out.planes, st))
                if is_json:
                    _native.check(lib.tk_json_scan(ctx, slot, sc.max_len, sc.min_size, st))
                elif isinstance(sc, (JsonObject, JsonNested)):
                    _native.check(lib.tk_raw_scan(ctx, slot, 0xFFFFFFFFFFFFFFFF, st))     # TK_ANY_RECORD_BYTES
                else:
                    _native.check(lib.tk_raw_scan(ctx, slot, sc.record_bytes, st))
                _native.check(lib.tk_compact(ctx, slot, self.batch_size, st))
                _native.check(lib.tk_commit_offsets(ctx, slot, st))
                ev0 = torch.cuda.Event(enable_timing=True)
                ev1 = torch.cuda.Event(enable_timing=True)
                ev0.record()
                P = out.planes
                if is_json:
                    variable = sc.fixed_len is None
                    _native.check(lib.tk_json_decode(
                        ctx, slot, sc.kind, sc.max_len, sc.fixed_len or 0, C.byref(P[0]),
                        C.byref(P[1]) if variable else None, C.byref(P[2]) if variable else None, st))
                elif isinstance(sc, JsonObject):
                    _native.check(lib.tk_json_object_decode(ctx, slot, sc.kind, C.byref(P[0]), st))
                elif isinstance(sc, JsonNested):
                    dims = (C.c_uint32 * len(sc.shape))(*sc.shape)
                    _native.check(lib.tk_json_nested_decode(ctx, slot, sc.kind, len(sc.shape), dims, C.byref(P[0]), st))
                elif isinstance(sc, ImageU8):
                    _native.check(lib.tk_collate_image_u8_bf16_nchw(ctx, slot, sc.height, sc.width,
                                                                    sc.channels, C.byref(P[0]), st))
                elif isinstance(sc, StructBytes):
                    for k, (off, nb) in enumerate(zip(sc.field_offsets, sc.field_bytes)):
                        _native.check(lib.tk_collate_field(ctx, slot, sc.record_bytes, off, nb, C.byref(P[k]), st))
                else:
                    _native.check(lib.tk_collate_raw(ctx, slot, sc.record_bytes, C.byref(P[0]), st))
                ev1.record()
                if timers is not None:
                    timers.append((ev0, ev1))
                _native.check(lib.tk_slot_fetch_result(ctx, slot, st))
            per_slot = LAUNCHES_PER_SLOT_RAW if not is_json else (
                LAUNCHES_PER_SLOT_JSON_FUSED if (self.json_fused and timers is None) else LAUNCHES_PER_SLOT_JSON)
            self.launches += per_slot \
                - (1 if prev_out is None else 0) + (len(out.planes) - 1 if isinstance(sc, StructBytes) else 0)
            out.done = torch.cuda.Event()
            out.done.record(self.stream)
        return out

    def set_json_fused(self, enabled: bool):
        """JSON arrays through the single-pass kernel (True) or the three-kernel chain (False)."""
        self.json_fused = bool(enabled)
        _native.check(self.lib.tk_ctx_set_json_fused(self.ctx, 1 if enabled else 0))

    def set_timing(self, enabled: bool):
        """Bracket the decode / collate kernels of every following ``process`` call with CUDA
        events inside the library (``decode_ms`` reads them)."""
        _native.check(self.lib.tk_ctx_set_timing(self.ctx, 1 if enabled else 0))

    def decode_ms(self, slot: int) -> float:
        ms = C.c_float()
        _native.check(self.lib.tk_slot_decode_ms(self.ctx, slot, C.byref(ms)))
        return float(ms.value)

    def publish(self, out: "SlotOutput"):
        """Make the slot's tensors safe to use on the caller's current stream."""
        cur = torch.cuda.current_stream(self.device)
        if cur != self.stream:
            cur.wait_event(out.done)
            for t in out.tensors:
                t.record_stream(cur)

    def assume_resident(self, slot: int, n: int, nbytes: int):
        self._slot_payload[slot] = nbytes
        _native.check(self.lib.tk_slot_assume_resident(self.ctx, slot, n, nbytes))

    def device_views(self, slot: int):
        """Raw device pointers of the slot's twin (bytes, recs), for benchmarks that preload HBM."""
        pb, pr = C.c_void_p(), C.c_void_p()
        _native.check(self.lib.tk_slot_device_ptrs(self.ctx, slot, C.byref(pb), C.byref(pr)))
        return pb.value, pr.value

    # ------------------------------------------------------------------ results
    def result(self, slot: int) -> SlotResult:
        """Block until the slot's result block is on the host, return a copy."""
        hdr = C.POINTER(_native.TkResult)()
        be = C.POINTER(C.c_uint32)()
        bl = C.POINTER(C.c_uint32)()
        cm = C.POINTER(C.c_int64)()
        pitch = C.c_uint32()
        _native.check(self.lib.tk_slot_result(self.ctx, slot, C.byref(hdr), C.byref(be), C.byref(bl),
                                              C.byref(cm), C.byref(pitch)))
        h = hdr.contents
        nb = h.n_closed + 1
        batch_end = np.ctypeslib.as_array(be, shape=(nb,)).copy()
        batch_lmax = np.ctypeslib.as_array(bl, shape=(nb,)).copy()
        commit = np.ctypeslib.as_array(cm, shape=(self.n_partitions, pitch.value))[:, :nb].copy()
        return SlotResult(h, batch_end, batch_lmax, commit)

    def record_debug(self, slot: int, n: int):
        """(rec_len, rec_status, rec_kidx) of the last run on ``slot`` (synchronises the device)."""
        ln = np.empty(n, np.int32)
        st = np.empty(n, np.int32)
        kx = np.empty(n, np.int32)
        _native.check(self.lib.tk_slot_debug_records(self.ctx, slot, ln.ctypes.data, st.ctypes.data,
                                                     kx.ctypes.data, n))
        return ln, st, kx

    def batch_view(self, out: SlotOutput, first_row: int, n_rows: int, lmax: int):
        """Batch object over rows [first_row, first_row + n_rows) of a slot's tensors."""
        sc = self.schema
        t = out.tensors
        if isinstance(sc, JsonArray):
            if sc.fixed_len is not None:
                return t[0][first_row:first_row + n_rows]
            width = sc.max_len if sc.pad_to_max_len else max(int(lmax), 0)
            if width == sc.max_len:
                return sc.wrap(t[0][first_row:first_row + n_rows], t[1][first_row:first_row + n_rows],
                               t[2][first_row:first_row + n_rows])
            return sc.wrap(t[0][first_row:first_row + n_rows, :width],
                           t[1][first_row:first_row + n_rows],
                           t[2][first_row:first_row + n_rows, :width])
        if isinstance(sc, StructBytes):
            return sc.wrap(*[p[first_row:first_row + n_rows] for p in t])
        return t[0][first_row:first_row + n_rows]

    def batch_views(self, out: SlotOutput, first_row: int, n_batches: int, lmax):
        """All ``n_batches`` full batches of a slot at once (one unflatten + unbind per plane
        instead of one slice per batch and plane)."""
        B = self.batch_size
        sc = self.schema
        planes = [p[first_row:first_row + n_batches * B].unflatten(0, (n_batches, B)).unbind(0)
                  for p in out.tensors]
        if isinstance(sc, StructBytes):
            return [sc.wrap(*fields) for fields in zip(*planes)]
        if not (isinstance(sc, JsonArray) and sc.fixed_len is None):
            return list(planes[0])
        batches = []
        for j in range(n_batches):
            v, ln, m = planes[0][j], planes[1][j], planes[2][j]
            width = sc.max_len if sc.pad_to_max_len else int(lmax[j])
            if width != sc.max_len:
                v, m = v[:, :width], m[:, :width]
            batches.append(sc.wrap(v, ln, m))
        return batches

    def error_for(self, slot: int, res: SlotResult, partitions=None, offsets=None) -> PayloadError:
        """The error of the slot's first bad record.  ``partitions`` / ``offsets`` are the host's
        copy of the record coordinates; a slot split on the device (record batches) has none and
        the record's header is read back instead."""
        e = int(res.first_error_rec)
        _, status, _ = self.record_debug(slot, res.n_records)
        code = int(status[e])
        if partitions is None:
            rec = _native.TkRec()
            _native.check(self.lib.tk_slot_record(self.ctx, slot, e, C.byref(rec)))
            part, off = int(rec.partition) & ~_native.REC_FLAGS, int(rec.offset)
        else:
            part, off = int(partitions[e]), int(offsets[e])
        return PayloadError(
            f"record {e} of the slot (partition index {part}, offset {off}): "
            f"{_native.STATUS_NAMES.get(code, code)}",
            partition=part, offset=off, status=code)

    # ------------------------------------------------------------------ record batches
    def submit_recordbatches(self, slot: int, n_bytes: int, blobs, host_ptrs=None) -> int:
        """Raw RecordBatch v2 bytes are in the slot's pinned buffer ``[0, n_bytes)``; ``blobs`` is a
        sequence of (start, length, dense partition, fetch_offset, offset_delta).  Returns the
        number of record entries of the slot (known from the batch headers); the copy to the
        device and the split kernels are asynchronous.  With ``host_ptrs`` (one address per blob)
        the bytes are not in the slot's buffer but wherever the consumer keeps them, and go to the
        device straight from there (``tk_slot_submit_recordbatches_ext``)."""
        self._slot_payload.pop(slot, None)
        arr = (_native.TkBlob * max(len(blobs), 1))()
        for k, (start, length, part, fetch_offset, delta) in enumerate(blobs):
            b = arr[k]
            b.start, b.length, b.partition = int(start), int(length), int(part)
            b.fetch_offset, b.offset_delta = int(fetch_offset), int(delta)
        n = C.c_uint32()
        if host_ptrs is None:
            _native.check(self.lib.tk_slot_submit_recordbatches(self.ctx, slot, C.c_uint64(int(n_bytes)), arr,
                                                                len(blobs), C.byref(n)))
        else:
            ptrs = (C.c_void_p * max(len(blobs), 1))(*[int(p) for p in host_ptrs])
            _native.check(self.lib.tk_slot_submit_recordbatches_ext(self.ctx, slot, C.c_uint64(int(n_bytes)), arr,
                                                                    ptrs, len(blobs), C.byref(n)))
        return int(n.value)

    def recordbatch_result(self, slot: int):
        """Wait for the split of ``slot``; returns (n_records, n_value_bytes, n_batches)."""
        r = _native.TkRbSummary()
        _native.check(self.lib.tk_slot_recordbatch_result(self.ctx, slot, C.byref(r)))
        return int(r.n_records), int(r.n_bytes), int(r.n_batches)


# ---------------------------------------------------------------------- engine pool
# Creating an engine pins host memory and allocates the device twins of every slot (hundreds of
# megabytes, ~0.2 s); a dataset that is iterated again -- a new epoch, a re-opened pipeline after
# a `break` -- would pay that each time.  Pipelines therefore take their engine from a small
# pool and hand it back when they close.  An engine in the pool is idle (its stream is
# synchronised) and reset (no open batch, no partition positions).
_POOL: "collections.OrderedDict" = None
_POOL_LOCK = None
#: engines kept for reuse (least recently released ones are destroyed first)
POOL_CAPACITY = 2


def _pool():
    global _POOL, _POOL_LOCK
    if _POOL is None:
        import atexit
        import collections
        import threading
        _POOL = collections.OrderedDict()
        _POOL_LOCK = threading.Lock()
        atexit.register(clear_pool)          # before the CUDA runtime shuts down
    return _POOL


def acquire(schema, batch_size: int, slot_bytes: int, n_slots: int, device: Optional[int],
            n_partitions: int, output_bytes: Optional[int] = None) -> SlotEngine:
    """An engine of this geometry: a pooled one if there is one, else a new one."""
    dev = torch.cuda.current_device() if (device is None and torch.cuda.is_available()) else device
    key = (dev, schema, int(batch_size), int(slot_bytes), int(n_slots), int(n_partitions), output_bytes)
    pool = _pool()
    try:
        with _POOL_LOCK:
            eng = pool.pop(key, None)
    except TypeError:                        # unhashable schema: no pooling
        eng, key = None, None
    if eng is None:
        eng = SlotEngine(schema, batch_size, slot_bytes=slot_bytes, n_slots=n_slots, device=device,
                         n_partitions=n_partitions, output_bytes=output_bytes)
    eng.pool_key = key
    return eng


def release(eng: SlotEngine):
    """Hand an engine back (or destroy it when it cannot be pooled)."""
    if getattr(eng, "ctx", None) is None or not eng.ctx.value:
        return
    key = eng.pool_key
    if key is None or POOL_CAPACITY <= 0:
        eng.close()
        return
    try:
        eng.stream.synchronize()
        eng.set_timing(False)
        eng.set_json_fused(bool(int(os.environ.get("TK_JSON_FUSED", "0") or 0)))
        eng.reset()
    except Exception:
        eng.close()
        return
    pool = _pool()
    evicted = []
    with _POOL_LOCK:
        old = pool.pop(key, None)
        if old is not None:
            evicted.append(old)
        pool[key] = eng
        while len(pool) > POOL_CAPACITY:
            evicted.append(pool.popitem(last=False)[1])
    for e in evicted:
        e.close()


def clear_pool():
    if _POOL is None:
        return
    try:
        _clear_arenas()
    except Exception:
        pass
    with _POOL_LOCK:
        engines = list(_POOL.values())
        _POOL.clear()
    for e in engines:
        try:
            e.close()
        except Exception:
            pass


# ---------------------------------------------------------------------- page-locked arenas
# A consumer that keeps fetched bytes in long-lived buffers of its own can have them page-locked
# once (cudaHostRegister); slots are then filled by DMA straight out of them.  The registry holds a
# reference to every registered buffer (it must not be freed while registered) and forgets the
# least recently used ones.
_ARENAS: "collections.OrderedDict" = None
ARENA_CAPACITY = 8


def register_arena(buf) -> bool:
    """Page-lock ``buf`` (anything with the buffer protocol / ``__array_interface__``).  Returns
    False when the driver refuses (the copies then go through the driver's staging buffer)."""
    global _ARENAS
    import collections
    arr = np.asarray(buf)
    if arr.nbytes == 0:
        return False
    addr = arr.ctypes.data
    _pool()                                  # installs the atexit hook
    if _ARENAS is None:
        _ARENAS = collections.OrderedDict()
    with _POOL_LOCK:
        if addr in _ARENAS:
            _ARENAS.move_to_end(addr)
            return True
    lib = _native.load()
    if lib.tk_host_register(C.c_void_p(addr), C.c_uint64(arr.nbytes)) != 0:
        return False
    drop = []
    with _POOL_LOCK:
        _ARENAS[addr] = arr
        while len(_ARENAS) > ARENA_CAPACITY:
            drop.append(_ARENAS.popitem(last=False))
    for a, _keep in drop:
        lib.tk_host_unregister(C.c_void_p(a))
    return True


def _clear_arenas():
    if not _ARENAS:
        return
    lib = _native.load()
    with _POOL_LOCK:
        items = list(_ARENAS.items())
        _ARENAS.clear()
    for a, _keep in items:
        lib.tk_host_unregister(C.c_void_p(a))
