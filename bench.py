#!/usr/bin/env python3
"""bench.py -- records/s of the ingest hot path on BASELINE.json config 2.

Workload ("cfg2"): variable-length JSON float arrays (16..256 elements,
uniform(-1e3, 1e3), json.dumps default separators), min_size=32 filter, float32
output, pad + length + mask collate, batch 1024, 64 partitions, commit offsets per
batch.  One *step* = one batch of 1024 kept records (~1096 consumed records,
~2.9 MB of payload).  With N ranks (torchrun), rank r owns the partitions p % N == r
 and ingests its own log of the same volume: weak scaling, no collective on
the data path.

    value   consumed records/s, slot bytes already resident in HBM, CUDA events on the stream
            the kernels are launched on.  The pipeline's unit of work is a staging slot of
            SLOT_BATCHES batches, so the K steps are repeated R times (R a multiple of
            SLOT_BATCHES / gcd(K, SLOT_BATCHES), and large enough for a >= 0.5 s window) and
            the K * R batches are issued as K * R / SLOT_BATCHES full slot passes; the
            warm-up runs the same passes (same shapes: the caching allocator is warm).
    e2e     the same metric through the public API, auto_commit(DataLoader(dataset)), from
            host ConsumerRecord objects.  The clock starts BEFORE the iteration (and with it
            the dataset's pipeline: poll, pack into pinned memory, H2D, kernels) is started
            and stops after the last batch of the stream was delivered, its per-step result
            (the int32 length vector) copied to the host and everything synchronised, so no
            read-ahead lies outside the window.  The stream is long enough for a window of
            seconds (raw / packed arms) or as long as the pre-built record objects allow
            (poll arm).  A separate warm-up iteration runs first.  link_frac > 1 is an error.
    roofline  dominant kernel (k_json_parse: pad/length/mask rows + thread-per-element parse,
              the one kernel of tk_json_decode) against the measured HBM copy bandwidth
    cpu_baseline  the oracle (reference algorithm: json.loads + torch.tensor + pad collate)
                  timed on this box's host cores on a bounded sample

`--impl reference` times the reference's CPU implementation of the same path
(oracle port; the reference is pure Python over kafka-python, which exists on no
machine here) with a multi-worker DataLoader on all usable host cores.
"""
import argparse
import gc
import json
import math
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

BATCH = 1024
MAX_LEN = 256
MIN_SIZE = 32
SLOT_BATCHES = int(os.environ.get("TK_SLOT_BATCHES", "32"))
PARTITIONS = 64
SEED = 2026
METRIC = "records/sec"
WORKLOAD = "cfg2: JSON float arrays len 16-256, min_size=32, f32 pad+mask collate, batch 1024"
VALUE_MIN_SECONDS = float(os.environ.get("TK_VALUE_SECONDS", "0.5"))
E2E_MIN_SECONDS = float(os.environ.get("TK_E2E_SECONDS", "2.5"))
#: ConsumerRecord objects the poll arm may hold (they exist before the clock starts, as they do
#: for the CPU arm): ~200 bytes each
POLL_MAX_OBJECTS = int(os.environ.get("TK_POLL_OBJECTS", str(12_000_000)))


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def shared_config(world):
    """`config` of the JSON line: the same for the GPU arm and the reference arm."""
    return {"workload": WORKLOAD, "batch": BATCH, "partitions": PARTITIONS,
            "sharding": f"rank r of {world} owns the partitions p % {world} == r; per-rank volume fixed (weak)",
            "l2": "device-resident leg: 4 distinct ~94 MB slots cycled (> 126 MB L2), inputs come from HBM; "
                  "end-to-end leg: every slot arrives over the link"}


# --------------------------------------------------------------------------- workload
def build_base_log(n_batches: int, rank: int = 0, world: int = 1):
    """Seeded cfg2 records of one rank, cut so that exactly n_batches * BATCH of them pass the
    filter.  Rank r's records lie in its own partitions (p % world == r)."""
    import numpy as np
    from torchkafka_b200.testing import synthetic as syn
    from torchkafka_b200.testing.stub_kafka import RecordLog
    want = n_batches * BATCH
    vals = syn.cfg2_values(int(want / 0.93) + 64, seed=SEED + rank)
    kept = 0
    cut = None
    for i, v in enumerate(vals):
        if v.count(b",") + 1 >= MIN_SIZE:
            kept += 1
            if kept == want:
                cut = i + 1
                break
    assert cut is not None
    vals = vals[:cut]
    i = np.arange(cut)
    owned = np.asarray([p for p in range(PARTITIONS) if p % world == rank], dtype=np.int32)
    # Kafka-like arrival: runs of 500 records per partition
    blk = i // 500
    part = owned[blk % len(owned)]
    off = (blk // len(owned)) * 500 + (i % 500)
    return RecordLog(vals, part, off)


class Clocks:
    """Samples nvidia-smi clocks / throttle reasons while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.samples = []
        self._stop = threading.Event()
        self._thr = None

    def _loop(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(
                    ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                     "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                parts = [p.strip() for p in out.strip().split(",")]
                if len(parts) >= 7:
                    self.samples.append(parts)
            except Exception:
                pass
            self._stop.wait(0.1)

    def __enter__(self):
        self._thr = threading.Thread(target=self._loop, daemon=True)
        self._thr.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        self._thr.join(timeout=6)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm = [float(s[0]) for s in self.samples if s[0].replace(".", "").isdigit()]
        mx = [float(s[1]) for s in self.samples if s[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[k] for s in self.samples for k in range(4) if s[3 + k].lower() == "active"})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.samples)}


# --------------------------------------------------------------------------- GPU arm
def slot_arrays(log, a, b):
    """CSR arrays of records [a, b) of a RecordLog."""
    import numpy as np
    base = log.starts[a]
    return (log.values[base:log.starts[b]], (log.starts[a:b + 1] - base).astype(np.uint32),
            log.partitions[a:b], log.offsets[a:b])


def run_device_resident(base_log, steps, warmup, device):
    """`value`: slots already in HBM, timed with CUDA events on the launching stream."""
    import torch
    from torchkafka_b200.engine import SlotEngine
    from torchkafka_b200.schema import JsonArray
    schema = JsonArray(torch.float32, max_len=MAX_LEN, min_size=MIN_SIZE)
    n_full = len(base_log)               # the base log holds exactly SLOT_BATCHES batches
    full = slot_arrays(base_log, 0, n_full)
    n_bytes = int(base_log.starts[n_full])
    slot_bytes = n_bytes + 4096
    n_slots = 4          # 4 slots of distinct addresses, together > 126 MB L2: every pass reads from HBM
    eng = SlotEngine(schema, BATCH, slot_bytes=slot_bytes, slot_records=n_full, n_slots=n_slots,
                     device=device, n_partitions=PARTITIONS)
    torch.cuda.set_device(device)
    # preload every slot's device twin once (H2D outside the timed region)
    for s in range(n_slots):
        n, nb = eng.fill_packed(s, *full)
        eng.submit(s, n, nb)
    torch.cuda.synchronize()

    def run(n_passes, dec_ms=None):
        """n_passes slot passes, one tk_process_slot call each, exactly as the pipeline issues
        them.  With dec_ms, the library's event pair around the decode kernel of a slot is read
        right before the slot is reused (n_slots passes later)."""
        prev_slot, prev_out = None, None
        outs = []
        for k in range(n_passes):
            s = k % n_slots
            if dec_ms is not None and k >= n_slots:
                dec_ms.append(eng.decode_ms(s))
            eng.assume_resident(s, n_full, n_bytes)
            out = eng.process(s, n_full, prev_slot, prev_out, resident=True)
            outs.append(out)
            if len(outs) > 2:
                outs.pop(0)
            prev_slot, prev_out = s, out

    def timed(n_passes, dec_ms=None):
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record(eng.stream)            # events on the stream the kernels are launched on
        run(n_passes, dec_ms)
        e1.record(eng.stream)
        torch.cuda.synchronize()
        return e0.elapsed_time(e1)

    # K steps repeated R times = K * R / SLOT_BATCHES full slot passes
    unit = SLOT_BATCHES // math.gcd(steps, SLOT_BATCHES)
    # warm-up: at least `warmup` steps and two rounds over the slots, same pass shape as the timed
    # region (same output sizes: no allocator miss inside the clock)
    warm_passes = max(-(-max(warmup, 3) // SLOT_BATCHES), 2 * n_slots)
    run(warm_passes)
    torch.cuda.synchronize()
    probe_ms = timed(2 * n_slots) / (2 * n_slots)                 # one slot pass
    eng.reset()
    per_rep_ms = probe_ms * steps / SLOT_BATCHES
    R = unit * max(1, math.ceil(VALUE_MIN_SECONDS * 1e3 / per_rep_ms / unit))
    n_passes = steps * R // SLOT_BATCHES

    # Region 1 (`value`): no events inside the kernel chain.
    launches0 = eng.launches
    ms = timed(n_passes)
    launches = eng.launches - launches0
    res = eng.result((n_passes - 1) % n_slots)
    assert res.n_closed == SLOT_BATCHES and res.n_errors == 0 and res.carry_out == 0, \
        (res.n_closed, res.n_errors, res.carry_out)
    # Region 2 (`roofline`): the same passes once more with the library's timing events around the
    # decode kernel of every slot.  The events cut the programmatic-launch chain at both ends of
    # the kernel, so this region runs a little slower than region 1; the kernel's share of the
    # step is taken against this region's own time.
    eng.reset()
    eng.set_timing(True)
    dec_ms = []
    ms_timed = timed(n_passes, dec_ms)
    dec_ms += [eng.decode_ms(k % n_slots) for k in range(max(0, n_passes - n_slots), n_passes)]
    assert len(dec_ms) == n_passes
    eng.set_timing(False)
    # algorithmic HBM bytes = (v + 16) read per consumed record + row written
    # per kept record (MAX_LEN f32 + int32 length); the bool mask plane is extra, not counted.
    algo = n_bytes + 16 * n_full + SLOT_BATCHES * BATCH * (MAX_LEN * 4 + 4)
    info = {
        "ms": ms, "ms_timed_region": ms_timed, "records": n_full * n_passes, "bytes": n_bytes * n_passes,
        "launches": launches, "repeats": R, "passes": n_passes,
        "decode_ms_total": sum(dec_ms), "decode_launches": len(dec_ms),
        "decode_algo_bytes_total": algo * n_passes, "algo_bytes_per_launch": algo,
        "slot_bytes": slot_bytes, "slot_records": n_full,
    }
    info["h2d_peak_gbs"] = eng.h2d_peak_gbs(256 << 20, 5)
    eng.close()
    return info


RAW_BATCH_BYTES = int(os.environ.get("RAW_BATCH_BYTES", str(64 << 10)))   # producer batch.size of the raw arm
E2E_SLOT_BYTES = 48 << 20
E2E_SLOTS = 3


def make_dataset_cls(log, mode, repeat=1):
    """mode: "poll" = poll() objects, "packed" = poll_packed() arrays, "raw" = poll_raw() record
    batches, "raw_into" = poll_raw_into() (record batches received straight into the pinned slot)."""
    import torch
    from torchkafka_b200 import KafkaDataset
    from torchkafka_b200.schema import JsonArray
    from torchkafka_b200.testing import stub_kafka as sk
    consumer_cls = sk.PollOnlyConsumer if mode == "poll" else sk.KafkaConsumer
    input_mode = {"poll": "records", "packed": "packed", "raw": "raw", "raw_into": "raw_into",
                  "raw_dma": "raw_dma"}[mode]

    class Cfg2Dataset(KafkaDataset):
        payload_schema = JsonArray(torch.float32, max_len=MAX_LEN, min_size=MIN_SIZE)
        device_slot_bytes = E2E_SLOT_BYTES
        device_slots = E2E_SLOTS
        device_producers = int(os.environ.get("TK_PRODUCERS", "2"))
        max_partitions = PARTITIONS
        device_input = input_mode

        @classmethod
        def new_consumer(cls, *a, **k):
            k.pop("enable_auto_commit", None)
            return consumer_cls(*a, log=log, repeat=repeat, raw_batch_bytes=RAW_BATCH_BYTES, keep_commit_log=False, **k)

    return Cfg2Dataset


def _iterate(cls, device, host_results):
    """One complete iteration of auto_commit(DataLoader(dataset)) with the per-step D2H.
    Returns (seconds, batches, stats); the clock covers everything from before the iteration is
    created to the final synchronize."""
    import torch
    from torch.utils.data import DataLoader
    from torchkafka_b200 import auto_commit
    cls.device_index = device
    ds = cls("topic", group_id="bench")
    loader = DataLoader(ds, batch_size=BATCH)
    cap = host_results.shape[0]
    torch.cuda.synchronize()
    seen = 0
    t0 = time.perf_counter()
    for batch in auto_commit(loader):
        # every step's result (the batch's int32 length vector) is read back to the host inside
        # the timed region: non-blocking D2H into pinned memory, consumed after the final
        # synchronize -- the way a training loop logs its per-step metric without stalling the GPU
        row = host_results[seen % cap]
        row[:batch.lengths.shape[0]].copy_(batch.lengths, non_blocking=True)
        seen += 1
    torch.cuda.synchronize()
    checksum = int(host_results[:min(seen, cap)].sum())
    t1 = time.perf_counter()
    stats = dict(ds.last_pipeline_stats)
    ds.close()
    return t1 - t0, seen, stats, checksum


def run_e2e(rlog, mode, device, min_seconds, rate_guess, max_records=None):
    """`e2e` of one input protocol.  Warm-up: a complete iteration over the base log (its engine --
    pinned slots, device twins -- goes back to the pool and serves the timed iteration).  Timed:
    a complete iteration over `reps` repetitions of the log, sized from `rate_guess` for a window
    of `min_seconds` and never shorter than 20x the pipeline's read-ahead."""
    import torch
    torch.cuda.set_device(device)
    n = len(rlog)
    slot_records = E2E_SLOT_BYTES / (rlog.nbytes / n)
    readahead = (E2E_SLOTS + 2) * slot_records                    # every slot of the pipeline
    want = max(min_seconds * rate_guess, 20 * readahead)
    if max_records:
        want = min(want, max_records)
    reps = max(2, int(math.ceil(want / n)))
    if mode == "poll":
        for r in range(reps):            # ConsumerRecord objects exist before the clock starts,
            rlog.rep_records(r)           # exactly as they do for the CPU arm
        gc.collect()
        gc.freeze()                      # (millions of live tuples: keep them out of later GC passes)
    host_results = torch.zeros((reps * SLOT_BATCHES + 2, BATCH), dtype=torch.int32).pin_memory()
    _iterate(make_dataset_cls(rlog, mode, 1), device, host_results)           # warm-up iteration
    seconds, batches, stats, checksum = _iterate(make_dataset_cls(rlog, mode, reps), device, host_results)
    # the iteration ran to the end of the stream: exactly reps * n records were consumed
    if stats.get("records") != reps * n:
        log(f"warning: pipeline counted {stats.get('records')} records, the stream holds {reps * n}")
    log(f"e2e {mode}: {seconds:.3f} s, {batches} batches, checksum {checksum},",
        {k: (round(v, 4) if isinstance(v, float) else v) for k, v in stats.items()})
    return {"seconds": seconds, "batches": batches, "records": reps * n,
            "bytes": int(stats["bytes"]), "reps": reps, "stats": stats}


def run_commit_barrier(rank, world, device, n_batches=600, warm=16):
    """BASELINE config 5 with its cross-rank commit barrier: 64 partitions x 64-byte records, every
    rank ingests its own partitions (p % world == rank) in batches of 256 through
    auto_commit(loader, commit_barrier=True): after each batch the ranks exchange their commit
    vectors (all_gather, host tensors over gloo), then commit.  Latency-bound by design
    (SURVEY.md §8d row 5): reports the barrier's p50 / p99 and batches/s per rank."""
    import torch
    from torch.utils.data import DataLoader
    from torchkafka_b200 import KafkaDataset, RawBytes, auto_commit
    from torchkafka_b200.testing import stub_kafka as sk
    from torchkafka_b200.testing import synthetic as syn
    import numpy as np
    B = 256
    n = (n_batches + warm) * B
    owned = np.asarray([p for p in range(PARTITIONS) if p % world == rank], dtype=np.int32)
    i = np.arange(n)
    clog = sk.RecordLog(syn.cfg5_values(n, seed=5 + rank), owned[i % len(owned)], i // len(owned))

    class Cfg5(KafkaDataset):
        payload_schema = RawBytes((64,), torch.uint8)
        device_slot_bytes = 64 * B * 8              # small slots: latency, not throughput
        device_index = device
        max_partitions = PARTITIONS

        @classmethod
        def new_consumer(cls, *a, **k):
            k.pop("enable_auto_commit", None)
            return sk.KafkaConsumer(*a, log=clog, keep_commit_log=False, **k)

    ds = Cfg5("topic", group_id="cfg5")
    seen, t0 = 0, None
    for _batch in auto_commit(DataLoader(ds, batch_size=B), commit_barrier=True):
        if seen == warm:
            torch.cuda.synchronize()
            t0 = time.perf_counter()
        seen += 1
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    lat = sorted(list(ds.commit_barrier_us)[warm:warm + n_batches])
    ds.close()
    pct = lambda q: lat[min(len(lat) - 1, int(q * len(lat)))] if lat else float("nan")   # noqa: E731
    return {"p50_us": pct(0.5), "p99_us": pct(0.99), "batches_per_s": (seen - warm) / dt, "batches": seen - warm}


# --------------------------------------------------------------------------- CPU arms (oracle)
def cpu_process_factory():
    import torch
    from oracle import reference_path as ref
    return ref.process_json(torch.float32, MIN_SIZE), ref.pad_collate


def cpu_baseline_for(schema, log, batch_size, seconds=6.0):
    """cpu_baseline leg for any schema (benchmarks/all_configs.py calls this: outside tests/ and
    smoke(), this file is the only place that executes the oracle): records/s of the oracle on one
    core, bounded to `seconds`."""
    import torch
    from oracle import reference_path as ref
    from torchkafka_b200.schema import JsonArray
    from torchkafka_b200.testing import stub_kafka as sk
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from helpers import oracle_process
    torch.set_num_threads(1)
    variable = isinstance(schema, JsonArray) and schema.fixed_len is None
    collate = ref.pad_collate if variable else ref.default_collate
    consumer = sk.KafkaConsumer("topic", log=log, repeat=1000)
    t0 = time.perf_counter()
    for _ in ref.run(consumer, oracle_process(schema), batch_size, False, collate):
        if time.perf_counter() - t0 > seconds:
            break
    return consumer.consumed / (time.perf_counter() - t0)


def run_cpu_single(log, seconds):
    """Oracle, one process, one thread: the reference's own num_workers=0 configuration, bounded
    to about `seconds` of CPU work."""
    import torch
    from oracle import reference_path as ref
    from torchkafka_b200.testing import stub_kafka as sk
    torch.set_num_threads(1)
    process, collate = cpu_process_factory()
    reps = 64
    for r in range(8):
        log.rep_records(r)
    consumer = sk.KafkaConsumer("topic", log=log, repeat=reps)
    n_batches = 0
    t0 = time.perf_counter()
    for _batch, _commit in ref.run(consumer, process, BATCH, False, collate):
        n_batches += 1
        if time.perf_counter() - t0 >= seconds:
            break
    dt = time.perf_counter() - t0
    return {"records": consumer.consumed, "seconds": dt, "batches": n_batches}


def run_cpu_workers(base_log, steps, warmup, num_workers):
    """Oracle multi-worker arm: DataLoader(num_workers=W) + SIGUSR1 commits, partitions split
    across workers.  Returns after `steps` timed batches."""
    import torch
    from torch.utils.data import DataLoader
    from oracle import ref_dataset
    from torchkafka_b200.testing import stub_kafka as sk
    from torchkafka_b200.testing import synthetic as syn
    torch.set_num_threads(1)
    process, collate = cpu_process_factory()
    W = max(num_workers, 1)
    warm = max(warmup, 2 * num_workers)
    # every worker must be able to serve its share (+ prefetch) without running dry
    per_worker_batches = (warm + steps) // W + 4
    reps = max(1, -(-per_worker_batches * BATCH // max(int(len(base_log) * 0.93), 1)) + 1)
    worker_logs = []
    for w in range(W):
        sub = base_log.select_partitions(lambda p, w=w: p % W == w) if W > 1 else base_log
        sub = syn.tile_log(sub, reps * (W if W > 1 else 1)) if reps * W > 1 else sub
        sub.records()
        worker_logs.append(sub)
    gc.collect()
    gc.freeze()

    def factory(worker_id, n):
        return sk.KafkaConsumer("topic", log=worker_logs[worker_id], block_at_end=True)

    ds = ref_dataset.OracleWorkerDataset(process, factory)
    loader = DataLoader(ds, batch_size=BATCH, num_workers=num_workers, collate_fn=collate,
                        **({"prefetch_factor": 2} if num_workers else {}))
    t0 = None
    kept = 0
    seen = 0
    for batch in ref_dataset.iterate_with_commits(loader, warm + steps):
        if seen == warm:
            t0 = time.perf_counter()
        if seen >= warm:
            kept += int(batch[1].shape[0])
        seen += 1
    dt = time.perf_counter() - t0
    # consumed = kept / keep_rate of the workload (trailing skipped records are not observable
    # from the main process); keep rate measured on the base log
    keep_rate = sum(v.count(b",") + 1 >= MIN_SIZE for v in base_log._byte_values) / len(base_log)
    return {"seconds": dt, "kept": kept, "records": kept / keep_rate, "workers": num_workers,
            "warmup_batches": warm}


# --------------------------------------------------------------------------- main
def dist_env():
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")),
            int(os.environ.get("WORLD_SIZE", "1")))


def main_reference(args):
    rank, _, world = dist_env()
    if rank != 0:
        return 0
    cores = len(os.sched_getaffinity(0))
    workers = max(1, min(cores - 1, 32))
    base = build_base_log(SLOT_BATCHES)
    # a reference step is a bounded sample of whole batches: one batch when K is large, more
    # when K is small, so that the timed region always spans >= 512 batches (every worker
    # delivers ~30+) and the arm is measured in its steady state, not during worker start-up
    per_step = max(1, -(-512 // max(args.steps, 1)))
    r = run_cpu_workers(base, args.steps * per_step, max(args.warmup * per_step, 4 * workers), workers)
    value = r["records"] / r["seconds"]
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "records/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * r["seconds"] / (args.steps * per_step), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64->f32", "data": "synthetic",
        "config": shared_config(max(world, args.gpus, 1)),
        "setup": {"num_workers": workers, "warmup_batches": r["warmup_batches"], "batches_per_step": per_step,
                  "timed_region_s": r["seconds"]},
        "cpu_baseline": {"value": value, "unit": "records/s", "cores": workers + 1, "kind": "port",
                         "sample": f"{args.steps * per_step} batches of {BATCH} kept records via DataLoader("
                                   f"num_workers={workers}) + SIGUSR1 commits; oracle/ref_dataset.py"},
        "e2e": {"value": value, "unit": "records/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "host_cores": cores,
    }
    print(json.dumps(line), flush=True)
    sys.stdout.flush()
    os._exit(0)      # workers are parked inside blocking stub consumers; skip the 5 s/worker join


def main_gpu(args):
    import torch
    rank, local_rank, world = dist_env()
    if world > 1:
        import torch.distributed as dist
        # One rank per GPU.  The data path has no collective (partition sharding); the only
        # exchanges here are the barrier and the max-over-ranks of the timings, done on host
        # tensors (gloo) so that no NCCL banner can leak into the single JSON line on stdout.
        dist.init_process_group("cpu:gloo,cuda:nccl")
    device = local_rank if world > 1 else 0
    torch.cuda.set_device(device)
    steps, warmup = max(args.steps, 1), max(args.warmup, 0)
    base = build_base_log(SLOT_BATCHES, rank, world)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            import torch.distributed as dist
            dist.all_reduce(torch.zeros(1))          # host-side barrier (gloo)
        torch.cuda.synchronize()

    # ---- value: device-resident
    barrier()
    with Clocks(device) as clk:
        dev = run_device_resident(base, steps, warmup, device)
    barrier()
    clocks = clk.summary()

    # ---- e2e: public API from host records, one arm per consumer protocol
    arms = {}
    if not args.skip_e2e:
        from torchkafka_b200.testing import stub_kafka as sk
        base.encoded(RAW_BATCH_BYTES)              # the broker's log segment exists before the clock starts
        modes = [m for m in os.environ.get("TK_E2E_ARMS", "poll,packed,raw,raw_dma").split(",") if m]
        for mode in modes:
            if mode == "raw_dma":
                base.encoded(RAW_BATCH_BYTES).pin()       # the consumer's receive arena is page-locked memory
            barrier()
            arms[mode] = run_e2e(base, mode, device, E2E_MIN_SECONDS, 20e6,
                                 max_records=POLL_MAX_OBJECTS if mode == "poll" else None)
    barrier()
    cb = None
    if world > 1 and not args.skip_e2e:
        cb = run_commit_barrier(rank, world, device)
        barrier()

    order = ["poll", "packed", "raw", "raw_into", "raw_dma"]
    cols = [dev["ms"] / 1e3, float(dev["records"])]
    for m in order:
        a = arms.get(m)
        cols += [a["seconds"], float(a["records"]), float(a["bytes"])] if a else [0.0, 0.0, 0.0]
    cols += [cb["p50_us"], cb["p99_us"], cb["batches_per_s"]] if cb else [0.0, 0.0, 0.0]
    t = torch.tensor(cols, dtype=torch.float64)
    tmax, tsum = t.clone(), t.clone()
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
    # whole-job throughput: the units all ranks processed / the slowest rank's time
    value = float(tsum[1]) / float(tmax[0])
    ms = float(tmax[0]) * 1e3

    rc = 0
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
        achieved = dev["decode_algo_bytes_total"] / (dev["decode_ms_total"] * 1e6)
        traffic = None
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "roofline_traffic.json")))["bytes_per_launch"]
        except Exception:
            pass
        rec_per_step = len(base) / SLOT_BATCHES
        bytes_per_rec = (base.nbytes + 16 * len(base)) / len(base)
        d2h_per_step = 4 * BATCH + (64 + 8 * (SLOT_BATCHES + 1) * (PARTITIONS + 1)) // SLOT_BATCHES
        link_peak = dev["h2d_peak_gbs"]

        def arm_line(k, mode, path):
            secs, recs, byts = float(tmax[2 + 3 * k]), float(tsum[3 + 3 * k]), float(tsum[4 + 3 * k])
            if not secs:
                return None
            v = recs / secs
            # link bytes: payload + 16 B of record header per record (CSR arms), or the raw
            # RecordBatch bytes (raw arms) -- as counted by the pipeline over the whole window
            link_bytes = byts + (16 * recs if mode in ("poll", "packed") else 0)
            gbps_per_gpu = link_bytes / secs / 1e9 / world
            a = arms[mode]
            return {"value": v, "unit": "records/s", "path": path,
                    "h2d_bytes_per_step": int(link_bytes / recs * rec_per_step), "d2h_bytes_per_step": d2h_per_step,
                    "timed_region_s": secs, "steps_timed": int(a["batches"]) * world, "log_repeats": a["reps"],
                    "link_GBps_per_gpu": gbps_per_gpu, "link_peak_GBps": link_peak,
                    "link_frac": gbps_per_gpu / link_peak}

        paths = {
            "poll": "auto_commit(DataLoader(KafkaDataset with payload_schema)) over ConsumerRecord objects from "
                    "consumer.poll(); clock from before the iteration starts to the last batch + synchronize",
            "packed": "same API over consumer.poll_packed() (flat arrays, no per-record objects)",
            "raw": "same API over consumer.poll_raw(): RecordBatch v2 blobs of the fetch response, "
                   f"batch.size {RAW_BATCH_BYTES}; CRC-32C, varint framing and value gather on the GPU",
            "raw_into": "same API over consumer.poll_raw_into(): the consumer receives the fetch response's "
                        "RecordBatch v2 bytes straight into the slot's pinned buffer (its copy is the recv)",
            "raw_dma": "same API over consumer.poll_raw() with fetch_arenas(): the consumer's receive buffer is "
                       "page-locked memory and every blob goes to the device straight out of it (no host copy)",
        }
        lines = {m: arm_line(k, m, paths[m]) for k, m in enumerate(order) if m in arms}
        line = {
            "metric": METRIC, "value": value, "unit": "records/s", "n_gpus": world, "steps": steps,
            "warmup": warmup, "ms_per_step": ms / (steps * dev["repeats"]), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64->f32", "data": "synthetic",
            "config": shared_config(world),
            "repeats": dev["repeats"], "timed_region_s": ms / 1e3,
            "setup": {"slot_batches": SLOT_BATCHES, "slot_bytes": dev["slot_bytes"],
                      "passes": dev["passes"], "seed": f"{SEED} + rank"},
            "batch_ready_GBps": value * bytes_per_rec / 1e9,
            "gpu_launches": dev["launches"],
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic,
                         "algo_bytes_per_launch": dev["algo_bytes_per_launch"],
                         "kernel": "tk_json_decode = k_json_parse<f32> (thread per element parse + pad/length/mask rows, one kernel)",
                         "share_of_step": dev["decode_ms_total"] / dev["ms_timed_region"],
                         "peak_source": peak_src, "launches_timed": dev["decode_launches"],
                         "timed_in": "a second pass over the same K*R steps with the library's CUDA events "
                                     "around the kernel (tk_ctx_set_timing); share_of_step is against that pass"},
            "clocks": clocks,
        }
        if "poll" in lines:
            line["e2e"] = lines["poll"]
        for m in ("packed", "raw", "raw_into", "raw_dma"):
            if m in lines:
                line["e2e_" + m] = lines[m]
        if cb:
            k = 2 + 3 * len(order)
            line["commit_barrier"] = {
                "workload": "cfg5: 64 partitions x 64-byte records, batch 256, auto_commit(loader, commit_barrier=True)",
                "ranks": world, "transport": "torch.distributed all_gather of int64[65] host tensors (gloo)",
                "p50_us_max_over_ranks": float(tmax[k]), "p99_us_max_over_ranks": float(tmax[k + 1]),
                "batches_per_s_per_rank": float(tsum[k + 2]) / world, "batches_timed": cb["batches"]}
        bad = [m for m, l in lines.items() if l["link_frac"] > 1.0]
        if bad:
            log(f"INVALID: link_frac > 1 for {bad}: the timed region skipped work")
            line["invalid"] = f"link_frac > 1 for {bad}"
            rc = 3
        if world == 1 and not args.no_cpu:
            c = run_cpu_single(base, args.cpu_seconds)
            line["cpu_baseline"] = {
                "value": c["records"] / c["seconds"], "unit": "records/s", "cores": 1, "kind": "port",
                "sample": f"{c['records']} records ({c['seconds']:.1f} s) of the same workload, oracle/reference_path.py "
                          f"(json.loads + torch.tensor + pad collate + commit), 1 process, 1 thread"}
        print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(torch.zeros(1))
        dist.destroy_process_group()
    return rc


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2048)
    ap.add_argument("--warmup", type=int, default=128)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the in-line cpu_baseline leg")
    ap.add_argument("--cpu-seconds", type=float, default=10.0, help="bound of the cpu_baseline sample")
    ap.add_argument("--skip-e2e", action="store_true", help="device-resident leg only (profiling runs)")
    args = ap.parse_args()
    if args.impl == "reference":
        return main_reference(args)
    return main_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
