#!/usr/bin/env python3
"""Per-config measurements for the five BASELINE.json configs (SURVEY.md §8d reporting).

For every config:
  device   records/s and payload GB/s of the slot kernel sequence with the slot already in HBM
           (CUDA events on the engine stream, 4 distinct slots cycled so inputs come from HBM)
  e2e      records/s through the public API (auto_commit(DataLoader(dataset))) from host
           ConsumerRecord objects (poll) and from flat arrays (poll_packed), with link fraction
           against the measured H2D peak
  cpu      the oracle (reference algorithm) on one core, bounded sample

Writes one JSON object per config to stdout / --out.  `bench.py` stays the contract benchmark
(config 2); this script is the wider table that goes into profiles/.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np      # noqa: E402
import torch            # noqa: E402

from torchkafka_b200 import KafkaDataset, auto_commit                                   # noqa: E402
from torchkafka_b200.engine import SlotEngine                                           # noqa: E402
from torchkafka_b200.schema import ImageU8, JsonArray, JsonObject, RawBytes, TokenList  # noqa: E402
from torchkafka_b200.testing import stub_kafka as sk                                    # noqa: E402
from torchkafka_b200.testing import synthetic as syn                                    # noqa: E402

CONFIGS = {
    1: dict(name="cfg1 README 8 floats", schema=JsonArray(torch.float32, fixed_len=8), batch=4,
            n=60000, slot_bytes=8 << 20),
    2: dict(name="cfg2 varlen floats, min_size=32", schema=JsonArray(torch.float32, max_len=256, min_size=32),
            batch=1024, n=17600, slot_bytes=48 << 20),
    3: dict(name="cfg3 token lists <= 2048", schema=TokenList(torch.int32, max_len=2048), batch=512,
            n=6144, slot_bytes=48 << 20),
    4: dict(name="cfg4 224x224x3 images -> bf16 NCHW", schema=ImageU8(224, 224, 3), batch=256,
            n=256, slot_bytes=256 * 150528),
    5: dict(name="cfg5 64-byte records, commit offsets", schema=RawBytes((64,), torch.uint8), batch=256,
            n=262144, slot_bytes=17 << 20),
    # not a BASELINE config: SURVEY §8f-4, JSON objects with 8 named numeric members + 2 ignored ones
    6: dict(name="objects, 8 named members of 10", schema=JsonObject(syn.OBJECT_FIELDS, torch.float32), batch=1024,
            n=65536, slot_bytes=20 << 20),
}


def arrays(log):
    return log.values, log.starts.astype(np.uint32), log.partitions, log.offsets


def device_resident(cfg, log, passes=24):
    eng = SlotEngine(cfg["schema"], cfg["batch"], slot_bytes=max(cfg["slot_bytes"], log.nbytes + 4096),
                     slot_records=len(log), n_slots=4, device=0, n_partitions=64)
    for s in range(4):
        n, nb = eng.fill_packed(s, *arrays(log))
        eng.submit(s, n, nb)
    torch.cuda.synchronize()

    def run(k):
        prev_slot, prev_out = None, None
        for i in range(k):
            s = i % 4
            eng.assume_resident(s, len(log), log.nbytes)
            out = eng.process(s, len(log), prev_slot, prev_out, resident=True)
            prev_slot, prev_out = s, out
    run(4)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(eng.stream)
    run(passes)
    e1.record(eng.stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / passes
    res = eng.result((passes - 1) % 4)
    assert res.n_errors == 0
    peak = eng.h2d_peak_gbs(256 << 20, 5)
    eng.close()
    return {"ms_per_slot": ms, "records_per_s": len(log) / ms * 1e3,
            "payload_GBps": (log.nbytes + 16 * len(log)) / ms / 1e6}, peak


RAW_BATCH_BYTES = int(os.environ.get("RAW_BATCH_BYTES", str(64 << 10)))
_rc = os.environ.get("RAW_CODEC", "0")                  # codec of the raw arm: 0, 2 (snappy), 3 (LZ4), "lz4i", ...
RAW_CODEC = int(_rc) if _rc.isdigit() else _rc


def e2e(cfg, log, packed, repeat):
    """packed: False = poll() objects, True = poll_packed() arrays, "raw" = poll_raw() record batches,
    "raw_dma" = poll_raw() blobs copied to the device out of the consumer's page-locked buffer.

    A complete iteration over `repeat` repetitions of the log, timed from before the iteration (and
    the dataset's pipeline) is started until its last batch was delivered and the device is idle;
    a warm-up iteration over one repetition runs first (bench.py's e2e leg, for any schema)."""
    from torch.utils.data import DataLoader
    consumer_cls = sk.KafkaConsumer if packed else sk.PollOnlyConsumer
    raw = packed in ("raw", "raw_dma")

    def make(reps):
        class DS(KafkaDataset):
            payload_schema = cfg["schema"]
            device_slot_bytes = cfg["slot_bytes"] + (cfg["slot_bytes"] >> 2 if raw else 0)   # framing
            device_slots = int(os.environ.get("DEVICE_SLOTS", "3"))
            device_producers = int(os.environ.get("TK_PRODUCERS", "2"))
            device_input = packed if raw else "auto"

            @classmethod
            def new_consumer(cls, *a, **k):
                k.pop("enable_auto_commit", None)
                return consumer_cls(*a, log=log, repeat=reps, raw_codec=RAW_CODEC, keep_commit_log=False,
                                    raw_batch_bytes=max(RAW_BATCH_BYTES, 2 * cfg.get("record_bytes", 0)), **k)
        return DS("topic", group_id="cfg")

    def run(reps):
        ds = make(reps)
        sink = torch.zeros(256, dtype=torch.uint8).pin_memory()
        torch.cuda.synchronize()
        seen = 0
        t0 = time.perf_counter()
        for batch in auto_commit(DataLoader(ds, batch_size=cfg["batch"])):
            seen += 1
            if seen % 64 == 0:          # a slice of every 64th batch is read back (per-step D2H of 4-record
                t = batch[0] if isinstance(batch, tuple) else batch          # batches would time cudaMemcpyAsync)
                flat = t.reshape(-1).view(torch.uint8)
                k = min(256, flat.numel())
                sink[:k].copy_(flat[:k], non_blocking=True)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        stats = dict(ds.last_pipeline_stats)
        ds.close()
        return dt, stats

    run(1)                                  # cold: context creation, first allocations
    pilot = max(1, min(4, repeat))
    dtp, _ = run(pilot)                     # warm pilot: its rate sizes the timed iteration (~2.5 s)
    repeat = max(2, min(repeat, int(2.5 * pilot / dtp) + 1))
    dt, stats = run(repeat)
    return repeat * len(log) / dt, stats, repeat


def cpu_oracle(cfg, log, seconds=6.0):
    """The CPU arm is bench.py's cpu_baseline leg (the one place outside the tests that runs the oracle)."""
    import bench
    return bench.cpu_baseline_for(cfg["schema"], log, cfg["batch"], seconds)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="1,2,3,4,5")
    ap.add_argument("--out", default=None)
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    rows = []
    for c in [int(x) for x in args.configs.split(",")]:
        cfg = CONFIGS[c]
        # runs of 500 records per partition; images (150 KB each) in runs of 4 so that a fetch of one
        # slot still spans all 64 partitions, as a real fetch response does
        log = syn.make_log(c, cfg["n"], seed=100 + c, partitions=64, layout="chunked", chunk=4 if c == 4 else 500)
        dev, peak = device_resident(cfg, log)
        link_bytes = (log.nbytes + 16 * len(log)) / len(log)
        slot_recs = max(1, int(cfg["slot_bytes"] // max(log.nbytes / len(log), 1)))
        # the poll arm is bounded by the record objects it may hold (they exist before the clock
        # starts, as they do for the CPU arm); the other arms run for ~2 s
        link_rate = peak * 1e9 / link_bytes                       # records/s the link could carry
        obj_budget = int(os.environ.get("TK_POLL_OBJECTS", "4000000"))
        row = {"config": c, "name": cfg["name"], "records": len(log), "avg_value_bytes": log.nbytes / len(log),
               "link_bytes_per_record": link_bytes, "device": dev, "h2d_peak_GBps": peak}
        for packed in (False, True, "raw", "raw_dma"):
            if packed in ("raw", "raw_dma"):
                enc = log.encoded(max(RAW_BATCH_BYTES, 2 * cfg.get("record_bytes", 0)), RAW_CODEC)   # before the clock
                lb = enc.nbytes / len(log)
            else:
                lb = link_bytes
            want = max(2.0 * min(link_rate, 60e6), 20 * 5 * slot_recs)
            if packed is False:
                want = min(want, obj_budget)
            repeat = max(2, int(want // len(log)) + 1)
            if packed is False:
                for r in range(repeat):
                    log.rep_records(r)
                import gc
                gc.collect()
                gc.freeze()
            rps, stats, repeat = e2e(cfg, log, packed, repeat)
            key = {False: "e2e_poll", True: "e2e_packed", "raw": "e2e_raw", "raw_dma": "e2e_raw_dma"}[packed]
            row[key] = {"records_per_s": rps, "link_GBps": rps * lb / 1e9, "link_frac": rps * lb / 1e9 / peak,
                        "seconds": repeat * len(log) / rps, "log_repeats": repeat}
            assert row[key]["link_frac"] <= 1.0, (key, row[key])
        if not args.no_cpu:
            row["cpu_oracle_1core_records_per_s"] = cpu_oracle(cfg, log)
        rows.append(row)
        print(json.dumps(row), flush=True)
    if args.out:
        with open(args.out, "w") as f:
            json.dump(rows, f, indent=1)


if __name__ == "__main__":
    main()
