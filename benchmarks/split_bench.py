#!/usr/bin/env python3
"""Time of the record-batch step alone (H2D of the raw fetch bytes + decompression + CRC-32C +
split into the slot's CSR) for one slot of cfg2 records, per codec and batch size.

    python benchmarks/split_bench.py [--codecs 0,2,3,lz4i,1,4] [--batch-bytes 16384,65536] [--out f.json]
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

from torchkafka_b200.engine import SlotEngine                      # noqa: E402
from torchkafka_b200.schema import JsonArray                       # noqa: E402
from torchkafka_b200.testing import synthetic as syn               # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--codecs", default="0,2,3,lz4i,1,4")
    ap.add_argument("--batch-bytes", default="16384,65536")
    ap.add_argument("--records", type=int, default=17600)
    ap.add_argument("--iters", type=int, default=8)
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    log = syn.make_log(2, args.records, seed=102, partitions=64, layout="chunked", chunk=500)
    rows = []
    for codec in args.codecs.split(","):
        codec = int(codec) if codec.isdigit() else codec
        for bb in [int(x) for x in args.batch_bytes.split(",")]:
            enc = log.encoded(bb, codec)
            eng = SlotEngine(JsonArray(torch.float32, max_len=256, min_size=32), 1024,
                             slot_bytes=max(log.nbytes, enc.nbytes) + (1 << 20), slot_records=len(log) + 16,
                             n_slots=2, device=0, n_partitions=64)
            hb, _ = eng.host_views(0)
            hb[:enc.nbytes] = enc.data
            # one blob per run of batches of the same partition
            blobs = []
            b = 0
            while b < enc.n_batches:
                e = int(enc.run_end[b])
                blobs.append((int(enc.start[b]), int(enc.start[e] - enc.start[b]), int(enc.partition[b]), 0, 0))
                b = e
            for _ in range(2):
                n = eng.submit_recordbatches(0, enc.nbytes, blobs)
                eng.recordbatch_result(0)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(args.iters):
                n = eng.submit_recordbatches(0, enc.nbytes, blobs)
                nrec, nbytes, nb = eng.recordbatch_result(0)
            dt = (time.perf_counter() - t0) / args.iters
            assert nrec == len(log) and nbytes == log.nbytes, (nrec, nbytes)
            row = {"codec": codec, "batch_bytes": bb, "batches": nb, "compressed_MB": enc.nbytes / 1e6,
                   "ms_per_slot": dt * 1e3, "records_per_s": nrec / dt, "decompressed_GBps": nbytes / dt / 1e9}
            rows.append(row)
            print(json.dumps(row), flush=True)
            eng.close()
    if args.out:
        with open(args.out, "w") as f:
            json.dump(rows, f, indent=1)


if __name__ == "__main__":
    main()
