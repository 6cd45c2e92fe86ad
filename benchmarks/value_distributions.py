#!/usr/bin/env python3
"""Device-resident rate of the JSON array chain (cfg2 shape: 16-256 floats per record, float32,
pad + mask collate, batch 1024) for value distributions other than BASELINE's uniform(-1e3, 1e3):
how much of the stream the number fast lane takes decides the parse kernel's speed."""
import json
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "benchmarks"))

import numpy as np                                            # noqa: E402
import all_configs as ac                                     # noqa: E402
from torchkafka_b200.testing.stub_kafka import RecordLog     # noqa: E402

DISTS = {
    "uniform(-1e3, 1e3)": lambda r: r.uniform(-1e3, 1e3),
    "uniform(-1, 1)": lambda r: r.uniform(-1, 1),
    "gauss(0, 1)": lambda r: r.gauss(0, 1),
    "gauss(0, 1e-3)": lambda r: r.gauss(0, 1e-3),
    "lognormal(0, 6)": lambda r: r.lognormvariate(0, 6),
    "round(x, 4)": lambda r: round(r.uniform(-1e3, 1e3), 4),
    "ints": lambda r: r.randrange(-10 ** 6, 10 ** 6),
}


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 17600
    cfg = dict(ac.CONFIGS[2], n=n)
    for name, draw in DISTS.items():
        rng = random.Random(7)
        vals = [json.dumps([draw(rng) for _ in range(rng.randint(16, 256))]).encode() for _ in range(n)]
        i = np.arange(n)
        log = RecordLog(vals, (i // 500 % 64).astype(np.int32), i // (500 * 64) * 500 + i % 500)
        dev, _ = ac.device_resident(cfg, log, passes=16)
        text = b"".join(vals[:200])
        print(json.dumps({"values": name, "records_per_s": dev["records_per_s"], "ms_per_slot": dev["ms_per_slot"],
                          "payload_GBps": dev["payload_GBps"], "bytes_per_element": len(text) / (text.count(b",") + 200),
                          "with_exponent": text.count(b"e") / (text.count(b",") + 200)}), flush=True)


if __name__ == "__main__":
    main()
