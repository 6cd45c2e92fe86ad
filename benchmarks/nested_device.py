#!/usr/bin/env python3
"""Device-resident timing of the nested-array chain (JsonNested) for small records
(thread per record out of the shared-memory window) and for records larger than the window (one
per warp); the numbers quoted in DESIGN.md section 3."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "benchmarks"))
import all_configs as ac
from torchkafka_b200.schema import JsonNested
from torchkafka_b200.testing.stub_kafka import RecordLog
for shape, n in [((4, 3), 65536), ((64, 24), 2048)]:
    rng = np.random.default_rng(1)
    vals = [json.dumps(rng.uniform(-1e3, 1e3, shape).tolist()).encode() for _ in range(min(n, 4096))]
    vals = (vals * (n // len(vals) + 1))[:n]
    i = np.arange(n)
    log = RecordLog(vals, i % 64, i // 64)
    cfg = dict(schema=JsonNested(shape, torch.float32), batch=256, slot_bytes=log.nbytes + (1 << 20))
    dev, _ = ac.device_resident(cfg, log, passes=8)
    print(shape, n, log.nbytes // n, dev)
