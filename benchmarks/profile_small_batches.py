#!/usr/bin/env python3
"""cProfile of the consuming thread for cfg1 (8 floats, batch 4): where the per-batch Python time goes."""
import cProfile
import os
import pstats
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "benchmarks"))
from all_configs import CONFIGS, e2e          # noqa: E402
from torchkafka_b200.testing import synthetic as syn   # noqa: E402

cfg = CONFIGS[1]
log = syn.make_log(1, cfg["n"], seed=101, partitions=64, layout="chunked", chunk=500)
for r in range(6):
    log.rep_records(r)
pr = cProfile.Profile()
pr.enable()
rps = e2e(cfg, log, True, 6, 60000, 2000)
pr.disable()
print("records/s", round(rps))
pstats.Stats(pr).sort_stats("tottime").print_stats(18)
