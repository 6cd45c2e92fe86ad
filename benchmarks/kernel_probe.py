#!/usr/bin/env python3
"""Run the slot kernel sequence of one config a few times with the slot resident in HBM
(profiling target for ncu: `ncu ... python benchmarks/kernel_probe.py 4`)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "benchmarks"))

import numpy as np   # noqa: E402
import torch         # noqa: E402

from all_configs import CONFIGS, arrays                       # noqa: E402
from torchkafka_b200.engine import SlotEngine                 # noqa: E402
from torchkafka_b200.testing import synthetic as syn          # noqa: E402


def main():
    c = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    passes = int(sys.argv[2]) if len(sys.argv) > 2 else 6
    cfg = CONFIGS[c]
    n = {4: 512, 5: 1 << 20}.get(c, cfg["n"])
    log = syn.make_log(c, n, seed=100 + c, partitions=64, layout="chunked", chunk=500)
    eng = SlotEngine(cfg["schema"], cfg["batch"], slot_bytes=log.nbytes + 4096, slot_records=len(log),
                     n_slots=4, device=0, n_partitions=64)
    for s in range(4):
        k, nb = eng.fill_packed(s, *arrays(log))
        eng.submit(s, k, nb)
    torch.cuda.synchronize()
    prev_slot, prev_out = None, None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for i in range(passes):
        if i == 2:
            e0.record(eng.stream)
        s = i % 4
        eng.assume_resident(s, len(log), log.nbytes)
        out = eng.process(s, len(log), prev_slot, prev_out, resident=True)
        prev_slot, prev_out = s, out
    e1.record(eng.stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / max(passes - 2, 1)
    print(f"cfg{c}: {len(log)} records, {log.nbytes / 1e6:.1f} MB per slot, {ms * 1e3:.1f} us per slot, "
          f"{log.nbytes / ms / 1e6:.1f} GB/s payload")
    eng.close()


if __name__ == "__main__":
    main()
