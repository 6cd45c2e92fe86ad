#!/usr/bin/env python3
"""Device-resident timing of the JSON-object chain (scan, plan, k_json_object, commit) on slots of
--records records; the command the ncu capture of k_json_object in profiles/ ran."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "benchmarks"))

import all_configs as ac                                     # noqa: E402
from torchkafka_b200.testing import synthetic as syn         # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--records", type=int, default=65536)
ap.add_argument("--passes", type=int, default=24)
args = ap.parse_args()
cfg = dict(ac.CONFIGS[6], n=args.records)
log = syn.make_log(6, args.records, seed=106, partitions=64, layout="chunked", chunk=500)
dev, _ = ac.device_resident(cfg, log, passes=args.passes)
dev["records"] = args.records
dev["bytes_per_record"] = log.nbytes / len(log)
print(json.dumps(dev))
