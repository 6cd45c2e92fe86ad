#!/usr/bin/env python3
"""Condense a bench.py JSON line (stdin) to: label, records/s, us per slot, decode GB/s."""
import json
import sys

d = json.loads(sys.stdin.read().strip().splitlines()[-1])
sb = d.get("setup", d["config"]).get("slot_batches", 32)
print(sys.argv[1] if len(sys.argv) > 1 else "", round(d["value"] / 1e6, 2), "Mrec/s",
      round(d["ms_per_step"] * sb * 1e3, 1), "us/slot", round(d["roofline"]["achieved"], 1), "GB/s decode")
