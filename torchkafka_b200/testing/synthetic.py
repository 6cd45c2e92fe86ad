"""Seeded synthetic record generators for the five BASELINE.json configs.

Shapes and distributions follow SURVEY.md §8(d): ``random.Random(seed)``,
``json.dumps`` with default separators, 64 partitions unless noted.  Two
interleavings are offered: ``"roundrobin"`` (record i -> partition i % P, the
survey's layout) and ``"chunked"`` (runs of ``chunk`` records per partition,
which is how a real fetch response arrives).  Both give per-partition offsets
that increase in consumption order.
"""
from __future__ import annotations

import json
import random
from typing import List, Tuple

import numpy as np

from .stub_kafka import RecordLog

IMAGENET_MEAN = (0.485, 0.456, 0.406)
IMAGENET_STD = (0.229, 0.224, 0.225)


def _coords(n: int, partitions: int, layout: str, chunk: int) -> Tuple[np.ndarray, np.ndarray]:
    i = np.arange(n, dtype=np.int64)
    if layout == "roundrobin":
        return (i % partitions).astype(np.int32), i // partitions
    if layout == "chunked":
        blk = i // chunk
        part = (blk % partitions).astype(np.int32)
        off = (blk // partitions) * chunk + (i % chunk)
        return part, off
    raise ValueError(f"unknown layout {layout!r}")


def _log(values: List[bytes], partitions: int, layout: str, chunk: int) -> RecordLog:
    part, off = _coords(len(values), partitions, layout, chunk)
    return RecordLog(values, part, off)


def cfg1_values(n: int, seed: int = 0) -> List[bytes]:
    """README MyDataset payload: JSON array of 8 floats, uniform(-1e3, 1e3)."""
    rng = random.Random(seed)
    u = rng.uniform
    return [json.dumps([u(-1e3, 1e3) for _ in range(8)]).encode() for _ in range(n)]


def cfg2_values(n: int, seed: int = 0, lo: int = 16, hi: int = 256) -> List[bytes]:
    """Variable-length JSON float arrays, len randint(16, 256)."""
    rng = random.Random(seed)
    u = rng.uniform
    return [json.dumps([u(-1e3, 1e3) for _ in range(rng.randint(lo, hi))]).encode()
            for _ in range(n)]


def cfg3_values(n: int, seed: int = 0, max_tokens: int = 2048, vocab: int = 128000) -> List[bytes]:
    """JSON int token-id lists, len randint(1, max_tokens), ids randrange(vocab)."""
    rng = random.Random(seed)
    rr = rng.randrange
    return [json.dumps([rr(0, vocab) for _ in range(rng.randint(1, max_tokens))]).encode()
            for _ in range(n)]


def cfg4_values(n: int, seed: int = 0, h: int = 224, w: int = 224, c: int = 3) -> List[bytes]:
    """Raw HWC uint8 images."""
    rng = np.random.default_rng(seed)
    return [rng.integers(0, 256, h * w * c, dtype=np.uint8).tobytes() for _ in range(n)]


def cfg5_values(n: int, seed: int = 0, size: int = 64) -> List[bytes]:
    """64-byte opaque records."""
    rng = np.random.default_rng(seed)
    blob = rng.integers(0, 256, n * size, dtype=np.uint8).tobytes()
    return [blob[i * size:(i + 1) * size] for i in range(n)]


def object_values(n: int, seed: int = 0, fields=("x", "y", "z", "w", "vx", "vy", "vz", "t"),
                  extras: bool = True) -> List[bytes]:
    """JSON objects with named numeric members: floats uniform(-1e3, 1e3) and a
    few ints under `fields`, plus (``extras``) an id string and a tag list the schema ignores."""
    rng = random.Random(seed)
    u = rng.uniform
    out = []
    for i in range(n):
        d = {}
        if extras:
            d["id"] = "sensor-%06d" % rng.randrange(10 ** 6)
        for k, f in enumerate(fields):
            d[f] = rng.randrange(-1000, 1000) if k % 4 == 3 else u(-1e3, 1e3)
        if extras:
            d["tags"] = ["a", {"n": i % 7}]
        out.append(json.dumps(d).encode())
    return out


OBJECT_FIELDS = ("x", "y", "z", "w", "vx", "vy", "vz", "t")
_GEN = {1: cfg1_values, 2: cfg2_values, 3: cfg3_values, 4: cfg4_values, 5: cfg5_values, 6: object_values}


def make_log(cfg: int, n: int, seed: int = 0, partitions: int = 64,
             layout: str = "roundrobin", chunk: int = 500, **kw) -> RecordLog:
    """RecordLog for BASELINE config ``cfg`` (1..5) with ``n`` records (6: JSON objects with the
    named numeric members OBJECT_FIELDS, SURVEY.md §8f-4 -- not a BASELINE config)."""
    return _log(_GEN[cfg](n, seed, **kw), partitions, layout, chunk)


def tile_log(log: RecordLog, times: int) -> RecordLog:
    """Repeat a log ``times`` times (offsets keep increasing per partition).

    Generating millions of JSON records with ``json.dumps`` takes minutes;
    benchmarks build a modest seeded log once and tile it.
    """
    n = len(log)
    vals = log._byte_values * times
    part = np.tile(log.partitions, times)
    per_part_count = np.bincount(log.partitions, minlength=int(log.partitions.max()) + 1)
    # offset of a record in repetition r = original offset + r * (records of that partition per rep)
    # (requires the original offsets of each partition to be 0..count-1 in order, true for _coords)
    reps = np.repeat(np.arange(times, dtype=np.int64), n)
    off = np.tile(log.offsets, times) + reps * per_part_count[part]
    return RecordLog(vals, part, off, log.topic)
