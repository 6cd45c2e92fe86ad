"""_tkpack.harvest_groups: flattening a poll() result into (address, length, offset, partition)
arrays, byte and record budgets included, GIL-free fast path and general path."""
import ctypes
import random

import numpy as np
import pytest

import helpers  # noqa: F401
from torchkafka_b200 import _tkpack
from torchkafka_b200.testing.stub_kafka import ConsumerRecord, TopicPartition


def make_groups(n_groups, per_group, seed=0):
    rng = random.Random(seed)
    groups, dense, flat = [], [], []
    for g in range(n_groups):
        recs = [ConsumerRecord("t", g, o, 0, 0, None, bytes(rng.getrandbits(8) for _ in range(rng.randrange(0, 40))),
                               [], None, -1, 0, -1) for o in range(100 * g, 100 * g + per_group)]
        groups.append((TopicPartition("t", g), recs))
        dense.append(g + 3)
        flat += [(r.value, r.offset, g + 3) for r in recs]
    return groups, dense, flat


def run(groups, dense, max_records, max_bytes):
    cap = max(max_records, 1)
    ptrs, lens = np.zeros(cap, np.uint64), np.zeros(cap, np.uint32)
    offs, parts = np.zeros(cap, np.int64), np.zeros(cap, np.int32)
    vi, oi = ConsumerRecord._fields.index("value"), ConsumerRecord._fields.index("offset")
    taken, tot, g_done, partial, keep = _tkpack.harvest_groups(
        groups, dense, vi, oi, max_records, max_bytes, ptrs.ctypes.data, lens.ctypes.data,
        offs.ctypes.data, parts.ctypes.data)
    vals = [ctypes.string_at(int(p), int(n)) for p, n in zip(ptrs[:taken], lens[:taken])]
    return taken, tot, g_done, partial, list(zip(vals, offs[:taken].tolist(), parts[:taken].tolist()))


@pytest.mark.parametrize("n_groups,per_group", [(1, 50), (8, 100), (16, 600), (64, 300)])
def test_everything_fits(n_groups, per_group):
    groups, dense, flat = make_groups(n_groups, per_group)
    taken, tot, g_done, partial, got = run(groups, dense, len(flat) + 10, 1 << 30)
    assert (taken, g_done, partial) == (len(flat), n_groups, 0)
    assert got == flat and tot == sum(len(v) for v, _, _ in flat)


@pytest.mark.parametrize("frac", [0.0, 0.37, 0.5, 0.99])
def test_byte_budget_cuts_inside_a_group(frac):
    groups, dense, flat = make_groups(16, 600, seed=1)
    budget = int(sum(len(v) for v, _, _ in flat) * frac)
    taken, tot, g_done, partial, got = run(groups, dense, len(flat), budget)
    # reference: take records while they fit
    t, want = 0, 0
    for v, _, _ in flat:
        if t + len(v) > budget:
            break
        t += len(v)
        want += 1
    assert (taken, tot) == (want, t) and got == flat[:want]
    assert g_done == want // 600 and partial == want % 600


def test_record_budget_cuts_inside_a_group():
    groups, dense, flat = make_groups(16, 600, seed=2)
    taken, tot, g_done, partial, got = run(groups, dense, 5000, 1 << 30)
    assert taken == 5000 and got == flat[:5000] and (g_done, partial) == (8, 200)


def test_unusual_records_fall_back():
    groups, dense, flat = make_groups(16, 600, seed=3)

    class Rec:                                                    # attribute access, not a tuple
        def __init__(self, value, offset):
            self.value, self.offset = value, offset
    tp, recs = groups[5]
    groups[5] = (tp, [Rec(r.value, r.offset) for r in recs])
    taken, tot, g_done, partial, got = run(groups, dense, len(flat), 1 << 30)
    assert taken == len(flat) and got == flat
