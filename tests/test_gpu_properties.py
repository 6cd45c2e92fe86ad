"""Full-size runs checked through size-independent properties (the oracle would need minutes
at these sizes): the result must not depend on how the stream is cut into staging slots or on
the consumer protocol; batch counts, kept counts and the final committed offsets must equal
what the record log implies; decoding is idempotent."""
import numpy as np
import pytest
import torch
from torch.utils.data import DataLoader

from torchkafka_b200 import KafkaDataset, auto_commit
from torchkafka_b200.schema import JsonArray, JsonObject, RawBytes
from torchkafka_b200.testing import stub_kafka as sk
from torchkafka_b200.testing import synthetic as syn

pytestmark = pytest.mark.gpu


def run(log, schema, batch_size, repeat, slot_bytes, packed, producers=2, raw_codec=None, raw_batch_bytes=65536):
    """packed: False = poll() objects, True = poll_packed() arrays; raw_codec not None: poll_raw()
    record batches (0 plain, 3 / "lz4i" LZ4, 2 snappy)."""
    consumer_cls = sk.KafkaConsumer if packed else sk.PollOnlyConsumer

    class DS(KafkaDataset):
        payload_schema = schema
        device_slot_bytes = slot_bytes
        device_producers = producers
        device_input = "auto" if raw_codec is None else "raw"

        @classmethod
        def new_consumer(cls, *a, **k):
            k.pop("enable_auto_commit", None)
            return consumer_cls(*a, log=log, repeat=repeat, raw_codec=raw_codec or 0,
                                raw_batch_bytes=raw_batch_bytes, **k)

    ds = DS("topic", group_id="prop")
    sums, rows, n_batches = [], 0, 0
    for batch in auto_commit(DataLoader(ds, batch_size=batch_size)):
        vals = batch if isinstance(batch, torch.Tensor) else batch[0]
        if isinstance(batch, torch.Tensor):
            sums.append(vals.double().sum())
        else:
            sums.append(vals.double().sum() + batch.lengths.double().sum() * 1e-3)
        rows += vals.shape[0]
        n_batches += 1
    digest = torch.stack(sums).cpu().numpy()
    committed = dict(ds._consumer.committed_log[-1])
    calls = len(ds._consumer.committed_log)
    ds.close()
    return digest, rows, n_batches, committed, calls


def expected_positions(log, repeat):
    per = np.bincount(log.partitions, minlength=int(log.partitions.max()) + 1)
    return {p: int(c) * repeat for p, c in enumerate(per) if c}


def test_cfg1_one_million_records_slot_and_protocol_invariance():
    base = syn.make_log(1, 50000, seed=31, layout="chunked", chunk=500)
    repeat, B = 20, 4                                  # 1,000,000 records, BASELINE config 1
    sc = JsonArray(torch.float32, fixed_len=8)
    a = run(base, sc, B, repeat, 8 << 20, packed=True)
    b = run(base, sc, B, repeat, 3 << 20, packed=False, producers=1)
    assert a[1] == b[1] == len(base) * repeat
    assert a[2] == b[2] == len(base) * repeat // B
    assert np.array_equal(a[0], b[0])                  # every batch identical, bit for bit
    assert a[3] == b[3] == expected_positions(base, repeat)
    assert a[4] == a[2] and b[4] == b[2]               # one commit per batch
    # the same 50 000 records repeat 20 times: per-repetition digests repeat as well
    per_rep = a[0].reshape(repeat, -1)
    assert np.array_equal(per_rep, np.tile(per_rep[0], (repeat, 1)))


def test_cfg2_large_filter_counts_and_idempotence():
    base = syn.make_log(2, 20000, seed=32, layout="chunked", chunk=500)
    repeat, B = 6, 1024                                # 120 000 records, ~320 MB of JSON
    kept_base = int(np.sum([v.count(b",") + 1 >= 32 for v in base._byte_values]))
    sc = JsonArray(torch.float32, max_len=256, min_size=32)
    a = run(base, sc, B, repeat, 48 << 20, packed=True)
    b = run(base, sc, B, repeat, 13 << 20, packed=True)
    c = run(base, sc, B, repeat, 48 << 20, packed=True)
    assert a[1] == b[1] == kept_base * repeat
    assert a[2] == -(-kept_base * repeat // B)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[0], c[0])
    assert a[3] == b[3] == expected_positions(base, repeat)


def test_cfg5_raw_records_commit_equals_log_positions():
    base = syn.make_log(5, 200000, seed=33)
    a = run(base, RawBytes((64,), torch.uint8), 256, 5, 4 << 20, packed=True)
    assert a[1] == 1000000
    assert a[3] == expected_positions(base, 5)
    total = int(np.frombuffer(base.values, np.uint8).astype(np.int64).sum()) * 5
    assert int(a[0].sum()) == total                    # checksum of checksums over all bytes


def test_cfg2_record_batches_and_codecs_give_the_same_batches():
    """The stream served as ConsumerRecord arrays, as RecordBatch v2 bytes, and as LZ4 / snappy /
    gzip / zstd compressed batches (~160 MB of JSON): every batch and the final commit must be
    identical."""
    base = syn.make_log(2, 20000, seed=34, layout="chunked", chunk=300)
    repeat, B = 3, 1024
    sc = JsonArray(torch.float32, max_len=256, min_size=32)
    ref = run(base, sc, B, repeat, 48 << 20, packed=True)
    for codec, bb in [(0, 65536), (0, 1 << 20), ("lz4i", 16384), (3, 65536), (2, 65536), (1, 16384), (4, 16384),
                      ("zstds", 65536)]:
        got = run(base, sc, B, repeat, 56 << 20, packed=True, raw_codec=codec, raw_batch_bytes=bb)
        assert got[1] == ref[1] and got[2] == ref[2], codec
        assert np.array_equal(got[0], ref[0]), codec
        assert got[3] == ref[3] == expected_positions(base, repeat)


def test_objects_half_a_million_records_round_trip_the_generated_numbers():
    """JSON objects at a size the oracle would need minutes for: the numbers the
    generator wrote (`repr` of float64, i.e. shortest round-trip text) must come back bit for bit in
    float64, in schema order whatever the member order, and as their float32 rounding in float32;
    the final commit must equal the log's positions."""
    import json
    rng = np.random.default_rng(35)
    n, F = 50000, 6
    names = ["f%d" % k for k in range(F)]
    table = rng.uniform(-1e6, 1e6, (n, F))
    table[:, 5] = np.round(table[:, 5]) + 0.0                        # whole numbers written as ints (no -0.0)
    vals = []
    for i in range(n):
        order = names if i % 3 else names[::-1]
        d = {"id": "r%07d" % i}
        for k in order:
            j = int(k[1:])
            d[k] = int(table[i, j]) if j == 5 else float(table[i, j])
        d["tags"] = [i % 5, {"k": None}]
        vals.append(json.dumps(d).encode())
    i = np.arange(n)
    base = sk.RecordLog(vals, i % 16, i // 16)
    repeat, B = 10, 512

    def rows(dtype):
        class DS(KafkaDataset):
            payload_schema = JsonObject(names, dtype)
            device_slot_bytes = 24 << 20

            @classmethod
            def new_consumer(cls, *a, **k):
                k.pop("enable_auto_commit", None)
                return sk.KafkaConsumer(*a, log=base, repeat=repeat, **k)

        ds = DS("topic", group_id="prop")
        out = torch.cat([b for b in auto_commit(DataLoader(ds, batch_size=B))]).cpu().numpy()
        committed = dict(ds._consumer.committed_log[-1])
        ds.close()
        return out, committed

    got64, committed = rows(torch.float64)
    assert got64.shape == (n * repeat, F)
    assert np.array_equal(got64.view(np.uint64), np.tile(table, (repeat, 1)).view(np.uint64))
    assert committed == expected_positions(base, repeat)
    got32, _ = rows(torch.float32)
    assert np.array_equal(got32.view(np.uint32), np.tile(table.astype(np.float32), (repeat, 1)).view(np.uint32))
