"""GPU parity of the record-batch input: the same log served as RecordBatch v2
blobs (`poll_raw`, CRC-32C + varint framing on the device) must give bit-identical batches and
commits to the oracle iterating ConsumerRecords; malformed batches surface like kafka-python's
CorruptRecordException would, at the batch that contains them."""
import json

import numpy as np
import pytest
import torch

from helpers import assert_same, make_dataset, run_gpu, run_oracle
from oracle import record_batch as orb
from torchkafka_b200.engine import SlotEngine
from torchkafka_b200.schema import ImageU8, JsonArray, PayloadError, RawBytes, TokenList
from torchkafka_b200.testing import record_batch as rb
from torchkafka_b200.testing import synthetic as syn
from torchkafka_b200.testing.stub_kafka import RecordLog

pytestmark = pytest.mark.gpu


def check_raw(log, schema, batch_size, batch_bytes=16384, codec=0, **kw):
    gpu = run_gpu(log, schema, batch_size, device_input="raw",
                  consumer_kwargs={"raw_batch_bytes": batch_bytes, "raw_codec": codec}, **kw)
    cpu = run_oracle(log, schema, batch_size)
    assert_same(gpu, cpu, schema)
    return gpu


@pytest.mark.parametrize("layout,chunk,batch_bytes", [("chunked", 500, 16384), ("chunked", 37, 3000),
                                                      ("roundrobin", 1, 16384), ("chunked", 500, 1 << 20)])
def test_cfg2_from_record_batches(layout, chunk, batch_bytes):
    log = syn.make_log(2, 6000, seed=21, partitions=8, layout=layout, chunk=chunk)
    check_raw(log, JsonArray(torch.float32, max_len=256, min_size=32), 256, batch_bytes=batch_bytes)


def test_cfg2_small_slots_many_fetches():
    log = syn.make_log(2, 3000, seed=22, partitions=4, layout="chunked", chunk=100)
    check_raw(log, JsonArray(torch.float32, max_len=256, min_size=32), 64, slot_bytes=200_000, batch_bytes=8192)


def test_cfg3_tokens_from_record_batches():
    log = syn.make_log(3, 800, seed=23, partitions=8, layout="chunked", chunk=50)
    check_raw(log, TokenList(torch.int32, max_len=2048), 32, batch_bytes=65536)


def test_cfg5_small_records_from_record_batches():
    log = syn.make_log(5, 20000, seed=24, partitions=16, layout="chunked", chunk=500)
    check_raw(log, RawBytes((64,), torch.uint8), 256)


def test_cfg4_images_from_record_batches():
    log = syn.make_log(4, 24, seed=25, partitions=4, layout="chunked", chunk=3)
    check_raw(log, ImageU8(224, 224, 3), 8, batch_bytes=1 << 20, slot_bytes=24 * 150528 + (1 << 20))


# ---------------------------------------------------------------- the split itself, blob by blob
def split_on_gpu(blobs, slot_bytes=1 << 20, slot_records=4096, schema=None):
    """blobs: [(partition, bytes, fetch_offset, offset_delta)] -> (engine, n, nbytes) after the split."""
    eng = SlotEngine(schema or RawBytes((1,), torch.uint8), 4, slot_bytes=slot_bytes, slot_records=slot_records,
                     n_slots=2, device=0, n_partitions=16)
    hb, _ = eng.host_views(0)
    pos, desc = 0, []
    for p, data, fo, delta in blobs:
        hb[pos:pos + len(data)] = np.frombuffer(data, dtype=np.uint8)
        desc.append((pos, len(data), p, fo, delta))
        pos += len(data)
    n_host = eng.submit_recordbatches(0, pos, desc)
    n, nbytes, nbatches = eng.recordbatch_result(0)
    assert n == n_host
    return eng, n, nbytes, nbatches


def device_records(eng, n, nbytes):
    """(start, partition with flags, offset) of every record header of slot 0's device twin, plus
    the closing entry."""
    from torchkafka_b200 import _native
    import ctypes as C
    recs = []
    for i in range(n + 1):
        r = _native.TkRec()
        _native.check(eng.lib.tk_slot_record(eng.ctx, 0, i, C.byref(r)))
        recs.append((int(r.start), int(r.partition), int(r.offset)))
    return recs


def test_split_headers_match_the_oracle():
    rng = np.random.default_rng(5)
    vals = [bytes(rng.integers(0, 256, int(k), dtype=np.uint8)) for k in rng.integers(0, 700, 300)]
    blob_a = b"".join(rb.encode_batch(100 + 50 * j, vals[50 * j:50 * j + 50]) for j in range(3))
    blob_b = rb.encode_batch(7, vals[150:300], keys=[b"key%d" % i if i % 3 else None for i in range(150)],
                             headers=[[(b"h", b"x" * (i % 5))] if i % 4 == 0 else [] for i in range(150)])
    cut = blob_a + rb.encode_batch(250, [b"never complete"])[:-3]          # response ends mid-batch
    blobs = [(3, cut, 120, 0), (5, blob_b, 0, 1000)]
    eng, n, nbytes, nbatches = split_on_gpu(blobs)
    want = []
    for p, data, fo, delta in blobs:
        # the device keeps skipped records as flagged entries; the oracle drops them
        want += [(r.partition, r.offset, r.value) for r in orb.split_blob(data, p, 0, delta)]
    assert nbatches == 4 and n == len(want) == 300
    recs = device_records(eng, n, nbytes)
    from torchkafka_b200 import _native
    kept_bytes = 0
    for (start, part, off), (p, o, v), nxt in zip(recs, want, recs[1:]):
        dropped = p == 3 and o < 120
        assert part == (p | (0x40000000 if dropped else 0)) and off == o
        assert nxt[0] - start == (0 if dropped else len(v))
        kept_bytes += 0 if dropped else len(v)
    assert recs[-1][0] == nbytes == kept_bytes
    eng.close()


def run_blobs(blobs, schema, batch_size):
    """Drive the public API with a consumer that serves exactly these blobs in one fetch."""
    from torchkafka_b200.testing import stub_kafka as sk

    class OneFetch(sk.KafkaConsumer):
        def poll_raw(self, max_bytes, max_records):
            if getattr(self, "_served", False):
                return None
            self._served = True
            return sk.RawFetch([(p, np.frombuffer(d, dtype=np.uint8), fo, delta) for p, d, fo, delta in blobs])

    from torchkafka_b200 import KafkaDataset, auto_commit
    from torch.utils.data import DataLoader

    class DS(KafkaDataset):
        payload_schema = schema
        device_input = "raw"

        @classmethod
        def new_consumer(cls, *a, **k):
            k.pop("enable_auto_commit", None)
            return OneFetch(*a, log=RecordLog([], [], []), **k)

    ds = DS("topic", group_id="g")
    out = [b for b in auto_commit(DataLoader(ds, batch_size=batch_size))]
    commits = list(ds._consumer.committed_log)
    ds.close()
    return out, commits


def jvals(lo, hi):
    return [json.dumps([float(i), i + 0.5]).encode() for i in range(lo, hi)]


def test_fetch_offset_skips_records_and_control_batches_are_invisible():
    blob = (rb.encode_batch(10, jvals(10, 16)) + rb.encode_batch(16, [b"\x00\x00\x00\x01"], attributes=0x20) +
            rb.encode_batch(17, jvals(17, 20)))
    out, commits = run_blobs([(2, blob, 13, 0)], JsonArray(torch.float32, fixed_len=2), 3)
    got = torch.cat([b.cpu() for b in out])
    assert got[:, 0].tolist() == [13.0, 14.0, 15.0, 17.0, 18.0, 19.0]
    assert commits == [{2: 16}, {2: 20}]


def test_fetch_offset_skips_records_of_raw_and_struct_schemas_too():
    """Records below the fetch position are not errors for the collate kernels either."""
    from torchkafka_b200.schema import StructBytes
    vals = [bytes([i]) * 8 for i in range(10, 20)]
    blob = rb.encode_batch(10, vals[:6]) + rb.encode_batch(16, vals[6:])
    out, commits = run_blobs([(2, blob, 13, 0)], RawBytes((8,), torch.uint8), 3)
    assert torch.cat([b.cpu() for b in out])[:, 0].tolist() == [13, 14, 15, 16, 17, 18, 19]
    assert commits == [{2: 16}, {2: 19}, {2: 20}]
    out, _ = run_blobs([(2, blob, 13, 0)], StructBytes((("a", torch.uint8, (3,)), ("b", torch.uint8, (5,)))), 4)
    assert torch.cat([b.b.cpu() for b in out])[:, 4].tolist() == [13, 14, 15, 16, 17, 18, 19]


def test_objects_from_record_batches():
    from torchkafka_b200.schema import JsonObject
    fields = ("x", "y", "z", "w", "vx", "vy", "vz", "t")
    vals = syn.object_values(4000, 26, fields)
    part, off = syn._coords(4000, 8, "chunked", 50)         # noqa: SLF001
    log = RecordLog(vals, part, off)
    check_raw(log, JsonObject(fields, torch.float32), 128)
    check_raw(log, JsonObject(("t", "x"), torch.int64), 100, batch_bytes=4096, codec="lz4i")


@pytest.mark.parametrize("damage,status", [("crc", 7), ("varint", 7), ("magic", 7), ("count", 7), ("codec", 4), ("null", 4)])
def test_broken_batches_raise_at_their_batch(damage, status):
    good = rb.encode_batch(0, jvals(0, 4))
    bad = bytearray(rb.encode_batch(4, jvals(4, 8)))
    if damage == "crc":
        bad[70] ^= 0x40
    elif damage == "magic":
        bad[16] = 1
    elif damage == "codec":
        bad = bytearray(rb.encode_batch(4, jvals(4, 8), attributes=5))        # no such codec
    elif damage == "null":
        bad = bytearray(rb.encode_batch(4, [jvals(4, 5)[0], None, jvals(6, 7)[0]]))
    elif damage == "varint":                                  # framing broken, CRC recomputed to match
        bad[61] = 0xFF
        bad[17:21] = rb.crc32c(bytes(bad[21:])).to_bytes(4, "big")
    elif damage == "count":
        bad[57:61] = (9).to_bytes(4, "big")
        bad[17:21] = rb.crc32c(bytes(bad[21:])).to_bytes(4, "big")
    blob = good + bytes(bad) + rb.encode_batch(8, jvals(8, 12))
    from torchkafka_b200.testing import stub_kafka as sk   # noqa: F401
    with pytest.raises(PayloadError) as ei:
        run_blobs([(1, blob, 0, 0)], JsonArray(torch.float32, fixed_len=2), 2)
    assert ei.value.status == status and ei.value.partition == 1
    assert ei.value.offset == (5 if damage == "null" else 4 if damage != "magic" else 0)


def test_large_and_tiny_batches_and_unaligned_blobs():
    """CRC over > 512-byte rows, exact multiples, short tails; blobs starting at odd addresses."""
    rng = np.random.default_rng(8)
    blobs, want = [], []
    for p, nrec, size in [(0, 1, 0), (1, 1, 1), (2, 3, 450), (3, 1, 512 - 69), (4, 2, 1024), (5, 40, 3000), (6, 1, 300000)]:
        vals = [bytes(rng.integers(0, 256, size, dtype=np.uint8)) for _ in range(nrec)]
        blobs.append((p, rb.encode_batch(1000 * p, vals), 0, 0))
        want += vals
    eng, n, nbytes, nbatches = split_on_gpu(blobs, slot_bytes=2 << 20)
    assert (n, nbatches, nbytes) == (len(want), len(blobs), sum(map(len, want)))
    recs = device_records(eng, n, nbytes)
    assert all(not (part & 0x70000000) for _, part, _ in recs[:-1])      # every CRC verified
    assert [b[0] - a[0] for a, b in zip(recs, recs[1:])] == [len(v) for v in want]
    eng.close()


def test_slot_too_small_is_reported():
    blob = rb.encode_batch(0, [b"x" * 100] * 50)
    from torchkafka_b200 import _native
    with pytest.raises(_native.NativeError, match="more records in the slot than slot_records"):
        split_on_gpu([(0, blob, 0, 0)], slot_records=10)


# ---------------------------------------------------------------- LZ4-compressed batches (§8f-2)
@pytest.mark.parametrize("codec", [3, "lz4i", "lz4ic", "lz4if", 2, "snappy_raw", 1, "gzipx", "gzip0", 4, "zstds", "zstd19"])
@pytest.mark.parametrize("cfg,n,batch_bytes", [(2, 3000, 16384), (2, 3000, 200000), (3, 600, 65536), (5, 20000, 16384)])
def test_compressed_batches_match_the_oracle(cfg, n, batch_bytes, codec):
    """LZ4 frames with dependent blocks (liblz4's default), with independent blocks and checksums
    (Kafka's Java client), with blocks that end before they are full ("lz4if": the blocks of a frame
    are placed by the frame's maximal block size and moved together afterwards), snappy in xerial
    framing and raw.  cfg2/cfg3: text (matches; several
    blocks / chunks per batch with batch_bytes 200000); cfg5: random bytes (stored blocks, tiny
    matches).  gzip: one member per batch, dynamic and fixed Huffman blocks, with the optional
    header fields ("gzipx"), stored blocks only ("gzip0").  zstd: frames without a content size
    (the streaming form, size pass on the device), with one ("zstds"), level 19 ("zstd19")."""
    log = syn.make_log(cfg, n, seed=31 + cfg, partitions=8, layout="chunked", chunk=100)
    schema = {2: JsonArray(torch.float32, max_len=256, min_size=32), 3: TokenList(torch.int32, max_len=2048),
              5: RawBytes((64,), torch.uint8)}[cfg]
    check_raw(log, schema, 128, batch_bytes=batch_bytes, codec=codec)


@pytest.mark.parametrize("ca,cb", [(3, "lz4i"), (2, "snappy_raw"), ("lz4ic", 2), (1, "gzipx"), ("gzip", 3), (4, "zstds"), ("zstd19", 1)])
def test_codecs_mixed_with_plain_batches_and_long_overlapping_matches(ca, cb):
    vals_a = [json.dumps([1.5] * 200).encode() for _ in range(40)]          # one long period-5 match
    vals_b = jvals(0, 50)
    wide = [json.dumps([float(i % 7)] * 120).encode() for i in range(300)]   # > 64 KiB of records
    blob = (rb.encode_batch(0, vals_a, codec=ca) + rb.encode_batch(40, vals_b) + rb.encode_batch(90, wide, codec=cb))
    out, commits = run_blobs([(4, blob, 0, 0)], JsonArray(torch.float32, max_len=256), 64)
    want = [json.loads(v) for v in vals_a + vals_b + wide]
    got = []
    for b in out:
        vals, lens = b.values.cpu(), b.lengths.cpu()
        got += [vals[i, :int(lens[i])].tolist() for i in range(len(lens))]
    assert got == want and commits[-1] == {4: 390}


@pytest.mark.parametrize("damage", ["magic", "offset", "truncated", "size"])
def test_broken_lz4_frames_raise_at_their_batch(damage):
    good = rb.encode_batch(0, jvals(0, 4), codec=3)
    bad = bytearray(rb.encode_batch(4, [json.dumps([0.5] * 60).encode()] * 4, codec=3))
    if damage == "magic":
        bad[61] ^= 0xFF
    elif damage == "offset":                         # first match reaches before the start of the output
        k = 61 + 7 + 4                               # frame header, block size -> first token
        lit = bad[k] >> 4
        assert lit < 15
        bad[k + 1 + lit:k + 3 + lit] = (60000).to_bytes(2, "little")
    elif damage == "truncated":
        bad[61 + 7:61 + 11] = (len(bad)).to_bytes(4, "little")      # block longer than the batch
    elif damage == "size":
        bad[-4:] = b"\x01\x00\x00\x00"                              # garbage instead of the end mark
    bad[17:21] = rb.crc32c(bytes(bad[21:])).to_bytes(4, "big")
    with pytest.raises(PayloadError) as ei:
        run_blobs([(1, good + bytes(bad), 0, 0)], JsonArray(torch.float32, max_len=64), 2)
    assert ei.value.status == 7 and ei.value.partition == 1 and ei.value.offset == 4


@pytest.mark.parametrize("damage", ["magic", "method", "blocktype", "isize", "truncated", "two-members"])
def test_broken_gzip_batches_raise_at_their_batch(damage):
    good = rb.encode_batch(0, jvals(0, 4), codec=1)
    bad = bytearray(rb.encode_batch(4, [json.dumps([0.5 + i] * 60).encode() for i in range(4)], codec=1))
    status = 7
    if damage == "magic":
        bad[61] ^= 0xFF
    elif damage == "method":
        bad[63] = 7
        status = 4
    elif damage == "blocktype":
        bad[61 + 10] |= 0x06                         # first block header: reserved block type 3
    elif damage == "isize":
        bad[-4:] = (int.from_bytes(bad[-4:], "little") + 1).to_bytes(4, "little")
    elif damage == "truncated":
        del bad[-12:-8]                              # the stream loses its last bytes
        bad[8:12] = (len(bad) - 12).to_bytes(4, "big")
    elif damage == "two-members":                    # gzip allows it, this decoder follows one member
        bad += rb.gzip_member(b"")
        bad[8:12] = (len(bad) - 12).to_bytes(4, "big")
    bad[17:21] = rb.crc32c(bytes(bad[21:])).to_bytes(4, "big")
    with pytest.raises(PayloadError) as ei:
        run_blobs([(1, good + bytes(bad), 0, 0)], JsonArray(torch.float32, max_len=64), 2)
    assert ei.value.status == status and ei.value.partition == 1 and ei.value.offset == 4


def test_zstd_frames_without_content_size_placed_by_bound_or_sized_first(monkeypatch):
    """Frames without a content size: placed by their upper bound when the slot's bounds fit the
    decompressed buffer (which grows for them), otherwise sized by a first pass; both must agree
    with the oracle.  Small slots with many small batches exceed the bound budget on their own."""
    log = syn.make_log(2, 3000, seed=36, partitions=8, layout="chunked", chunk=100)
    schema = JsonArray(torch.float32, max_len=256, min_size=32)
    check_raw(log, schema, 128, batch_bytes=16384, codec=4)                       # bounds fit
    check_raw(log, schema, 128, batch_bytes=2048, codec=4, slot_bytes=300_000)    # 128 KiB bound per 2.6 KB batch: size pass
    monkeypatch.setenv("TK_ZSTD_SIZE_PASS", "1")
    check_raw(log, schema, 128, batch_bytes=16384, codec=4)
    check_raw(log, schema, 128, batch_bytes=200000, codec="zstd19")


@pytest.mark.parametrize("damage", ["magic", "blocktype", "truncated", "dictionary", "size"])
@pytest.mark.parametrize("codec", [4, "zstds"])
def test_broken_zstd_batches_raise_at_their_batch(damage, codec):
    good = rb.encode_batch(0, jvals(0, 4), codec=codec)
    bad = bytearray(rb.encode_batch(4, [json.dumps([0.5 + i] * 60).encode() for i in range(4)], codec=codec))
    status = 7
    head = 61 + (6 if codec == 4 else 5 + (1 if bad[61 + 4] >> 6 == 0 else 2))       # first block header
    if damage == "magic":
        bad[61] ^= 0xFF
    elif damage == "blocktype":
        bad[head] |= 0x06                            # reserved block type 3
    elif damage == "truncated":
        del bad[-3:]
        bad[8:12] = (len(bad) - 12).to_bytes(4, "big")
    elif damage == "dictionary":
        if codec != 4:
            pytest.skip("header layout of the sized frame")
        bad[61 + 4] |= 0x01
        bad[61 + 6:61 + 6] = b"\x09"
        bad[8:12] = (len(bad) - 12).to_bytes(4, "big")
        status = 4
    elif damage == "size":
        if codec == 4:
            pytest.skip("no content size in this frame")
        bad[61 + 5] ^= 0x01                          # Frame_Content_Size says another length
    bad[17:21] = rb.crc32c(bytes(bad[21:])).to_bytes(4, "big")
    with pytest.raises(PayloadError) as ei:
        run_blobs([(1, good + bytes(bad), 0, 0)], JsonArray(torch.float32, max_len=64), 2)
    assert ei.value.status == status and ei.value.partition == 1 and ei.value.offset == 4


@pytest.mark.parametrize("damage", ["preamble", "offset", "chunk"])
def test_broken_snappy_batches_raise_at_their_batch(damage):
    good = rb.encode_batch(0, jvals(0, 4), codec=2)
    bad = bytearray(rb.encode_batch(4, [json.dumps([0.5] * 60).encode()] * 4, codec=2))
    k = 61 + 16 + 4                                  # xerial header, chunk length -> the snappy block
    if damage == "preamble":
        bad[k] ^= 0x01                               # announces another length than it decodes to
    elif damage == "chunk":
        bad[61 + 16:61 + 20] = (len(bad)).to_bytes(4, "big")
    elif damage == "offset":
        pre = 1 if bad[k] < 0x80 else 2
        tag = bad[k + pre]
        assert tag & 3 == 0 and (tag >> 2) < 60      # a short literal first, then a copy
        c = k + pre + 1 + (tag >> 2) + 1
        assert bad[c] & 3 in (1, 2)
        if bad[c] & 3 == 1:
            bad[c] |= 0xE0
            bad[c + 1] = 0xFF
        else:
            bad[c + 1:c + 3] = b"\xff\xff"
    bad[17:21] = rb.crc32c(bytes(bad[21:])).to_bytes(4, "big")
    with pytest.raises(PayloadError) as ei:
        run_blobs([(1, good + bytes(bad), 0, 0)], JsonArray(torch.float32, max_len=64), 2)
    assert ei.value.status == 7 and ei.value.partition == 1 and ei.value.offset == 4


# ---------------------------------------------------------------- input modes without a gather copy
@pytest.mark.parametrize("mode", ["raw_into", "raw_dma"])
@pytest.mark.parametrize("codec", [0, 3])
def test_cfg2_received_into_the_slot_or_copied_from_the_arena(mode, codec):
    """`poll_raw_into` (the consumer writes the response into the pinned slot) and `raw_dma` (blobs go to
    the device out of the consumer's page-locked arena) give the same batches and commits."""
    log = syn.make_log(2, 5000, seed=31, partitions=8, layout="chunked", chunk=200)
    schema = JsonArray(torch.float32, max_len=256, min_size=32)
    gpu = run_gpu(log, schema, 128, device_input=mode, slot_bytes=1 << 20,
                  consumer_kwargs={"raw_batch_bytes": 16384, "raw_codec": codec})
    assert_same(gpu, run_oracle(log, schema, 128), schema)


@pytest.mark.parametrize("codec", [3, 4])
def test_decompressed_values_beyond_the_slot_are_reported_not_written(codec):
    """Compressed batches whose values need more room than the slot has: nothing may be written
    outside the slot's byte buffer; the batch that does not fit raises PayloadError(status 9).
    (LZ4: the decompressed buffer is as large as the slot and refuses the batch; zstd frames
    without content size are placed by their upper bounds in a buffer that grows, so the values
    reach the gather step, which refuses them.)"""
    n = 400
    vals = [json.dumps([0.0] * 200).encode() for _ in range(n)]          # ~1 KB each, compresses > 20x
    log = RecordLog(vals, np.zeros(n, np.int32), np.arange(n))
    schema = JsonArray(torch.float32, max_len=256, min_size=0)

    from torchkafka_b200.pipeline import IngestPipeline
    ds = make_dataset(log, schema, slot_bytes=64 << 10, device_input="raw",
                      consumer_kwargs={"raw_batch_bytes": 32768, "raw_codec": codec})
    # defeat the fetch budget's expansion estimate: ask for whole slots of compressed bytes
    orig = IngestPipeline._gather_raw

    def greedy(self, slot, scratch):
        self._raw_expansion = 1.0
        return orig(self, slot, scratch)

    IngestPipeline._gather_raw = greedy
    try:
        got, err = 0, None
        try:
            for batch in ds.device_batches(16):
                got += batch[0].shape[0]
        except PayloadError as e:
            err = e
        assert err is not None and err.status == 9, (got, err)
        # everything in front of the batch that did not fit was delivered intact
        assert got > 0 and got % 16 == 0
    finally:
        IngestPipeline._gather_raw = orig
        ds.close()


def test_device_split_of_kafka_python_golden_batches():
    """Bytes of kafka-python's own test suite (tests/golden/kafka_python_record_batches.py), not written
    by this repository's encoder: CRC check, header walk and record framing on the device."""
    from golden import kafka_python_record_batches as kp
    eng, n, nbytes, nbatches = split_on_gpu([(5, kp.STREAM, 0, 0), (6, b"".join(kp.BATCHES[1:]), 102, 100)])
    assert nbatches == 5 and n == 4 + 3
    recs = device_records(eng, n, nbytes)
    want = [(5, o, v, False) for o, v in kp.RECORDS] + [(6, 101, b"", True), (6, 102, b"", False), (6, 103, b"123", False)]
    for (start, part, off), (p, o, v, dropped), nxt in zip(recs, want, recs[1:]):
        assert part == (p | (0x40000000 if dropped else 0)) and off == o, (part, off, p, o)
        assert nxt[0] - start == (0 if dropped else len(v))
    assert nbytes == 3 + 3 + 3
    # the gathered value bytes, read through the slot's device pointer
    ptr, _ = eng.device_views(0)

    class _Slot:
        __cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}

    assert bytes(torch.as_tensor(_Slot(), device="cuda").cpu().numpy()) == b"123" * 3
    eng.close()
