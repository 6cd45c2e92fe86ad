"""CPU tests of the record-batch step: the oracle restatement of the RecordBatch v2
format against the format's published known answers, the stub broker's encoder against the
oracle, and the library's host-side CRC-32C.  No GPU."""
import random
import struct

import numpy as np
import pytest

import helpers  # noqa: F401  (sys.path)
from oracle import record_batch as orb
from torchkafka_b200.testing import record_batch as rb
from torchkafka_b200.testing import stub_kafka as sk
from torchkafka_b200.testing import synthetic as syn

# RFC 3720 (iSCSI) appendix B.4 CRC-32C test vectors + the classic check value
CRC_KAT = [
    (b"123456789", 0xE3069283),
    (bytes(32), 0x8A9136AA),
    (b"\xff" * 32, 0x62A8AB43),
    (bytes(range(32)), 0x46DD794E),
    (bytes(range(31, -1, -1)), 0x113FDB5C),
    (b"", 0x00000000),
]


@pytest.mark.parametrize("data,want", CRC_KAT)
def test_crc32c_known_answers(data, want):
    assert orb.crc32c(data) == want
    assert rb.crc32c(data) == want


def test_crc32c_host_library_matches_oracle_on_random_data_and_continues():
    rng = random.Random(3)
    for n in [1, 7, 8, 9, 63, 64, 65, 1000, 4099]:
        data = bytes(rng.getrandbits(8) for _ in range(n))
        assert rb.crc32c(data) == orb.crc32c(data)
        k = rng.randrange(n + 1)
        assert rb.crc32c(data[k:], rb.crc32c(data[:k])) == orb.crc32c(data)      # incremental form
        # unaligned start
        arr = np.frombuffer(b"x" + data, dtype=np.uint8)[1:]
        assert rb.crc32c(arr) == orb.crc32c(data)


# zig-zag varints: the examples of the protobuf encoding guide (sint32/sint64)
VARINT_KAT = [(0, b"\x00"), (-1, b"\x01"), (1, b"\x02"), (-2, b"\x03"), (2147483647, b"\xfe\xff\xff\xff\x0f"),
              (-2147483648, b"\xff\xff\xff\xff\x0f"), (150, b"\xac\x02"), (-64, b"\x7f"), (64, b"\x80\x01")]


@pytest.mark.parametrize("value,wire", VARINT_KAT)
def test_varint_known_answers(value, wire):
    assert rb.zigzag_varint(value) == wire
    assert orb.read_varint(wire, 0, len(wire)) == (value, len(wire))


def test_varint_errors():
    with pytest.raises(orb.CorruptRecordError):
        orb.read_varint(b"\x80\x80", 0, 2)                     # runs off the end
    with pytest.raises(orb.CorruptRecordError):
        orb.read_varint(b"\xff" * 11, 0, 11)                   # more than 10 bytes


def test_batch_layout_matches_the_protocol_guide():
    """Field offsets of the 61-byte header (Kafka protocol guide, Record Batch, magic 2)."""
    blob = rb.encode_batch(1234567890123, [b"ab", b"cde"])
    base, length, epoch, magic, crc = struct.unpack_from(">qiibI", blob, 0)
    attrs, last, t0, t1, pid, pepoch, seq, count = struct.unpack_from(">hiqqqhii", blob, 21)
    assert (base, magic, attrs, last, count) == (1234567890123, 2, 0, 1, 2)
    assert length == len(blob) - 12 and (pid, pepoch, seq) == (-1, -1, -1)
    assert crc == orb.crc32c(blob[21:])
    # first record: length 8, attributes 0, timestampDelta 0, offsetDelta 0, key null, value "ab", no headers
    assert blob[61:70] == bytes([2 * 8, 0, 0, 0, 1, 2 * 2]) + b"ab" + b"\x00"


def test_roundtrip_with_keys_headers_nulls_and_offset_gaps():
    rng = random.Random(9)
    values = [None if rng.random() < 0.1 else bytes(rng.getrandbits(8) for _ in range(rng.choice([0, 1, 5, 200, 5000])))
              for _ in range(40)]
    keys = [None if rng.random() < 0.5 else b"k" * rng.randrange(0, 300) for _ in values]
    headers = [[(b"h", b"v" * rng.randrange(3)), (b"n", None)] if rng.random() < 0.3 else [] for _ in values]
    deltas = sorted(rng.sample(range(200), len(values)))        # compacted topic: gaps
    blob = rb.encode_batch(1000, values, offset_deltas=deltas, keys=keys, headers=headers)
    recs = orb.split_blob(blob, 5)
    assert [r.value for r in recs] == values
    assert [r.offset for r in recs] == [1000 + d for d in deltas]
    assert all(r.partition == 5 for r in recs)


def test_fetch_offset_partial_tail_control_and_errors():
    a = rb.encode_batch(10, [b"a", b"b", b"c"])
    b = rb.encode_batch(13, [b"d"])
    ctl = rb.encode_batch(14, [b"\x00\x00\x00\x00"], attributes=0x20)
    c = rb.encode_batch(15, [b"e"])
    blob = a + b + ctl + c
    assert [r.value for r in orb.split_blob(blob, 0)] == [b"a", b"b", b"c", b"d", b"e"]
    assert [r.offset for r in orb.split_blob(blob, 0, fetch_offset=12)] == [12, 13, 15]
    assert [r.offset for r in orb.split_blob(blob, 0, offset_delta=100)] == [110, 111, 112, 113, 115]
    for cut in (1, 11, 12, 30, len(c) - 1):                     # response cut inside the last batch
        assert [r.value for r in orb.split_blob(blob[:len(blob) - cut], 0)] == [b"a", b"b", b"c", b"d"]
    bad = bytearray(blob)
    bad[len(a) + 63] ^= 1                                      # inside the second batch's record
    with pytest.raises(orb.CorruptRecordError):
        orb.split_blob(bytes(bad), 0)
    with pytest.raises(orb.UnsupportedBatch):
        orb.split_blob(rb.encode_batch(0, [b"x"], attributes=5), 0)         # no such codec
    v1 = bytearray(a)
    v1[16] = 1
    with pytest.raises(orb.CorruptRecordError):
        orb.split_blob(bytes(v1), 0)


@pytest.mark.parametrize("layout,chunk", [("chunked", 7), ("roundrobin", 1), ("chunked", 500)])
def test_poll_raw_serves_the_log_in_iteration_order(layout, chunk):
    log = syn.make_log(2, 400, seed=2, partitions=5, layout=layout, chunk=chunk)
    c = sk.KafkaConsumer("topic", log=log, repeat=2, raw_batch_bytes=6000)
    got = []
    while True:
        f = c.poll_raw(50000, 64)
        if f is None:
            break
        assert len({p for p, *_ in f.blobs}) == len(f.blobs)    # a partition at most once per fetch
        for p, data, fetch_offset, delta in f.blobs:
            got += orb.split_blob(bytes(data), p, fetch_offset, delta)
    it = sk.KafkaConsumer("topic", log=log, repeat=2)
    want = list(it)
    assert [(r.partition, r.offset, r.value) for r in got] == [(r.partition, r.offset, r.value) for r in want]
    assert c._positions == it._positions


def test_poll_raw_after_partial_consumption_returns_the_batch_whole():
    log = syn.make_log(2, 60, seed=4, partitions=1, layout="chunked", chunk=60)
    c = sk.KafkaConsumer("topic", log=log, raw_batch_bytes=1 << 20)      # one batch
    first = [next(c) for _ in range(25)]
    f = c.poll_raw(1 << 22, 1000)
    (p, data, fetch_offset, delta), = f.blobs
    assert fetch_offset == first[-1].offset + 1
    recs = orb.split_blob(bytes(data), p, fetch_offset, delta)
    assert [r.offset for r in recs] == list(range(25, 60))
    assert len(orb.split_blob(bytes(data), p, 0, delta)) == 60


# ---------------------------------------------------------------- LZ4 (codec 3)
def test_lz4_oracle_decodes_frames_of_an_independent_compressor():
    """The oracle's LZ4 frame/block restatement against frames written by pyarrow's liblz4:
    incompressible data (stored blocks), long overlapping matches, text, several blocks."""
    import os
    rng = random.Random(6)
    cases = [b"", b"a", os.urandom(200_000), b"abc" * 70_000, b"\x00" * 300_000,
             json_like(rng, 150_000), os.urandom(10) * 9_000 + os.urandom(5_000)]
    for data in cases:
        assert orb.lz4_frame_decode(rb.lz4_frame(data)) == data


def json_like(rng, n):
    import json
    out = json.dumps([rng.randrange(128000) for _ in range(n // 6)]).encode()
    return out[:n]


def test_lz4_errors():
    z = bytearray(rb.lz4_frame(b"hello world, hello world, hello world" * 20))
    with pytest.raises(orb.CorruptRecordError):
        orb.lz4_frame_decode(bytes(z[:-1]))
    bad = bytearray(z)
    bad[0] ^= 1
    with pytest.raises(orb.CorruptRecordError):
        orb.lz4_frame_decode(bytes(bad))
    with pytest.raises(orb.CorruptRecordError):                  # offset beyond the start of the output
        orb.lz4_frame_decode(b"\x04\x22\x4d\x18\x60\x40\x82" + (7).to_bytes(4, "little") + b"\x10a\x05\x00\x50bc" + bytes(4))


def test_lz4_batches_roundtrip_through_the_oracle():
    log = syn.make_log(3, 120, seed=8, partitions=3, layout="chunked", chunk=20)
    c = sk.KafkaConsumer("topic", log=log, raw_batch_bytes=30000, raw_codec=3)
    got = []
    while True:
        f = c.poll_raw(1 << 22, 1000)
        if f is None:
            break
        for p, data, fetch_offset, delta in f.blobs:
            assert bytes(data[21:23]) == b"\x00\x03"            # attributes: LZ4
            got += orb.split_blob(bytes(data), p, fetch_offset, delta)
    want = list(sk.KafkaConsumer("topic", log=log))
    assert [(r.partition, r.offset, r.value) for r in got] == [(r.partition, r.offset, r.value) for r in want]
    assert log.encoded(30000, 3).nbytes < 0.9 * log.encoded(30000).nbytes


def test_independent_block_frames_are_valid_lz4_frames():
    """The stub's Kafka-style frame writer (independent blocks, optional xxHash checksums) against
    liblz4 itself (pyarrow), which verifies the header and content checksums."""
    import os
    import pyarrow as pa
    rng = random.Random(12)
    for data in [json_like(rng, 200_000), os.urandom(70_000), b"xyz" * 50_000]:
        for kw in ({}, {"block_checksum": True, "content_checksum": True}):
            z = rb.lz4_frame_independent(data, **kw)
            assert pa.Codec("lz4").decompress(z, decompressed_size=len(data)).to_pybytes() == data
            assert orb.lz4_frame_decode(z) == data


def test_snappy_oracle_decodes_blocks_of_an_independent_compressor():
    import os
    rng = random.Random(13)
    cases = [b"", b"a", os.urandom(100_000), b"abc" * 70_000, b"\x00" * 200_000, json_like(rng, 150_000)]
    for data in cases:
        assert orb.snappy_decode(rb.snappy_raw(data)) == data
        assert orb.snappy_decode(rb.snappy_xerial(data)) == data
    # format_description.txt: literal "abcd" then a copy of 4 bytes at offset 4 (tag 01, length 4)
    assert orb.snappy_block_decode(bytes([8, 0x0C]) + b"abcd" + bytes([0x01, 0x04])) == b"abcdabcd"
    with pytest.raises(orb.CorruptRecordError):
        orb.snappy_block_decode(bytes([8, 0x0C]) + b"abcd" + bytes([0x01, 0x05]))     # offset before the start
    with pytest.raises(orb.CorruptRecordError):
        orb.snappy_block_decode(bytes([9, 0x0C]) + b"abcd" + bytes([0x01, 0x04]))     # length mismatch


@pytest.mark.parametrize("codec", [2, "snappy_raw", "lz4i", "lz4ic", 1, "gzipx", 4, "zstds"])
def test_compressed_batches_roundtrip_through_the_oracle(codec):
    log = syn.make_log(2, 200, seed=9, partitions=2, layout="chunked", chunk=50)
    c = sk.KafkaConsumer("topic", log=log, raw_batch_bytes=100000, raw_codec=codec)
    got = []
    while True:
        f = c.poll_raw(1 << 22, 1000)
        if f is None:
            break
        for p, data, fetch_offset, delta in f.blobs:
            got += orb.split_blob(bytes(data), p, fetch_offset, delta)
    want = list(sk.KafkaConsumer("topic", log=log))
    assert [(r.partition, r.offset, r.value) for r in got] == [(r.partition, r.offset, r.value) for r in want]


# ---------------------------------------------------------------- bytes of kafka-python's own test suite
def _independent_crc32c(data: bytes) -> int:
    """A CRC-32C that shares no code with this repository (TensorBoard's record writer)."""
    from tensorboard.compat.tensorflow_stub.pywrap_tensorflow import crc32c
    return crc32c(data)


def test_kafka_python_golden_batches_are_intact_and_every_crc_agrees():
    from golden import kafka_python_record_batches as kp
    assert len(kp.STREAM) == kp.SIZE_IN_BYTES and sum(len(b) for b in kp.BATCHES) == kp.VALID_BYTES
    for blob, want in zip(kp.BATCHES, kp.CRCS):
        assert struct.unpack(">I", blob[17:21])[0] == want
        assert struct.unpack(">i", blob[8:12])[0] + 12 == len(blob)             # batchLength
        covered = blob[21:]                                                       # attributes .. end
        assert _independent_crc32c(covered) == want      # the hand-entered bytes are the producer's bytes
        assert orb.crc32c(covered) == want               # oracle restatement
        assert rb.crc32c(covered) == want                # library host CRC (tk_crc32c)


def test_oracle_reads_kafka_python_golden_batches_like_kafka_python_does():
    from golden import kafka_python_record_batches as kp
    recs = orb.split_blob(kp.STREAM, 5)                  # the 4 stray bytes are an incomplete batch: ignored
    assert [(r.partition, r.offset, r.value) for r in recs] == [(5, o, v) for o, v in kp.RECORDS]
    # a consumer positioned inside the second batch gets the batch whole and skips below its position
    recs = orb.split_blob(b"".join(kp.BATCHES[1:]), 5, fetch_offset=2)
    assert [(r.offset, r.value) for r in recs] == [(2, b""), (3, b"123")]
    # the repository's own encoder writes the same bytes for the same records (timestamps aside):
    # header layout and record framing agree field by field where both are defined
    mine = rb.encode_batch(3, [b"123"])
    theirs = kp.BATCHES[2]
    assert len(mine) == len(theirs) and mine[:8] == theirs[:8] and mine[8:12] == theirs[8:12]
    assert mine[16] == theirs[16] == 2 and mine[21:27] == theirs[21:27] and mine[57:] == theirs[57:]
