"""RecordBatch v2 bytes that were NOT produced by this repository's encoder.

`record_batch_data_v2` of kafka-python's own test suite (test/record/test_records.py of
kafka-python >= 1.4, the dependency the reference pins at `setup.py:9`): three batches written by
a real producer (offset 0: value b"123"; offsets 1, 2: value b""; offset 3: value b"123"; keys
null, no headers).  kafka-python's test reads them with `MemoryRecords(b"".join(...) + b"\\x00" * 4)`
and expects 218 valid bytes of 222 and exactly those records.

kafka-python is on no machine this repository runs on, so the strings were entered by hand.  Their
integrity does not rest on that: every batch carries its own CRC-32C over everything behind the
checksum field, and `tests/test_record_batch_cpu.py` recomputes it with an implementation that is
independent of this repository (TensorBoard's pure-Python `crc32c`), so one wrong byte would fail.
They pin the oracle's (and through it the device's) reading of the header layout, the checksum
range, the zig-zag varints and the record framing to bytes of the reference's dependency.
"""

BATCHES = [
    # baseOffset 0, one record: value b"123", timestamp 1503229838908
    (b'\x00\x00\x00\x00\x00\x00\x00\x00\x00\x00\x00;\x00\x00\x00\x01\x02\x03'
     b'\x18\xa2p\x00\x00\x00\x00\x00\x00\x00\x00\x01]\xff{\x06<\x00\x00\x01]'
     b'\xff{\x06<\xff\xff\xff\xff\xff\xff\xff\xff\xff\xff\xff\xff\xff\xff'
     b'\x00\x00\x00\x01\x12\x00\x00\x00\x01\x06123\x00'),
    # baseOffset 1, two records: value b"" twice (the second with a timestamp delta of two varint bytes)
    (b'\x00\x00\x00\x00\x00\x00\x00\x01\x00\x00\x00@\x00\x00\x00\x02\x02\xc8'
     b'\\\xbd#\x00\x00\x00\x00\x00\x01\x00\x00\x01]\xff|\xddl\x00\x00\x01]\xff'
     b'|\xde\x14\xff\xff\xff\xff\xff\xff\xff\xff\xff\xff\xff\xff\xff\xff\x00'
     b'\x00\x00\x02\x0c\x00\x00\x00\x01\x00\x00\x0e\x00\xd0\x02\x02\x01\x00'
     b'\x00'),
    # baseOffset 3, one record: value b"123"
    (b'\x00\x00\x00\x00\x00\x00\x00\x03\x00\x00\x00;\x00\x00\x00\x02\x02.\x0b'
     b'\x85\xb7\x00\x00\x00\x00\x00\x00\x00\x00\x01]\xff|\xe7\x9d\x00\x00\x01]'
     b'\xff|\xe7\x9d\xff\xff\xff\xff\xff\xff\xff\xff\xff\xff\xff\xff\xff\xff'
     b'\x00\x00\x00\x01\x12\x00\x00\x00\x01\x06123\x00'),
]

#: (offset, value) of every record, in order
RECORDS = [(0, b"123"), (1, b""), (2, b""), (3, b"123")]
#: CRC-32C stored in each batch (bytes 17..20)
CRCS = [0x0318A270, 0xC85CBD23, 0x2E0B85B7]
#: what kafka-python's test feeds MemoryRecords: the batches and four stray bytes
STREAM = b"".join(BATCHES) + b"\x00" * 4
VALID_BYTES, SIZE_IN_BYTES = 218, 222
