"""ORACLE (test infrastructure, never on the product path): CPU restatement of the step *before*
the hot path on real topics -- splitting Kafka RecordBatch v2 blobs into records.

The reference delegates this to kafka-python (`kafka/record/default_records.py`,
`kafka/consumer/fetcher.py`; dependency `kafka-python>=2.0.2`, /root/reference/setup.py:9), which
is not vendored in /root/reference and is absent from this image.  PARITY UNPINNED against that
package: this file restates the published wire format (Apache Kafka protocol guide, "Record Batch",
magic 2; KIP-98) and is pinned only by the format's own known answers (CRC-32C check values of
RFC 3720 B.4, zig-zag varint examples of the protobuf encoding guide) in
tests/test_record_batch_cpu.py.

Semantics restated from kafka-python's fetcher:
  * a batch that is not completely inside the blob (fetch responses end mid-batch) is ignored;
  * records whose offset is below the fetch position are skipped (batches are returned whole);
  * CRC mismatch, magic != 2, malformed varints raise CorruptRecordError;
  * compressed batches are outside this restatement and raise UnsupportedBatch;
  * control batches (attributes bit 5) carry no user records and are skipped.
Pure Python, byte at a time: for small cases.
"""
from __future__ import annotations

import struct
from typing import List, NamedTuple, Optional, Tuple

POLY = 0x82F63B78
_T0: List[int] = []
for _b in range(256):
    _c = _b
    for _ in range(8):
        _c = (_c >> 1) ^ (POLY if _c & 1 else 0)
    _T0.append(_c)

HEADER_BYTES = 61          # baseOffset .. recordCount
CRC_FROM = 21              # the checksum covers attributes .. end of batch


class CorruptRecordError(Exception):
    pass


class UnsupportedBatch(Exception):
    pass


def crc32c(data: bytes, crc: int = 0) -> int:
    """CRC-32C (Castagnoli), reflected, init/xorout 0xFFFFFFFF; `crc` continues a previous value."""
    c = crc ^ 0xFFFFFFFF
    for ch in data:
        c = _T0[(c ^ ch) & 0xFF] ^ (c >> 8)
    return c ^ 0xFFFFFFFF


def read_varint(buf: bytes, pos: int, end: int) -> Tuple[int, int]:
    """Zig-zag base-128 varint (32- or 64-bit) at buf[pos:]; returns (value, next position)."""
    shift = 0
    raw = 0
    while True:
        if pos >= end:
            raise CorruptRecordError("varint runs past the end of the batch")
        b = buf[pos]
        pos += 1
        raw |= (b & 0x7F) << shift
        if not b & 0x80:
            break
        shift += 7
        if shift > 63:
            raise CorruptRecordError("varint longer than 10 bytes")
    return (raw >> 1) ^ -(raw & 1), pos


class Record(NamedTuple):
    partition: int
    offset: int
    value: Optional[bytes]


def split_blob(blob: bytes, partition: int, fetch_offset: int = 0, offset_delta: int = 0,
               check_crc: bool = True) -> List[Record]:
    """Records of one partition's blob of a fetch response, in consumption order."""
    out: List[Record] = []
    pos, n = 0, len(blob)
    while n - pos >= 12:
        base_offset, batch_len = struct.unpack_from(">qi", blob, pos)
        end = pos + 12 + batch_len
        if batch_len < HEADER_BYTES - 12:
            raise CorruptRecordError(f"batch length {batch_len} below the header size")
        if end > n:
            break                                   # partial trailing batch: not part of the response
        magic = blob[pos + 16]
        if magic != 2:
            raise CorruptRecordError(f"magic {magic}: only RecordBatch v2 is supported")
        (crc,) = struct.unpack_from(">I", blob, pos + 17)
        if check_crc and crc32c(blob[pos + CRC_FROM:end]) != crc:
            raise CorruptRecordError(f"CRC mismatch in the batch at byte {pos}")
        (attributes,) = struct.unpack_from(">h", blob, pos + 21)
        (count,) = struct.unpack_from(">i", blob, pos + 57)
        if attributes & 0x07:
            raise UnsupportedBatch(f"compression codec {attributes & 7}")
        control = bool(attributes & 0x20)
        p = pos + HEADER_BYTES
        if count < 0:
            raise CorruptRecordError("negative record count")
        for _ in range(count):
            length, p = read_varint(blob, p, end)
            if length < 0 or p + length > end:
                raise CorruptRecordError("record length outside the batch")
            rec_end = p + length
            q = p + 1                                # attributes (int8)
            _ts, q = read_varint(blob, q, rec_end)
            od, q = read_varint(blob, q, rec_end)
            klen, q = read_varint(blob, q, rec_end)
            if klen > 0:
                q += klen
            vlen, q = read_varint(blob, q, rec_end)
            if q > rec_end or (vlen > 0 and q + vlen > rec_end):
                raise CorruptRecordError("record fields outside the record")
            value = None if vlen < 0 else bytes(blob[q:q + vlen])
            offset = base_offset + od + offset_delta
            if not control and offset >= fetch_offset:
                out.append(Record(partition, offset, value))
            p = rec_end
        if p != end:
            raise CorruptRecordError("records do not fill the batch")
        pos = end
    return out
