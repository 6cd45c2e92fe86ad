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
  * LZ4-compressed batches (codec 3) are decompressed with the restatement of the LZ4 frame and
    block formats below (lz4_Frame_format.md / lz4_Block_format.md of the LZ4 project; xxHash
    checksums are skipped, the batch CRC covers the same bytes); snappy batches (codec 2) with the
    restatement of snappy's format_description.txt, raw or in snappy-java's xerial framing (what
    kafka-python's snappy_decode accepts); gzip batches (codec 1) with CPython's own gzip module,
    which is what kafka-python's gzip_decode calls; zstd batches (codec 4) with libzstd (through
    pyarrow's stream reader; kafka-python calls the same library through `zstandard`);
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


def lz4_block_decode(src: bytes, out: bytearray, window_start: int) -> None:
    """Append the decoded LZ4 block `src` to `out`; matches may reach back to out[window_start:]."""
    ip, n = 0, len(src)
    while ip < n:
        token = src[ip]
        ip += 1
        lit = token >> 4
        if lit == 15:
            while True:
                if ip >= n:
                    raise CorruptRecordError("lz4: literal length runs past the block")
                b = src[ip]
                ip += 1
                lit += b
                if b != 255:
                    break
        if ip + lit > n:
            raise CorruptRecordError("lz4: literals run past the block")
        out += src[ip:ip + lit]
        ip += lit
        if ip >= n:
            break                                    # the last sequence has no match
        if ip + 2 > n:
            raise CorruptRecordError("lz4: truncated offset")
        offset = src[ip] | (src[ip + 1] << 8)
        ip += 2
        if offset == 0 or offset > len(out) - window_start:
            raise CorruptRecordError("lz4: offset outside the window")
        mlen = token & 15
        if mlen == 15:
            while True:
                if ip >= n:
                    raise CorruptRecordError("lz4: match length runs past the block")
                b = src[ip]
                ip += 1
                mlen += b
                if b != 255:
                    break
        mlen += 4
        start = len(out) - offset
        for i in range(mlen):                        # byte by byte: the match may overlap its output
            out.append(out[start + i])


def lz4_frame_decode(src: bytes) -> bytes:
    """One LZ4 frame (what Kafka writes per compressed batch) -> the bytes it holds."""
    if len(src) < 7 or src[:4] != b"\x04\x22\x4d\x18":
        raise CorruptRecordError("lz4: bad frame magic")
    flg, _bd = src[4], src[5]
    if flg >> 6 != 1:
        raise CorruptRecordError("lz4: unknown frame version")
    if flg & 1:
        raise UnsupportedBatch("lz4: dictionary frames")
    independent, block_checksum = bool(flg & 0x20), bool(flg & 0x10)
    pos = 6 + (8 if flg & 0x08 else 0) + 1            # content size, header checksum
    out = bytearray()
    while True:
        if pos + 4 > len(src):
            raise CorruptRecordError("lz4: truncated frame")
        (size,) = struct.unpack_from("<I", src, pos)
        pos += 4
        if size == 0:
            break
        raw_block, size = bool(size >> 31), size & 0x7FFFFFFF
        if pos + size > len(src):
            raise CorruptRecordError("lz4: block runs past the frame")
        if raw_block:
            out += src[pos:pos + size]
        else:
            lz4_block_decode(src[pos:pos + size], out, len(out) if independent else 0)
        pos += size + (4 if block_checksum else 0)
    pos += 4 if flg & 0x04 else 0                     # content checksum
    if pos != len(src):
        raise CorruptRecordError("lz4: bytes behind the frame")
    return bytes(out)


def snappy_block_decode(src: bytes) -> bytes:
    """One raw snappy block: varint32 length, then literal / copy elements."""
    n, ip, shift, want = len(src), 0, 0, 0
    while True:
        if ip >= n or shift > 28:
            raise CorruptRecordError("snappy: bad length preamble")
        b = src[ip]
        ip += 1
        want |= (b & 0x7F) << shift
        if not b & 0x80:
            break
        shift += 7
    out = bytearray()
    while ip < n:
        tag = src[ip]
        ip += 1
        kind = tag & 3
        if kind == 0:
            length = tag >> 2
            if length >= 60:
                nb = length - 59
                if ip + nb > n:
                    raise CorruptRecordError("snappy: truncated literal length")
                length = int.from_bytes(src[ip:ip + nb], "little")
                ip += nb
            length += 1
            if ip + length > n:
                raise CorruptRecordError("snappy: literal runs past the block")
            out += src[ip:ip + length]
            ip += length
            continue
        if kind == 1:
            if ip >= n:
                raise CorruptRecordError("snappy: truncated copy")
            length, offset = 4 + ((tag >> 2) & 7), ((tag >> 5) << 8) | src[ip]
            ip += 1
        else:
            nb = 2 if kind == 2 else 4
            if ip + nb > n:
                raise CorruptRecordError("snappy: truncated copy")
            length, offset = 1 + (tag >> 2), int.from_bytes(src[ip:ip + nb], "little")
            ip += nb
        if offset == 0 or offset > len(out):
            raise CorruptRecordError("snappy: offset outside the output")
        start = len(out) - offset
        for i in range(length):
            out.append(out[start + i])
    if len(out) != want:
        raise CorruptRecordError("snappy: length differs from the preamble")
    return bytes(out)


XERIAL_MAGIC = b"\x82SNAPPY\x00"


def snappy_decode(src: bytes) -> bytes:
    """Raw snappy block, or snappy-java ("xerial") framing: magic, two int32, {int32 length, block}*."""
    if src[:8] == XERIAL_MAGIC and len(src) >= 16:
        pos, out = 16, bytearray()
        while pos < len(src):
            if pos + 4 > len(src):
                raise CorruptRecordError("snappy: truncated chunk header")
            (size,) = struct.unpack_from(">I", src, pos)
            pos += 4
            if pos + size > len(src):
                raise CorruptRecordError("snappy: chunk runs past the batch")
            out += snappy_block_decode(src[pos:pos + size])
            pos += size
        return bytes(out)
    return snappy_block_decode(src)


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
        codec = attributes & 0x07
        if codec == 3:
            body = lz4_frame_decode(bytes(blob[pos + HEADER_BYTES:end]))
        elif codec == 2:
            body = snappy_decode(bytes(blob[pos + HEADER_BYTES:end]))
        elif codec == 1:
            import gzip
            try:
                body = gzip.decompress(bytes(blob[pos + HEADER_BYTES:end]))
            except Exception as e:                           # noqa: BLE001 (zlib.error, BadGzipFile, EOFError)
                raise CorruptRecordError(f"gzip: {e}") from e
        elif codec == 4:
            import pyarrow as pa
            try:
                body = pa.CompressedInputStream(pa.BufferReader(bytes(blob[pos + HEADER_BYTES:end])), "zstd").read()
            except Exception as e:                           # noqa: BLE001
                raise CorruptRecordError(f"zstd: {e}") from e
        elif codec:
            raise UnsupportedBatch(f"compression codec {codec}")
        else:
            body = blob[pos + HEADER_BYTES:end]
        control = bool(attributes & 0x20)
        p, body_end = 0, len(body)
        if count < 0:
            raise CorruptRecordError("negative record count")
        for _ in range(count):
            length, p = read_varint(body, p, body_end)
            if length < 0 or p + length > body_end:
                raise CorruptRecordError("record length outside the batch")
            rec_end = p + length
            q = p + 1                                # attributes (int8)
            _ts, q = read_varint(body, q, rec_end)
            od, q = read_varint(body, q, rec_end)
            klen, q = read_varint(body, q, rec_end)
            if klen > 0:
                q += klen
            vlen, q = read_varint(body, q, rec_end)
            if q > rec_end or (vlen > 0 and q + vlen > rec_end):
                raise CorruptRecordError("record fields outside the record")
            value = None if vlen < 0 else bytes(body[q:q + vlen])
            offset = base_offset + od + offset_delta
            if not control and offset >= fetch_offset:
                out.append(Record(partition, offset, value))
            p = rec_end
        if p != body_end:
            raise CorruptRecordError("records do not fill the batch")
        pos = end
    return out
