"""Kafka RecordBatch v2 encoder for the in-memory stub broker (tests and benchmarks).

What a broker would put into the `records` field of a fetch response for one partition:
uncompressed, non-transactional batches (Apache Kafka protocol guide, "Record Batch", magic 2).
Record keys are null, headers empty, timestamps zero -- the reference's `_process` hook only
reads `record.value`, `.partition` and `.offset`.

The checksum comes from the library's host-side `tk_crc32c` (include/torchkafka_b200.h).
"""
from __future__ import annotations

import ctypes as C
import struct
from typing import List, Optional, Sequence

import numpy as np

HEADER_BYTES = 61
CRC_FROM = 21


def zigzag_varint(v: int) -> bytes:
    """Signed integer -> zig-zag base-128 varint."""
    u = (v << 1) ^ (v >> 63)
    u &= (1 << 64) - 1
    out = bytearray()
    while True:
        b = u & 0x7F
        u >>= 7
        if u:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


_VARINT_SMALL = [zigzag_varint(i) for i in range(4096)]
_VARINT_M1 = zigzag_varint(-1)


def _vi(v: int) -> bytes:
    return _VARINT_SMALL[v] if 0 <= v < 4096 else zigzag_varint(v)


def crc32c(data, crc: int = 0) -> int:
    from .. import _native
    lib = _native.load()
    buf = np.frombuffer(data, dtype=np.uint8) if not isinstance(data, np.ndarray) else data
    return int(lib.tk_crc32c(C.c_void_p(buf.ctypes.data), C.c_uint64(buf.size), C.c_uint32(crc)))


def encode_record(offset_delta: int, value: Optional[bytes], key: Optional[bytes] = None,
                  timestamp_delta: int = 0, headers: Sequence = ()) -> bytes:
    body = bytearray(b"\x00")                                  # attributes
    body += _vi(timestamp_delta) if timestamp_delta >= 0 else zigzag_varint(timestamp_delta)
    body += _vi(offset_delta)
    if key is None:
        body += _VARINT_M1
    else:
        body += _vi(len(key)) + key
    if value is None:
        body += _VARINT_M1
    else:
        body += _vi(len(value))
        body += value
    body += _vi(len(headers))
    for hk, hv in headers:
        body += _vi(len(hk)) + hk
        body += _VARINT_M1 if hv is None else _vi(len(hv)) + hv
    return _vi(len(body)) + bytes(body)


def lz4_frame(data: bytes) -> bytes:
    """LZ4 frame of `data` as liblz4 writes it by default: block-dependent, 64 KiB blocks, no
    checksums (pyarrow's codec: test infrastructure, as the oracle is)."""
    import pyarrow as pa
    return pa.Codec("lz4").compress(data, asbytes=True)


def lz4_frame_independent(data: bytes, block_bytes: int = 65536, block_checksum: bool = False,
                          content_checksum: bool = False, short_blocks: Sequence[int] = ()) -> bytes:
    """LZ4 frame with independent blocks, the way Kafka's Java client writes it
    (KafkaLZ4BlockOutputStream: FLG = version 01 | block independence, BD = 64 KiB).
    `short_blocks`: sizes of leading blocks that end before they are full (a writer that flushes;
    the format allows any block up to the maximum)."""
    import pyarrow as pa
    import xxhash
    raw = pa.Codec("lz4_raw")
    flg = 0x60 | (0x10 if block_checksum else 0) | (0x04 if content_checksum else 0)
    bd = {65536: 0x40, 262144: 0x50, 1 << 20: 0x60, 4 << 20: 0x70}[block_bytes]
    head = bytes([flg, bd])
    out = bytearray(b"\x04\x22\x4d\x18" + head + bytes([(xxhash.xxh32(head, seed=0).intdigest() >> 8) & 0xFF]))
    cuts = [0]
    for n in short_blocks:
        if cuts[-1] + n < len(data):
            cuts.append(cuts[-1] + min(int(n), block_bytes))
    while cuts[-1] < len(data):
        cuts.append(min(cuts[-1] + block_bytes, len(data)))
    for a, b in zip(cuts, cuts[1:]):
        chunk = data[a:b]
        z = raw.compress(chunk, asbytes=True)
        if len(z) >= len(chunk):                               # stored block
            out += struct.pack("<I", len(chunk) | 0x80000000) + chunk
            z = chunk
        else:
            out += struct.pack("<I", len(z)) + z
        if block_checksum:
            out += struct.pack("<I", xxhash.xxh32(z, seed=0).intdigest())
    out += b"\x00\x00\x00\x00"
    if content_checksum:
        out += struct.pack("<I", xxhash.xxh32(data, seed=0).intdigest())
    return bytes(out)


def snappy_raw(data: bytes) -> bytes:
    import pyarrow as pa
    return pa.Codec("snappy").compress(data, asbytes=True)


def snappy_xerial(data: bytes, chunk_bytes: int = 32768) -> bytes:
    """snappy-java framing (what Kafka's Java client writes): magic, version, compatible version,
    then {int32 length, raw snappy block} per 32 KiB of input."""
    out = bytearray(b"\x82SNAPPY\x00" + struct.pack(">ii", 1, 1))
    for k in range(0, len(data), chunk_bytes):
        z = snappy_raw(data[k:k + chunk_bytes])
        out += struct.pack(">i", len(z)) + z
    return bytes(out)


def gzip_member(data: bytes, level: int = 6, name: bytes = None, extra: bytes = None, comment: bytes = None,
                header_crc: bool = False) -> bytes:
    """One gzip member (RFC 1952); without the optional fields this is what
    java.util.zip.GZIPOutputStream (Kafka's Java client) writes."""
    import zlib
    flg = (2 if header_crc else 0) | (4 if extra is not None else 0) | (8 if name is not None else 0) | \
        (16 if comment is not None else 0)
    head = bytearray(b"\x1f\x8b\x08" + bytes([flg]) + b"\x00\x00\x00\x00\x00\x00")
    if extra is not None:
        head += struct.pack("<H", len(extra)) + extra
    if name is not None:
        head += name + b"\x00"
    if comment is not None:
        head += comment + b"\x00"
    if header_crc:
        head += struct.pack("<H", zlib.crc32(bytes(head)) & 0xFFFF)
    co = zlib.compressobj(level, zlib.DEFLATED, -15)
    body = co.compress(data) + co.flush()
    return bytes(head) + body + struct.pack("<II", zlib.crc32(data) & 0xFFFFFFFF, len(data) & 0xFFFFFFFF)


def zstd_frame(data: bytes, level: int = 3, content_size: bool = False) -> bytes:
    """One Zstandard frame (libzstd through pyarrow).  Kafka's Java client compresses through a
    stream and so writes no Frame_Content_Size: that is the default here -- the field is taken
    out of the one-shot frame and a window descriptor put in place of the single-segment flag
    (the frame header carries no checksum)."""
    import pyarrow as pa
    z = pa.Codec("zstd", compression_level=level).compress(data, asbytes=True)
    if content_size:
        return z
    fhd = z[4]
    single, fcs, did = (fhd >> 5) & 1, fhd >> 6, fhd & 3
    pos = 5 + (0 if single else 1) + (4 if did == 3 else did)
    fcs_bytes = (1 if single else 0) if fcs == 0 else (2, 4, 8)[fcs - 1]
    return z[:4] + bytes([fhd & 0x07, (27 - 10) << 3]) + z[5 + (0 if single else 1):pos] + z[pos + fcs_bytes:]


#: codec argument of encode_batch -> (attribute bits, compressor)
CODECS = {
    4: (4, zstd_frame), "zstd": (4, zstd_frame), "zstds": (4, lambda d: zstd_frame(d, 1, content_size=True)),
    "zstd19": (4, lambda d: zstd_frame(d, 19)),
    1: (1, gzip_member), "gzip": (1, gzip_member), "gzip0": (1, lambda d: gzip_member(d, level=0)),
    "gzipx": (1, lambda d: gzip_member(d, 9, name=b"records.bin", extra=b"\x01\x02\x03", comment=b"c", header_crc=True)),
    3: (3, lz4_frame), "lz4": (3, lz4_frame), "lz4i": (3, lz4_frame_independent),
    "lz4ic": (3, lambda d: lz4_frame_independent(d, block_checksum=True, content_checksum=True)),
    "lz4if": (3, lambda d: lz4_frame_independent(d, short_blocks=(20000, 65536, 7))),
    2: (2, snappy_xerial), "snappy": (2, snappy_xerial), "snappy_raw": (2, snappy_raw),
}


def encode_batch(base_offset: int, values: Sequence[Optional[bytes]], attributes: int = 0,
                 offset_deltas: Optional[Sequence[int]] = None, keys: Optional[Sequence] = None,
                 headers: Optional[Sequence] = None, crc: Optional[int] = None, codec: int = 0) -> bytes:
    """One RecordBatch v2.  `offset_deltas` default to 0, 1, 2, ... (a compacted topic has gaps).
    `codec` (a key of CODECS: 3 / "lz4", "lz4i", "lz4ic", "lz4if", 2 / "snappy", "snappy_raw", 1 / "gzip", "gzip0",
    "gzipx", 4 / "zstd", "zstds", "zstd19") compresses the
    records and sets the attribute bits."""
    n = len(values)
    deltas = list(range(n)) if offset_deltas is None else list(offset_deltas)
    recs = b"".join(encode_record(deltas[k], values[k], None if keys is None else keys[k], 0,
                                  () if headers is None else headers[k]) for k in range(n))
    last = deltas[-1] if n else 0
    if codec:
        bits, compress = CODECS[codec]
        recs = compress(recs)
        attributes |= bits
    tail = struct.pack(">hiqqqhii", attributes, last, 0, 0, -1, -1, -1, n) + recs    # from byte 21 on
    if crc is None:
        crc = crc32c(tail)
    batch_len = (HEADER_BYTES - 12) + len(recs)
    return struct.pack(">qiibI", base_offset, batch_len, 0, 2, crc) + tail


class EncodedLog:
    """A RecordLog as record batches: every run of same-partition records of the log becomes one
    or more batches of at most `batch_bytes` of payload (at least one record)."""

    def __init__(self, log, batch_bytes: int = 16384, codec: int = 0):
        values = log._byte_values            # noqa: SLF001 (test helper of the same package)
        parts = log.partitions
        offs = log.offsets
        n = len(log)
        chunks: List[bytes] = []
        first_rec, part_of, pos_of = [], [], []
        pos = 0
        i = 0
        while i < n:
            p = int(parts[i])
            j, size = i, 0
            # a batch: consecutive records of one partition with consecutive offsets
            while j < n and int(parts[j]) == p and (j == i or (int(offs[j]) == int(offs[j - 1]) + 1
                                                               and size + len(values[j]) <= batch_bytes)):
                size += len(values[j])
                j += 1
            blob = encode_batch(int(offs[i]), values[i:j], codec=codec)
            chunks.append(blob)
            first_rec.append(i)
            part_of.append(p)
            pos_of.append(pos)
            pos += len(blob)
            i = j
        self.data = np.frombuffer(b"".join(chunks), dtype=np.uint8) if chunks else np.zeros(0, np.uint8)
        self.first_rec = np.asarray(first_rec + [n], dtype=np.int64)      # batch b holds records [first_rec[b], first_rec[b+1])
        self.partition = np.asarray(part_of, dtype=np.int32)
        self.start = np.asarray(pos_of + [pos], dtype=np.int64)
        self.n_batches = len(part_of)
        # run_end[b]: index of the first batch after b that belongs to another partition
        nbt = self.n_batches
        change = np.flatnonzero(self.partition[1:] != self.partition[:-1]) + 1 if nbt > 1 else np.zeros(0, np.int64)
        ends = np.concatenate([change, [nbt]]).astype(np.int64)
        self.run_end = ends[np.searchsorted(ends, np.arange(nbt), side="right")] if nbt else np.zeros(0, np.int64)

    def pin(self):
        """Move the segment into page-locked memory (a consumer that receives into pinned arenas):
        DMA straight out of it then runs at the link's full rate.  Needs CUDA; idempotent."""
        if not getattr(self, "_pinned", None):
            import torch
            t = torch.from_numpy(self.data.copy()).pin_memory()
            self._pinned = t                         # keeps the allocation alive
            self.data = t.numpy()
        return self

    @property
    def nbytes(self) -> int:
        return int(self.start[-1])
