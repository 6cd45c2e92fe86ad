"""The device-side Zstandard decoder (tk_zstd.cuh), compiled for the host, against libzstd
(through pyarrow's codec) on the same frames: every block and literal type, several levels, frames
without a content size (what Kafka's Java client writes), concatenated and skippable frames, and
damaged frames.  Runs without a GPU."""
import ctypes
import json
import os
import random
import struct
import subprocess

import numpy as np
import pyarrow as pa
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
CANARY = 4096


@pytest.fixture(scope="module")
def lib():
    subprocess.run(["make", "-C", os.path.join(HERE, "native")], check=True, capture_output=True)
    return ctypes.CDLL(os.path.join(HERE, "native", "_build", "libtk_host_decode.so"))


def unzstd(lib, z: bytes, cap: int, emit: bool = True):
    out = np.full(cap + CANARY, 0xA5, np.uint8)
    n = ctypes.c_uint32()
    st = lib.tk_host_zstd(z, len(z), out.ctypes.data_as(ctypes.c_void_p), cap, 1 if emit else 0, ctypes.byref(n))
    assert (out[cap:] == 0xA5).all(), "wrote past the output buffer"
    if emit:
        # the batched sequence decoder (tk_zstd_next_batch) with the device's execution order of a
        # batch's copies must agree with the sequence-by-sequence decoder: status and bytes
        out2 = np.full(cap + CANARY, 0xA5, np.uint8)
        n2 = ctypes.c_uint32()
        st2 = lib.tk_host_zstd_batched(z, len(z), out2.ctypes.data_as(ctypes.c_void_p), cap, ctypes.byref(n2))
        assert (out2[cap:] == 0xA5).all(), "batched decoder wrote past the output buffer"
        assert (st2 == 0) == (st == 0), (st, st2)
        if st == 0:
            assert n2.value == n.value and bytes(out2[:n2.value]) == bytes(out[:n.value])
    return st, (bytes(out[:n.value]) if emit else n.value)


def compress(data: bytes, level: int = 3) -> bytes:
    return pa.Codec("zstd", compression_level=level).compress(data, asbytes=True)


def without_content_size(z: bytes) -> bytes:
    """Same frame with the Frame_Content_Size field removed (and a window descriptor instead of
    the single-segment flag): the streaming form."""
    assert z[:4] == b"\x28\xb5\x2f\xfd"
    fhd = z[4]
    single, fcs, did = (fhd >> 5) & 1, fhd >> 6, fhd & 3
    pos = 5 + (0 if single else 1) + (4 if did == 3 else did)
    fcs_bytes = (1 if single else 0) if fcs == 0 else (2, 4, 8)[fcs - 1]
    body = z[pos + fcs_bytes:]
    new_fhd = fhd & 0x07                                   # no content size, not single segment
    return z[:4] + bytes([new_fhd, (27 - 10) << 3]) + z[5 + (0 if single else 1):pos] + body


def payload(rng, kind, n):
    if kind == 0:
        return bytes(rng.getrandbits(8) for _ in range(n))
    if kind == 1:
        return json.dumps([rng.uniform(-1e3, 1e3) for _ in range(n // 18 + 1)]).encode()[:n]
    if kind == 2:
        return b"\0" * n
    if kind == 3:
        return bytes(rng.choice(b"ab") for _ in range(n))
    if kind == 4:
        return (b"hello world " * (n // 12 + 1))[:n]
    if kind == 5:
        return bytes(rng.getrandbits(3) for _ in range(n))
    return json.dumps([rng.randrange(128000) for _ in range(n // 6 + 1)]).encode()[:n]


def test_frames_match_libzstd_with_and_without_content_size(lib):
    rng = random.Random(3)
    for _ in range(500):
        kind = rng.randrange(7)
        n = rng.choice([0, 1, 2, 5, 40, 100, 1000, 5000, 70000, 200000, 400000])
        if kind in (0, 3, 5):
            n = min(n, 70000)
        data = payload(rng, kind, n)
        z = compress(data, rng.choice([1, 1, 3, 3, 6, 12, 19, -5]))
        if rng.random() < 0.5:
            z = without_content_size(z)
        st, got = unzstd(lib, z, len(data))
        assert st == 0 and got == data, (kind, n)
        st, size = unzstd(lib, z, 0, emit=False)               # the size pass
        assert st == 0 and size == len(data)


def test_concatenated_skippable_and_checksummed_frames(lib):
    a, b = b"first frame " * 300, json.dumps(list(range(3000))).encode()
    za, zb = compress(a), without_content_size(compress(b, 9))
    skip = struct.pack("<II", 0x184D2A53, 5) + b"hello"
    z = skip + za + skip + zb
    st, got = unzstd(lib, z, len(a) + len(b))
    assert st == 0 and got == a + b
    # content checksum flag: four more bytes behind the last block (not verified)
    zc = bytearray(za)
    zc[4] |= 0x04
    st, got = unzstd(lib, bytes(zc) + b"\x01\x02\x03\x04" + zb, len(a) + len(b))
    assert st == 0 and got == a + b
    assert unzstd(lib, bytes(zc), len(a))[0] != 0               # flag set, checksum missing


def test_dictionary_frames_are_unsupported_and_garbage_is_an_error(lib):
    z = bytearray(without_content_size(compress(b"abc" * 100)))
    z[4] |= 0x01                                               # Dictionary_ID_flag 1: one byte follows the window byte
    z[6:6] = b"\x07"
    assert unzstd(lib, bytes(z), 300)[0] == 3
    assert unzstd(lib, b"\x00\x01\x02\x03\x04\x05", 64)[0] == 2
    assert unzstd(lib, compress(b"abc" * 100)[:-1], 300)[0] == 2


def test_damaged_frames_never_write_out_of_bounds_and_too_small_buffers_are_errors(lib):
    rng = random.Random(4)
    for _ in range(800):
        data = payload(rng, rng.choice([1, 3, 4, 5, 6]), rng.choice([60, 600, 6000, 150000]))
        z = bytearray(compress(data, rng.choice([1, 3, 9])))
        k = rng.randrange(4)
        if k == 0:
            z[rng.randrange(len(z))] ^= 1 << rng.randrange(8)
        elif k == 1:
            del z[rng.randrange(4, len(z)):]
        elif k == 2:
            z[rng.randrange(len(z))] = rng.getrandbits(8)
        cap = len(data) if k < 3 else rng.randrange(len(data))
        st, got = unzstd(lib, bytes(z), cap)                    # the canary check is the point
        if k == 3:
            assert st != 0
        unzstd(lib, bytes(z), 0, emit=False)
