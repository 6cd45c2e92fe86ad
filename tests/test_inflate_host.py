"""The device-side DEFLATE decoder (tk_inflate.cuh), compiled for the host, against zlib on the
same streams: stored, fixed and dynamic blocks, several blocks per stream, damaged streams.  Runs
without a GPU."""
import ctypes
import json
import os
import random
import subprocess
import zlib

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def lib():
    subprocess.run(["make", "-C", os.path.join(HERE, "native")], check=True, capture_output=True)
    return ctypes.CDLL(os.path.join(HERE, "native", "_build", "libtk_host_decode.so"))


def inflate(lib, z: bytes, cap: int):
    out = np.zeros(max(cap, 1), np.uint8)
    n, used = ctypes.c_uint32(), ctypes.c_uint32()
    st = lib.tk_host_inflate(z, len(z), out.ctypes.data_as(ctypes.c_void_p), cap, ctypes.byref(n), ctypes.byref(used))
    # the batched form (matches collected 32 at a time, copied in the device's least favourable
    # order) must agree with the serial one: status, bytes, input consumed
    out2 = np.zeros(max(cap, 1) + 64, np.uint8)
    out2[max(cap, 1):] = 0xA5
    n2, used2 = ctypes.c_uint32(), ctypes.c_uint32()
    st2 = lib.tk_host_inflate_batched(z, len(z), out2.ctypes.data_as(ctypes.c_void_p), cap, ctypes.byref(n2),
                                      ctypes.byref(used2))
    assert (out2[max(cap, 1):] == 0xA5).all(), "batched decoder wrote past the output buffer"
    assert st2 == st, (st, st2)
    if st == 0:
        assert n2.value == n.value and used2.value == used.value and bytes(out2[:n2.value]) == bytes(out[:n.value])
    return st, bytes(out[:n.value]), used.value


def payload(rng, kind, n):
    if kind == 0:
        return bytes(rng.getrandbits(8) for _ in range(n))
    if kind == 1:
        return json.dumps([rng.uniform(-1e3, 1e3) for _ in range(n // 18 + 1)]).encode()
    if kind == 2:
        return b"\0" * n
    if kind == 3:
        return bytes(rng.choice(b"ab") for _ in range(n))
    if kind == 4:
        return (b"hello world " * (n // 12 + 1))[:n]
    if kind == 6:
        # geometric over the byte values: literal codes of every length up to 15 bits, i.e. codes the
        # first-level table does not hold (the bit-by-bit walk, started behind the table's index bits)
        return bytes(min(255, int(rng.expovariate(0.35))) for _ in range(n))
    return bytes(rng.getrandbits(3) for _ in range(n))


def test_streams_of_every_block_type_match_zlib(lib):
    rng = random.Random(1)
    for _ in range(1500):
        data = payload(rng, rng.randrange(7), rng.choice([0, 1, 2, 5, 100, 1000, 5000, 70000]))
        level = rng.choice([0, 1, 6, 9])
        strategy = rng.choice([zlib.Z_DEFAULT_STRATEGY, zlib.Z_FIXED, zlib.Z_HUFFMAN_ONLY, zlib.Z_RLE, zlib.Z_FILTERED])
        co = zlib.compressobj(level, zlib.DEFLATED, -15, 9, strategy)
        z = co.compress(data)
        if rng.random() < 0.3:
            z += co.flush(zlib.Z_FULL_FLUSH) + co.compress(data[:100])
            data += data[:100]
        z += co.flush()
        st, got, used = inflate(lib, z, len(data))
        assert st == 0 and got == data and used == len(z), (level, strategy, len(data))


def test_damaged_streams_are_rejected_or_decode_like_zlib(lib):
    rng = random.Random(2)
    for _ in range(1500):
        data = payload(rng, rng.choice([1, 3, 4, 5, 6]), rng.choice([50, 500, 5000]))
        z = bytearray(zlib.compress(data, rng.choice([1, 6, 9]), wbits=-15))
        k = rng.randrange(3)
        if k == 0:
            z[rng.randrange(len(z))] ^= 1 << rng.randrange(8)
        elif k == 1:
            del z[rng.randrange(len(z)):]
        else:
            z[rng.randrange(len(z))] = rng.getrandbits(8)
        z = bytes(z)
        d = zlib.decompressobj(-15)
        try:
            want = d.decompress(z)
            ok = d.eof
        except zlib.error:
            ok = False
        st, got, used = inflate(lib, z, max(len(data) * 4, 64))
        if ok and len(want) <= max(len(data) * 4, 64):
            assert st == 0 and got == want and used == len(z) - len(d.unused_data)
        elif not ok:
            assert st != 0, (k, len(z))


def test_output_capacity_is_never_exceeded(lib):
    data = b"abcdefgh" * 1000
    z = zlib.compress(data, 6, wbits=-15)
    for cap in (0, 1, 7, 100, len(data) - 1):
        assert inflate(lib, z, cap)[0] != 0
    assert inflate(lib, z, len(data))[0] == 0
