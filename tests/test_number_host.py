"""The device-side decoder (tk_number.cuh / tk_json.cuh), compiled for the host, against
CPython: json.loads + float() / torch.tensor on the same bytes.  Runs without a GPU."""
import ctypes
import json
import os
import random
import struct
import subprocess

import numpy as np
import pytest
import torch

from golden_cases import KNOWN_ANSWER_PAYLOADS

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")


@pytest.fixture(scope="module")
def lib():
    subprocess.run(["make", "-C", os.path.join(HERE, "native")], check=True, capture_output=True)
    return ctypes.CDLL(os.path.join(HERE, "native", "_build", "libtk_host_decode.so"))


def strtod_many(lib, strs):
    bs = [s.encode() for s in strs]
    buf = b"".join(bs)
    starts = np.zeros(len(bs) + 1, np.uint32)
    starts[1:] = np.cumsum([len(b) for b in bs])
    bits = np.zeros(len(bs), np.uint64)
    st = np.zeros(len(bs), np.uint8)
    lib.tk_host_strtod_many(buf, starts.ctypes.data_as(ctypes.c_void_p), len(bs),
                            bits.ctypes.data_as(ctypes.c_void_p), st.ctypes.data_as(ctypes.c_void_p))
    return bits, st


def gen_decimal(rng):
    k = rng.randrange(9)
    if k == 0:
        return repr(rng.uniform(-1e3, 1e3))
    if k == 1:
        return repr(struct.unpack("<d", struct.pack("<Q", rng.getrandbits(64) & 0x7FEFFFFFFFFFFFFF))[0])
    if k == 2:
        ip = "".join(rng.choice("0123456789") for _ in range(rng.randint(1, 40))).lstrip("0") or "0"
        fp = "".join(rng.choice("0123456789") for _ in range(rng.randint(1, 40)))
        return ip + "." + fp + ("e%d" % rng.randint(-340, 300) if rng.random() < 0.5 else "")
    if k == 3:       # exact midpoint between two adjacent doubles, and neighbours of it
        from decimal import Decimal, getcontext
        from fractions import Fraction
        getcontext().prec = 1300
        b = rng.getrandbits(63) & 0x7FEFFFFFFFFFFFFF
        if rng.random() < 0.2:
            b &= 0x001FFFFFFFFFFFFF
        x = struct.unpack("<d", struct.pack("<Q", b))[0]
        y = struct.unpack("<d", struct.pack("<Q", b + 1))[0]
        if y == float("inf"):
            return "1.5"
        mid = (Fraction(x) + Fraction(y)) / 2
        s = format(Decimal(mid.numerator) / Decimal(mid.denominator), "f")
        if "." not in s:
            s += ".0"
        r = rng.random()
        if r < 0.33:
            return s
        if r < 0.66:
            return s + "1"
        t = s.rstrip("0")
        if t[-1] == ".":
            return s
        return t[:-1] + str(int(t[-1]) - 1) + "9" * rng.randint(1, 30)
    if k == 4:
        return "%d" % rng.getrandbits(rng.randint(1, 1100))
    if k == 5:
        return "%de%d" % (rng.getrandbits(rng.randint(1, 70)), rng.randint(-345, 310))
    if k == 6:
        return "0." + "0" * rng.randint(0, 330) + "%d" % rng.getrandbits(rng.randint(1, 80))
    if k == 7:
        return "%d.%de%d" % (rng.randint(0, 9), rng.getrandbits(60),
                             rng.choice([-324, -323, -322, -308, -307, 308, 307, -309, -310]))
    return repr(rng.random() * 10 ** rng.randint(-320, 308))


def test_decimal_to_double_matches_cpython_zero_ulp(lib):
    rng = random.Random(20261106)
    strs = [gen_decimal(rng) for _ in range(120000)]
    bits, st = strtod_many(lib, strs)
    bad = []
    for s, b, c in zip(strs, bits, st):
        is_int = not any(ch in s for ch in ".eE")
        if is_int:      # a Python int in the reference: OverflowError if it rounds to infinity
            try:
                f = float(int(s))
                exp, expst = (0 if f == 0 else struct.unpack("<Q", struct.pack("<d", f))[0]), 0
            except OverflowError:
                exp, expst = 0, 2
        else:
            exp, expst = struct.unpack("<Q", struct.pack("<d", float(s)))[0], 0
        if c != expst or (expst == 0 and int(b) != exp):
            bad.append((s[:60], int(c), hex(int(b)), hex(exp)))
    assert not bad, bad[:5]


FLAT = False     # which decomposition of the record grammar the helper below exercises


def decode(lib, payload: bytes, kind: int, max_len: int = 64):
    dt = [np.float32, np.float64, np.int32, np.int64][kind]
    out = np.zeros(max_len, dt)
    n = ctypes.c_int(0)
    fn = lib.tk_host_decode_flat if FLAT else lib.tk_host_decode
    st = fn(payload, len(payload), kind, max_len, out.ctypes.data_as(ctypes.c_void_p), ctypes.byref(n))
    return st, out[:min(n.value, max_len)], n.value


@pytest.fixture(params=["runs", "comma-split"])
def decomposition(request):
    global FLAT
    FLAT = request.param == "comma-split"
    yield request.param
    FLAT = False


def test_record_grammar_and_conversions_match_reference_known_answers(lib, decomposition):
    with open(os.path.join(GOLD, "known_answers.json")) as f:
        stored = json.load(f)["answers"]
    for a in stored:
        payload = a["payload"].encode("latin-1")
        for kind, dt in enumerate(("float32", "float64", "int32", "int64")):
            want = a[dt]
            st, vals, n = decode(lib, payload, kind)
            if "error" in want or "shape" in want:
                # the reference raised, or produced a non-1-D tensor (outside the schema)
                if "shape" in want and want["shape"] == [0]:
                    assert st == 0 and n == 0
                else:
                    assert st != 0, (a["payload"], dt)
                continue
            assert st == 0, (a["payload"], dt, st)
            if "bits" in want:
                it = np.uint32 if dt == "float32" else np.uint64
                mask = 0xFFFFFFFF if dt == "float32" else 0xFFFFFFFFFFFFFFFF
                assert [int(x) for x in vals.view(it)] == [b & mask for b in want["bits"]], (a["payload"], dt)
            else:
                assert [int(x) for x in vals] == want["values"], (a["payload"], dt)


def test_random_records_match_json_loads(lib, decomposition):
    rng = random.Random(99)
    for _ in range(3000):
        toks = [gen_decimal(rng) if rng.random() < 0.7 else str(rng.randint(-10**6, 10**6))
                for _ in range(rng.randint(0, 12))]
        sep = rng.choice([", ", ",", " ,\t", "\n, "])
        payload = (rng.choice(["", " ", "\r\n"]) + "[" + sep.join(toks) + "]" + rng.choice(["", " \n"])).encode()
        if rng.random() < 0.15:      # corrupt it
            cut = rng.randrange(len(payload) + 1)
            payload = payload[:cut] + rng.choice([b"", b",", b"]", b"x", b"-", b"."]) + payload[cut + 1:]
        try:
            ref = torch.tensor(json.loads(payload), dtype=torch.float64)
            ok = ref.dim() == 1
        except Exception:                                 # noqa: BLE001
            ok = False
        st, vals, n = decode(lib, payload, 1)
        if not ok:
            assert st != 0, payload
        else:
            assert st == 0, (payload, st)
            assert np.array_equal(vals.view(np.uint64), ref.numpy().view(np.uint64)), payload
