"""The fast lane of the JSON array kernels (csrc/tk_fastlane.cuh), compiled for the host, against
CPython on the same literals: float(text) for float64, numpy.float32(float(text)) for float32 (the
reference rounds twice: json.loads, then torch.tensor), int(text) for the integer targets.  The lane
must accept exactly  [' ']['-'](0|[1-9]d*)['.'d+][('e'|'E')['+'|'-']d{1,3}]  within its digit limits and
nothing else; what
it does not accept goes to the scalar grammar, which has its own tests.  Runs without a GPU."""
import ctypes
import json
import os
import random
import re
import struct
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
KINDS = {"f32": (0, np.float32), "f64": (1, np.float64), "i32": (2, np.int32), "i64": (3, np.int64)}
FORM = re.compile(rb"^ ?(-?)(0|[1-9][0-9]*)(?:\.([0-9]+))?(?:[eE]([+-]?)([0-9]{1,3}))?$")


@pytest.fixture(scope="module")
def lib():
    subprocess.run(["make", "-C", os.path.join(HERE, "native")], check=True, capture_output=True)
    return ctypes.CDLL(os.path.join(HERE, "native", "_build", "libtk_host_decode.so"))


def lay_out(literals, sep, prefix):
    """The literals as the elements of one array behind `prefix`: buffer, pos[], stop[]."""
    buf = bytearray(prefix) + b"["
    pos, stop = [], []
    for k, lit in enumerate(literals):
        if k:
            buf += sep[:1]
            pos.append(len(buf))
            buf += sep[1:]
        else:
            pos.append(len(buf))
        buf += lit
        stop.append(len(buf))
    buf += b"]" + b"\0" * 64
    return bytes(buf), np.asarray(pos, np.uint32), np.asarray(stop, np.uint32)


def run_lane(lib, buf, pos, stop, kind):
    code, dtype = KINDS[kind]
    out = np.zeros(len(pos), dtype)
    acc = np.zeros(len(pos), np.uint8)
    lib.tk_host_fastlane_many(buf, pos.ctypes.data_as(ctypes.c_void_p), stop.ctypes.data_as(ctypes.c_void_p),
                              ctypes.c_uint32(len(pos)), code, out.ctypes.data_as(ctypes.c_void_p),
                              acc.ctypes.data_as(ctypes.c_void_p))
    return out, acc.astype(bool)


def lane_takes(text: bytes, stop: int, kind: str) -> bool:
    """What the lane is specified to accept (text = the element's bytes, leading space included)."""
    m = FORM.match(text)
    if not m or len(text) > 30:
        return False
    ipart, frac, exp = m.group(2), m.group(3), m.group(5)
    # the window is anchored at the end of the mantissa
    mant_end = stop - (len(text) - m.end(3 if frac is not None else 2))
    if mant_end < 20:
        return False
    if frac is None:
        ok = len(ipart) <= 8
    else:
        body = len(ipart) + 1 + len(frac)
        ok = body <= 20 or (ipart == b"0" and (body == 21 or (body == 22 and frac[:1] == b"0")))
    if not ok:
        return False
    if kind in ("i32", "i64"):
        return frac is None and exp is None
    return True


def expected(text: bytes, kind: str):
    s = text.decode().strip()
    with np.errstate(over="ignore"):
        if kind == "f64":
            return np.float64(float(int(s)) if re.fullmatch(r"-?\d+", s) else float(s))
        if kind == "f32":
            return np.float32(float(int(s)) if re.fullmatch(r"-?\d+", s) else float(s))
    return KINDS[kind][1](int(s))


def check(lib, literals, kind, sep=b", ", prefix=b"x" * 24, must_take=None):
    buf, pos, stop = lay_out(literals, sep, prefix)
    out, acc = run_lane(lib, buf, pos, stop, kind)
    n_taken = 0
    for k, lit in enumerate(literals):
        text = buf[pos[k]:stop[k]]
        want_take = lane_takes(text, int(stop[k]), kind)
        assert bool(acc[k]) == want_take, (text, kind, bool(acc[k]))
        if acc[k]:
            n_taken += 1
            want = expected(text, kind)
            assert out[k].tobytes() == want.tobytes(), (text, kind, out[k], want)
    if must_take is not None:
        assert n_taken >= must_take * len(literals), (n_taken, len(literals))
    return n_taken


@pytest.mark.parametrize("kind", ["f32", "f64"])
@pytest.mark.parametrize("sep", [b", ", b","])
def test_what_json_dumps_writes(lib, kind, sep):
    rng = random.Random(11)
    for scale in (1e3, 1.0, 1e7, 30.0, 1e-3, 1e-9, 1e18, 1e300):
        vals = [rng.uniform(-scale, scale) for _ in range(30000)]
        lits = [json.dumps(v).encode() for v in vals]
        for shift in range(4):
            check(lib, lits, kind, sep=sep, prefix=b"y" * (24 + shift), must_take=0.999)


@pytest.mark.parametrize("kind", ["f32", "f64", "i32", "i64"])
def test_integers_and_mixed(lib, kind):
    rng = random.Random(5)
    lits = [str(rng.randrange(-10 ** rng.randrange(1, 12), 10 ** rng.randrange(1, 12))).encode() for _ in range(30000)]
    lits += [b"0", b"-0", b"7", b"-7", b"99999999", b"-99999999", b"100000000", b"0.0", b"-0.0", b"0.5", b"-0.5"]
    check(lib, lits, kind)
    check(lib, lits, kind, sep=b",", prefix=b"z" * 21)


def test_float32_rounding_boundaries(lib):
    """Doubles at and around the halfway points between two float32 values, written with 17
    significant digits and as their exact decimal expansions cut to the lane's 19 digits."""
    rng = random.Random(3)
    lits = []
    for _ in range(20000):
        f = np.float32(rng.uniform(-1e3, 1e3))
        bits = struct.unpack("<I", struct.pack("<f", f))[0]
        up = struct.unpack("<f", struct.pack("<I", bits + 1))[0]
        mid = (float(f) + float(up)) / 2.0                       # exact in double
        for delta in (0, 1, -1, 3, -3, 9, -9):
            q = struct.unpack("<d", struct.pack("<q", struct.unpack("<q", struct.pack("<d", mid))[0] + delta))[0]
            lits.append(repr(q).encode())
            lits.append(("%.17g" % q).encode())
        from decimal import Decimal
        exact = format(Decimal(mid), "f")
        for cut in (17, 18, 19, 20):
            digits = exact.replace("-", "")
            keep = cut + (1 if "." in digits[:cut + 1] else 0)
            lits.append((("-" if exact[0] == "-" else "") + digits[:keep]).rstrip(".").encode())
    lits = [x for x in lits if b"e" not in x and b"inf" not in x and b"nan" not in x]
    check(lib, lits, "f32")
    check(lib, lits, "f64")


def test_digit_limits_and_shapes(lib):
    rng = random.Random(8)
    lits = []
    for _ in range(60000):
        ni = rng.randrange(1, 14)
        nf = rng.randrange(0, 22)
        i = "0" if rng.random() < 0.2 else str(rng.randrange(10 ** (ni - 1), 10 ** ni))
        f = "".join(rng.choice("0123456789") for _ in range(nf))
        s = ("-" if rng.random() < 0.5 else "") + i + ("." + f if nf else "")
        lits.append(s.encode())
    for kind in ("f32", "f64", "i64"):
        check(lib, lits, kind)
        check(lib, lits, kind, sep=b",", prefix=b"9" * 23)


def test_everything_else_is_left_to_the_grammar(lib):
    odd = [b".5", b"5.", b"-.5", b"-", b"--1", b"+1", b"1..2", b"1.2.3", b"1e5", b"1.5E-3", b"1.5e+3", b"01", b"00.5",
           b"-01.5", b" 1", b"1 ", b"\t1", b"1.5 ", b"NaN", b"Infinity", b"-Infinity", b"true", b"null", b"\"1\"", b"1/2",
           b"1./5", b"1.\x2f5", b"1:5", b"0x10", b"1_000", b"1.5f", b"12345678901234567890", b"1.2345678901234567890123",
           b"", b"1,5"[:1] + b"\x00", b"\xff1", b"1\xff", b"1.\xff", b"\x80.5", b"3.1\x1f4", b"3\x1f.14",
           b"1e", b"1e+", b"1e-", b"e5", b"-e5", b"1.e5", b".5e5", b"1e5.5", b"1e1234", b"1e0005", b"1ee5", b"1e+-5", b"1E5", b"1E+05",
           b"1e-400", b"1e400", b"-1e400", b"0e0", b"-0e5", b"0.0e-3", b"1.5e300", b"9.999999999999999e22", b"1e23", b"2.5e-324",
           b"4.9e-324", b"2.2250738585072014e-308", b"1.7976931348623157e+308", b"1.7976931348623159e+308",
           b"0.000123456789012345", b"0.0001234567890123456", b"0.00012345678901234567", b"0.000123456789012345678",
           b"0.1234567890123456789", b"0.12345678901234567890", b"0.01234567890123456789", b"1.01234567890123456789",
           b"-0.00012345678901234567", b"0.00000000000000000001", b"0.00000000000000000000"]
    rng = random.Random(2)
    for _ in range(20000):
        n = rng.randrange(1, 24)
        odd.append(bytes(rng.choice(b"0123456789.-+eE ,/\x1f\x1e\x2f1234567890") for _ in range(n)).replace(b",", b"."))
    odd = [x for x in odd if b"," not in x]
    for kind in ("f32", "f64", "i32"):
        check(lib, odd, kind, sep=b",")
        check(lib, odd, kind, sep=b",", prefix=b"1" * 22)


def test_near_the_start_of_the_buffer(lib):
    """An element that ends less than 20 bytes into the buffer cannot be anchored at its end."""
    for prefix in (b"", b"[", b"x" * 5, b"x" * 17):
        check(lib, [b"1.5", b"2.25", b"-3.125", b"4.0625", b"5.03125", b"6.5", b"7.75"], "f64", prefix=prefix)
