"""The device-side JSON object scanner (tk_json_object.cuh), compiled for the host, against
CPython: ``torch.tensor([json.loads(v)[k] for k in fields], dtype)`` on the same bytes.  Runs
without a GPU."""
import ctypes
import json
import os
import random
import subprocess

import numpy as np
import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
DTYPES = [(torch.float32, np.float32, np.uint32), (torch.float64, np.float64, np.uint64),
          (torch.int32, np.int32, np.int32), (torch.int64, np.int64, np.int64)]


@pytest.fixture(scope="module")
def lib():
    subprocess.run(["make", "-C", os.path.join(HERE, "native")], check=True, capture_output=True)
    return ctypes.CDLL(os.path.join(HERE, "native", "_build", "libtk_host_decode.so"))


def field_table(fields):
    names = [f.encode() for f in fields]
    off = np.zeros(len(names) + 1, np.uint32)
    off[1:] = np.cumsum([len(b) for b in names])
    return b"".join(names) or b"\0", off


def decode(lib, payload: bytes, fields, kind: int):
    blob, off = field_table(fields)
    row = np.zeros(max(len(fields), 1), DTYPES[kind][1])
    st = lib.tk_host_decode_object(payload, len(payload), kind, blob, off.ctypes.data_as(ctypes.c_void_p),
                                   len(fields), row.ctypes.data_as(ctypes.c_void_p))
    return st, row[:len(fields)]


def reference(payload: bytes, fields, kind: int):
    try:
        d = json.loads(payload)
        t = torch.tensor([d[k] for k in fields], dtype=DTYPES[kind][0])
        # torch.tensor([[], 1.0]) is a [2, 0] tensor: not a row of numbers, outside the schema
        return t if t.dim() == 1 else None
    except Exception:                                     # noqa: BLE001
        return None


def check(lib, payload, fields, kinds=(0, 1, 2, 3)):
    for kind in kinds:
        want = reference(payload, fields, kind)
        st, row = decode(lib, payload, fields, kind)
        if want is None:
            assert st != 0, (payload, kind)
        else:
            assert st == 0, (payload, kind, st)
            bits = DTYPES[kind][2]
            assert np.array_equal(row.view(bits), want.numpy().view(bits)), (payload, kind, row, want)


KNOWN = [
    (b'{"a": 1.5, "b": -2, "c": 3e4}', 0),
    (b'{"c":3,"b":2,"a":1}', 0),
    (b'\xef\xbb\xbf \r\n{ "a" : 1 , "b" : 2 , "c" : 3 } \t', 0),
    (b'{"a": 1, "a": 2, "b": 0, "c": 0}', 0),                       # duplicate key: the last one wins
    (b'{"a": "x", "a": 2, "b": 0, "c": 0}', 0),                     # ... also over a bad earlier value
    (b'{"a": 1e999, "a": 2, "b": 1e999, "c": 0}', 0),
    (b'{"a": 2, "a": "x", "b": 0, "c": 0}', 4),
    (b'{"a": true, "b": false, "c": 1}', 0),                        # Python bools are ints
    (b'{"a": null, "b": 0, "c": 1}', 4),
    (b'{"a": [1], "b": 0, "c": 1}', 4),
    (b'{"a": {"a": 1}, "b": 0, "c": 1}', 4),
    (b'{"a": 1, "b": 2}', 8),                                       # KeyError
    (b'{}', 8),
    (b'{"a": 1, "b": 2, "c": 3,}', 1),
    (b'{"a": 1, "b": 2, "c": 3} x', 1),
    (b'{"a": 1, "b": 2 "c": 3}', 1),
    (b'{"a": 1, "b": 2, c: 3}', 1),
    (b"{'a': 1, 'b': 2, 'c': 3}", 1),
    (b'[1, 2, 3]', 4),
    (b'', 1),
    (b'{"a": 1, "b": 2, "c": 3, "note": "x\\ty \\u00e9 \\ud83d\\ude00 \\"q\\"", "n": null, "l": [1, [2, {"k": []}], {}]}', 0),
    (b'{"a": 1, "b": 2, "c": 3, "s": "\\x"}', 1),
    (b'{"a": 1, "b": 2, "c": 3, "s": "\\u12g4"}', 1),
    (b'{"a": 1, "b": 2, "c": 3, "s": "tab\there"}', 1),             # raw control character
    (b'{"a": 1, "b": 2, "c": 3, "s": "unterminated}', 1),
    ('{"a": 1, "b": 2, "c": 3, "s": "héllo 世界 \U0001F600"}'.encode(), 0),
    (b'{"a": 1, "b": 2, "c": 3, "s": "\xc0\xaf"}', 1),              # overlong UTF-8
    (b'{"a": 1, "b": 2, "c": 3, "s": "\xed\xa0\x80"}', 0),          # encoded surrogate: surrogatepass
    (b'{"a": 1, "b": 2, "c": 3, "s": "\xf4\x90\x80\x80"}', 1),      # above U+10FFFF
    (b'{"a": 1, "b": 2, "c": 3, "s": "\xe2\x82"}', 1),              # truncated sequence
    (b'{"a": NaN, "b": -Infinity, "c": Infinity}', None),
    (b'{"a": -0, "b": -0.0, "c": 1152921573326323713}', 0),
    (b'{"a": 1.0000000596046448, "b": 16777217, "c": 8.41e21}', None),
    (b'{"a": 2147483647, "b": -2147483648, "c": 1.5}', 0),
    (b'{"a": 2147483648, "b": 0, "c": 0}', None),
    (b'{"a": 01, "b": 0, "c": 0}', 1),
    (b'{"a": 1., "b": 0, "c": 0}', 1),
    (b'{"a": 1, "b": 2, "c": 3, "big": ' + b"9" * 4300 + b'}', 0),
    (b'{"a": 1, "b": 2, "c": 3, "big": ' + b"9" * 4301 + b'}', 1),  # int_max_str_digits
    (b'{"a": 1, "b": 2, "c": 3, "big": -' + b"9" * 4300 + b'}', 0),
    (b'{"a": 1, "b": 2, "c": 3, "big": ' + b"9" * 4301 + b'.0}', 0),
    (b'{"a": 1, "b": 2, "c": 3, "deep": ' + b"[" * 32 + b"]" * 32 + b'}', 0),
]


def test_known_records_match_cpython(lib):
    fields = ("a", "b", "c")
    for payload, f32_status in KNOWN:
        check(lib, payload, fields)
        if f32_status is not None:
            assert decode(lib, payload, fields, 0)[0] == f32_status, payload


def test_documented_limits_are_reported_not_guessed(lib):
    fields = ("a", "b", "c")
    # member name written with an escape: the reference compares the decoded name
    assert decode(lib, b'{"\\u0061": 1, "b": 2, "c": 3}', fields, 0)[0] == 4
    # nesting deeper than the scanner follows
    assert decode(lib, b'{"a": 1, "b": 2, "c": 3, "deep": ' + b"[" * 33 + b"]" * 33 + b'}', fields, 0)[0] == 4
    # UTF-16 payload (json.loads detects and decodes it)
    assert decode(lib, '{"a": 1, "b": 2, "c": 3}'.encode("utf-16"), fields, 0)[0] != 0


def gen_number(rng):
    k = rng.randrange(8)
    if k == 0:
        return repr(rng.uniform(-1e3, 1e3))
    if k == 1:
        return str(rng.randint(-10 ** 6, 10 ** 6))
    if k == 2:
        return "%d.%de%d" % (rng.randint(0, 9), rng.getrandbits(40), rng.randint(-30, 30))
    if k == 3:
        return str(rng.getrandbits(rng.randint(1, 70)) * rng.choice([1, -1]))
    if k == 4:
        return rng.choice(["true", "false", "NaN", "Infinity", "-Infinity", "-0", "-0.0", "0", "1E5", "2147483648"])
    if k == 5:
        return repr(rng.random() * 10 ** rng.randint(-40, 40))
    if k == 6:
        return rng.choice(["null", '"str"', "[1, 2]", '{"x": 1}', "[]", "{}"])
    return repr(float(rng.randint(-100, 100)))


def gen_other(rng, depth=0):
    k = rng.randrange(7 if depth < 3 else 4)
    if k == 0:
        return gen_number(rng)
    if k == 1:
        return json.dumps("".join(rng.choice(["a", " ", '"', "\\", "\n", "é", "世", "\U0001F600", ",", ":", "}", "]"])
                                  for _ in range(rng.randint(0, 8))), ensure_ascii=rng.random() < 0.5)
    if k == 2:
        return "null"
    if k == 3:
        return rng.choice(["true", "false"])
    if k == 4:
        return "[" + rng.choice([",", " , "]).join(gen_other(rng, depth + 1) for _ in range(rng.randint(0, 3))) + "]"
    return "{" + ", ".join('"k%d": %s' % (i, gen_other(rng, depth + 1)) for i in range(rng.randint(0, 3))) + "}"


def gen_record(rng, fields):
    members = []
    names = list(fields)
    if rng.random() < 0.3:
        rng.shuffle(names)
    for name in names:
        if rng.random() < 0.04:
            continue                                      # missing key
        members.append((name, gen_number(rng)))
        if rng.random() < 0.08:
            members.append((name, gen_number(rng)))       # duplicate
    for _ in range(rng.randint(0, 3)):
        key = rng.choice(["extra", "a_", "", "bb", "é", "x y", "c世"]) + str(rng.randint(0, 3))
        members.insert(rng.randint(0, len(members)), (key, gen_other(rng)))
    ws = lambda: rng.choice(["", "", " ", "\n", "\t ", "\r\n"])           # noqa: E731
    body = ",".join("%s%s%s:%s%s%s" % (ws(), json.dumps(k, ensure_ascii=False), ws(), ws(), v, ws()) for k, v in members)
    return (rng.choice(["", "", " ", "﻿"]) + "{" + (body or ws()) + "}" + ws()).encode()


def test_random_records_match_cpython(lib):
    rng = random.Random(41)
    fields = ("a", "b", "c", "weight", "été")
    for it in range(6000):
        payload = gen_record(rng, fields)
        if rng.random() < 0.2:      # damage it
            cut = rng.randrange(len(payload) + 1)
            payload = payload[:cut] + rng.choice([b"", b",", b"}", b"{", b'"', b"\\", b":", b"x", b"\xff", b"\x00", b"[", b"]"]) \
                + payload[cut + rng.randint(0, 1):]
        # escaped member names are outside the scanner's scope (documented; tested above)
        try:
            d = json.loads(payload)
            if isinstance(d, dict) and '\\' in payload.decode("utf-8", "surrogatepass") and \
                    any(json.dumps(k, ensure_ascii=False) != '"%s"' % k for k in d):
                continue
        except Exception:                                 # noqa: BLE001
            pass
        check(lib, payload, fields, kinds=(it % 4,))


def test_many_fields_and_schema_order(lib):
    rng = random.Random(5)
    fields = tuple("f%02d" % i for i in range(64))
    for _ in range(200):
        order = list(fields)
        if rng.random() < 0.5:
            rng.shuffle(order)
        payload = json.dumps({k: rng.uniform(-10, 10) for k in order}).encode()
        check(lib, payload, fields, kinds=(0, 1))


# ---------------------------------------------------------------- nested arrays of a fixed shape
def decode_nested(lib, payload: bytes, shape, kind: int, flat: bool = False):
    """flat: the comma-split decomposition the flat decode kernels run (tk_json.cuh) instead of the
    sequential scanner of the object kernel (tk_json_object.cuh)."""
    dims = np.asarray(shape, np.uint32)
    row = np.zeros(int(np.prod(shape)), DTYPES[kind][1])
    fn = lib.tk_host_decode_nested_flat if flat else lib.tk_host_decode_nested
    st = fn(payload, len(payload), kind, dims.ctypes.data_as(ctypes.c_void_p), len(shape),
            row.ctypes.data_as(ctypes.c_void_p))
    return st, row.reshape(shape)


def check_nested(lib, payload, shape, kinds=(0, 1, 2, 3)):
    for flat in (False, True):
        _check_nested(lib, payload, shape, kinds, flat)


def _check_nested(lib, payload, shape, kinds, flat):
    for kind in kinds:
        try:
            want = torch.tensor(json.loads(payload), dtype=DTYPES[kind][0])
            if tuple(want.shape) != tuple(shape):
                want = None                               # a tensor, but not a sample of this schema
        except Exception:                                 # noqa: BLE001
            want = None
        st, got = decode_nested(lib, payload, shape, kind, flat)
        if want is None:
            assert st != 0, (payload, shape, kind, flat)
        else:
            assert st == 0, (payload, shape, kind, st, flat)
            bits = DTYPES[kind][2]
            assert np.array_equal(got.view(bits), want.numpy().view(bits)), (payload, kind)


def test_nested_arrays_known_records(lib):
    for payload, shape, status in [
        (b"[[1, 2, 3], [4, 5, 6]]", (2, 3), 0),
        (b"\xef\xbb\xbf [ [1.5,-2e3 ,true] ,\r\n[NaN, -0, -0.0]\t] ", (2, 3), None),
        (b"[[[1], [2]], [[3], [4]], [[5], [6]]]", (3, 2, 1), 0),
        (b"[1, 2, 3, 4]", (4,), 0),
        (b"[[[[7]]]]", (1, 1, 1, 1), 0),
        (b"[[1, 2, 3], [4, 5]]", (2, 3), 6),              # ragged: torch.tensor raises
        (b"[[1, 2, 3], [4, 5, 6, 7]]", (2, 3), 6),
        (b"[[1, 2, 3]]", (2, 3), 6),                      # another shape: torch.stack would raise
        (b"[[1, 2, 3], [4, 5, 6], [7, 8, 9]]", (2, 3), 6),
        (b"[1, 2, 3]", (1, 3), 6),                        # not deep enough
        (b"[[[1, 2, 3]]]", (1, 3), 6),                    # too deep
        (b"[[1, 2, 3], [4, 5, 6]] x", (2, 3), 1),
        (b"[[1, 2, 3], [4, 5, 6]", (2, 3), 1),
        (b"[[1, 2, 3] [4, 5, 6]]", (2, 3), 1),
        (b"[[1, 2, 3], [4, 5, 6],]", (2, 3), 6),
        (b"[[1, 2, null], [4, 5, 6]]", (2, 3), 4),
        (b'[[1, 2, "3"], [4, 5, 6]]', (2, 3), 6),
        (b"[[1, 2, 1e999], [4, 5, 6]]", (2, 3), None),
        (b"", (2, 3), 1),
        (b"{}", (2, 3), 6),
    ]:
        check_nested(lib, payload, shape)
        if status is not None:
            assert decode_nested(lib, payload, shape, 0)[0] == status, payload


def test_nested_arrays_random_records(lib):
    rng = random.Random(43)
    for it in range(3000):
        shape = tuple(rng.randint(1, 4) for _ in range(rng.randint(1, 4)))
        true_shape = shape if rng.random() < 0.8 else tuple(rng.randint(1, 4) for _ in range(rng.randint(1, 4)))

        def build(dims):
            if not dims:
                tok = gen_number(rng)
                return tok if not tok.startswith(("[", "{")) or rng.random() < 0.02 else "7"
            sep = rng.choice([",", ", ", " ,\n"])
            return "[" + rng.choice(["", " "]) + sep.join(build(dims[1:]) for _ in range(dims[0])) + "]"
        payload = (rng.choice(["", " ", "﻿"]) + build(true_shape) + rng.choice(["", "\n"])).encode()
        if rng.random() < 0.15:
            cut = rng.randrange(len(payload) + 1)
            payload = payload[:cut] + rng.choice([b"", b",", b"]", b"[", b"x", b"1"]) + payload[cut + rng.randint(0, 1):]
        check_nested(lib, payload, shape, kinds=(it % 4,))


# ---------------------------------------------------------------- the fast lane (tk_object_fast.cuh)
def decode_ex(lib, payload: bytes, fields, kind: int):
    blob, off = field_table(fields)
    row = np.zeros(max(len(fields), 1), DTYPES[kind][1])
    fast = ctypes.c_int(-1)
    st = lib.tk_host_decode_object_ex(payload, len(payload), kind, blob, off.ctypes.data_as(ctypes.c_void_p),
                                      len(fields), row.ctypes.data_as(ctypes.c_void_p), ctypes.byref(fast))
    row2 = np.zeros(max(len(fields), 1), DTYPES[kind][1])
    st2 = lib.tk_host_decode_object_scanner(payload, len(payload), kind, blob, off.ctypes.data_as(ctypes.c_void_p),
                                            len(fields), row2.ctypes.data_as(ctypes.c_void_p))
    return st, row[:len(fields)], fast.value, st2, row2[:len(fields)]


def dumps_record(rng, fields, separators):
    d = {}
    names = list(fields)
    if rng.random() < 0.2:
        rng.shuffle(names)
    extras = ["id", "note", "ts", "k_%d" % rng.randrange(5)]
    for name in names:
        if rng.random() < 0.15:
            e = rng.choice(extras)
            d[e] = rng.choice([rng.randrange(10 ** 6), rng.uniform(-5, 5), "abc", "a b,c}d", ""])
        k = rng.randrange(6)
        d[name] = (rng.uniform(-1e3, 1e3) if k < 3 else rng.randrange(-10 ** 7, 10 ** 7) if k == 3
                   else float(rng.randrange(-50, 50)) if k == 4 else 0.1 + rng.random() * 10 ** rng.randrange(0, 9))
    return json.dumps(d, separators=separators).encode()


@pytest.mark.parametrize("separators", [None, (",", ":")])
def test_fast_lane_takes_what_json_dumps_writes_and_agrees_with_the_scanner(lib, separators):
    rng = random.Random(9)
    fields = ("x", "y", "label", "weight", "a1", "velocity_mps", "q", "z")
    taken = total = 0
    for it in range(4000):
        payload = dumps_record(rng, fields, separators)
        kind = it % 2                                     # float32 / float64 (ints: next test)
        st, row, fast, st2, row2 = decode_ex(lib, payload, fields, kind)
        assert st == st2 == 0, payload
        bits = DTYPES[kind][2]
        assert np.array_equal(row.view(bits), row2.view(bits)), (payload, row, row2)
        want = reference(payload, fields, kind)
        assert want is not None and np.array_equal(row.view(bits), want.numpy().view(bits)), payload
        taken += fast
        total += 1
    assert taken >= 0.97 * total, (taken, total)


def test_fast_lane_never_disagrees_on_variants_and_damage(lib):
    """Records around the edge of what the fast lane takes: the combined decoder (fast lane, else
    scanner) must answer as the scanner alone and as CPython do."""
    rng = random.Random(10)
    fields = ("x", "y", "n")
    base = [b'{"x": 1.5, "y": -2.25, "n": 7}', b'{"x":1.5,"y":-2.25,"n":7}', b'{"n": 7, "y": 2, "x": 0.5, "x": 9.75}',
            b'{"x": 1.5, "s": "plain string", "y": 2, "n": 3}', b'{"x": 1.5, "s": "esc \\" aped", "y": 2, "n": 3}',
            b'{"x": 1.5, "s": "h\xc3\xa9", "y": 2, "n": 3}', b'{"x": 1.5, "y": 2, "n": 3} ', b' {"x": 1.5, "y": 2, "n": 3}',
            b'{"x": 1.5,  "y": 2, "n": 3}', b'{"x" : 1.5, "y": 2, "n": 3}', b'{"x": 1.5, "y": 2, "n": true}',
            b'{"x": 1.5, "y": 2, "n": null}', b'{"x": 1.5, "y": [2], "n": 3}', b'{"x": 1.5, "y": 2}', b'{"x": 1e5, "y": 2, "n": 3}',
            b'{"x": 01.5, "y": 2, "n": 3}', b'{"x": -0, "y": -0.0, "n": 0}', b'{"x": 1.5, "y": 2, "n": 3, "long_member_name_of_30_bytes__": 1}',
            b'{"x": 1.5, "y": 2, "n": 3, "s": "' + b"v" * 40 + b'"}', b'{"x": 12345678901234567890, "y": 2, "n": 3}',
            b'{"x": 0.1234567890123456789012, "y": 2, "n": 3}', b'{"x": 1.5, "y": 2, "n": 3}x', b'{"x": 1.5, "y": 2, "n": 3',
            b'{"x": 1.5, "y": 2, "n": 3}}', b'{"x": 1.5, "y": 2,, "n": 3}', b'{"x": 1.5 "y": 2, "n": 3}', b'{"x": "1.5", "y": 2, "n": 3}',
            b'{"": 1, "x": 1.5, "y": 2, "n": 3}', b'{"x": 99999999, "y": 100000000, "n": 3}']
    for payload in base:
        for kind in range(4):
            st, row, fast, st2, row2 = decode_ex(lib, payload, fields, kind)
            assert st == st2, (payload, kind, st, st2, fast)
            if st == 0:
                bits = DTYPES[kind][2]
                assert np.array_equal(row.view(bits), row2.view(bits)), (payload, kind)
        check(lib, payload, fields)
    for it in range(6000):
        payload = bytearray(dumps_record(rng, fields, rng.choice([None, (",", ":")])))
        for _ in range(rng.randint(1, 2)):
            cut = rng.randrange(len(payload) + 1)
            payload[cut:cut + rng.randint(0, 1)] = rng.choice([b"", b",", b"}", b"{", b'"', b"\\", b":", b"x", b"\xff", b"\x00", b" ",
                                                               b".", b"-", b"e", b"1", b"\x7f", b"/"])
        payload = bytes(payload)
        kind = it % 4
        st, row, fast, st2, row2 = decode_ex(lib, payload, fields, kind)
        assert st == st2, (payload, kind, st, st2, fast)
        if st == 0:
            bits = DTYPES[kind][2]
            assert np.array_equal(row.view(bits), row2.view(bits)), (payload, kind)
