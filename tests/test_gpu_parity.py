"""GPU parity: the CUDA path (through the public API, C-ABI underneath) against the
oracle on identical seeded records.  Bit-exact for every dtype, including float
(0 ulp: the decimal->binary conversion is correctly rounded like CPython's)."""
import json
import random

import numpy as np
import pytest
import torch

from helpers import assert_same, bit_equal, make_dataset, run_gpu, run_oracle
from torchkafka_b200 import KafkaDataset, auto_commit
from torchkafka_b200.testing import stub_kafka as sk
from torchkafka_b200.schema import ImageU8, JsonArray, PayloadError, RawBytes, TokenList
from torchkafka_b200.testing import synthetic as syn
from torchkafka_b200.testing.stub_kafka import RecordLog

pytestmark = pytest.mark.gpu


def check(log, schema, batch_size, drop_last=False, **kw):
    gpu = run_gpu(log, schema, batch_size, drop_last, **kw)
    cpu = run_oracle(log, schema, batch_size, drop_last)
    assert_same(gpu, cpu, schema)
    return gpu


def test_cfg1_fixed_len_floats():
    log = syn.make_log(1, 5000, seed=1)
    check(log, JsonArray(torch.float32, fixed_len=8), 4)


def test_cfg2_variable_floats_filter():
    log = syn.make_log(2, 6000, seed=2)
    check(log, JsonArray(torch.float32, max_len=256, min_size=32), 1024)


def test_cfg2_small_slots_many_carries():
    # slots far smaller than a batch: the open batch is carried across many slots
    log = syn.make_log(2, 1500, seed=3)
    check(log, JsonArray(torch.float32, max_len=256, min_size=32), 256, slot_bytes=96 << 10)


def test_cfg2_same_thread_producer_and_poll_objects():
    log = syn.make_log(2, 1500, seed=13, layout="chunked", chunk=100)
    sc = JsonArray(torch.float32, max_len=256, min_size=32)
    check(log, sc, 256, slot_bytes=512 << 10, device_producer_thread=False)
    check(log, sc, 256, slot_bytes=512 << 10, use_packed=False)
    check(log, sc, 256, slot_bytes=512 << 10, use_packed=False, device_producer_thread=False)


def test_cfg2_float64():
    log = syn.make_log(2, 1200, seed=4)
    check(log, JsonArray(torch.float64, max_len=256, min_size=32), 128)


def test_cfg3_tokens():
    log = syn.make_log(3, 1500, seed=5)
    check(log, TokenList(torch.int32, max_len=2048), 512)


def test_cfg3_tokens_int64_chunked_poll_path():
    log = syn.make_log(3, 600, seed=6, layout="chunked", chunk=37)
    check(log, TokenList(torch.int64, max_len=2048), 64, use_packed=False)


def test_cfg4_images():
    log = syn.make_log(4, 70, seed=7)
    check(log, ImageU8(224, 224, 3), 32, slot_bytes=24 * 150528)


def test_cfg5_raw64_commit_offsets():
    log = syn.make_log(5, 20000, seed=8)
    check(log, RawBytes((64,), torch.uint8), 256)


def test_drop_last_and_tail():
    log = syn.make_log(2, 700, seed=9)
    sc = JsonArray(torch.float32, max_len=256, min_size=32)
    check(log, sc, 64, drop_last=True)
    check(log, sc, 64, drop_last=False)


def test_survey_commit_example():
    # 3 partitions round-robin, lengths below, min_size=4, batch_size=3
    lens = [4, 4, 2, 2, 2, 4, 2, 4, 4, 4, 2, 4, 4, 2]
    vals = [json.dumps([float(i)] * n).encode() for i, n in enumerate(lens)]
    i = np.arange(len(vals))
    log = RecordLog(vals, i % 3, i // 3)
    gpu = check(log, JsonArray(torch.float32, max_len=8, min_size=4), 3)
    commits = [c for _, c in gpu]
    assert commits == [{0: 2, 1: 2, 2: 2}, {0: 4, 1: 3, 2: 3}, {0: 5, 1: 5, 2: 4}]


ADVERSARIAL = [
    "0.1", "1e23", "8.41e21", "1.0000000596046448", "1152921573326323713", "16777217",
    "2.2250738585072011e-308", "4.9e-324", "2.4703282292062327e-324", "2.4703282292062328e-324",
    "1e400", "1E-400", "-0", "-0.0", "0", "1E+5", "1e-5", "NaN", "Infinity", "-Infinity",
    "123456789012345678901234567890", "9007199254740993", "9007199254740992.5",
    "1.7976931348623157e308", "1.7976931348623158e308", "0.000001", "5e-324", "3e-324",
    "0.30000000000000004", "179769313486231570000000000000000000000000000000000000000000000000000"
    "0000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000"
    "0000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000"
    "00000000000000000000000000000000000000000000000000000000",
    # exact halfway between 1.0 and nextafter(1.0): ties-to-even, and one digit above it
    "1.00000000000000011102230246251565404236316680908203125",
    "1.000000000000000111022302462515654042363166809082031251",
    "1.00000000000000011102230246251565404236316680908203124999",
]


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_adversarial_floats(dtype):
    vals = []
    for k in range(0, len(ADVERSARIAL), 5):
        vals.append(("[" + ", ".join(ADVERSARIAL[k:k + 5]) + "]").encode())
    vals.append(b"\xef\xbb\xbf[1]")
    vals.append(b"  [ 1 ,\t2 ,\n3 ]\r\n ")
    vals.append(b"[]")
    vals.append(b"[true, false, 2.5]")
    i = np.arange(len(vals))
    log = RecordLog(vals, i % 2, i // 2)
    check(log, JsonArray(dtype, max_len=8), 4)


def test_float32_double_rounding_boundaries():
    """float32 results are round_f32(round_f64(text)) like torch.tensor(json.loads(..)).  The
    parse kernel takes a one-FMA shortcut away from float32 rounding boundaries and the exact
    conversion next to them: decimal strings within a few double ulps of float32 halfway points
    (both sides, exact ties, 16/17 digits, with and without integer part) must come out bit-exact."""
    import random
    from decimal import Decimal, getcontext
    getcontext().prec = 60
    rng = random.Random(5)
    toks = []
    for _ in range(6000):
        e = rng.randint(-20, 26)
        m = rng.getrandbits(23) | (1 << 23)
        half = (Decimal(2 * m + 1) * Decimal(2) ** (e - 24))       # halfway between two floats
        ulp64 = Decimal(2) ** (e - 52)
        x = half + ulp64 * Decimal(rng.choice([0, 0, 1, -1, 2, -2, 3, -3, 5, -5, 9, -9, 40, -40])) \
            * Decimal(rng.choice([1, 1, 0.5, 0.25, 0.75]))
        digits = rng.choice([15, 16, 17, 17, 18, 19])
        s = format(x, "f")
        ip, _, fp = s.partition(".")
        keep = max(1, digits - (len(ip.lstrip("0")) or 0))
        fp = fp[:min(keep, 16)] or "0"
        if len(ip.lstrip("0")) > 8:
            continue
        toks.append(rng.choice(["", "-"]) + ip + "." + fp)
    vals = [("[" + ", ".join(toks[k:k + 48]) + "]").encode() for k in range(0, len(toks), 48)]
    i = np.arange(len(vals))
    log = RecordLog(vals, i % 4, i // 4)
    check(log, JsonArray(torch.float32, max_len=48), 16)
    check(log, JsonArray(torch.float64, max_len=48), 16)


def test_random_decimal_strings_float64():
    import random
    rng = random.Random(11)
    vals = []
    for _ in range(400):
        toks = []
        for _ in range(rng.randint(1, 40)):
            k = rng.randrange(5)
            if k == 0:
                toks.append(repr(rng.uniform(-1e3, 1e3)))
            elif k == 1:
                toks.append("%de%d" % (rng.getrandbits(rng.randint(1, 70)), rng.randint(-345, 308)))
            elif k == 2:
                toks.append("0." + "0" * rng.randint(0, 30) + str(rng.getrandbits(rng.randint(1, 90))))
            elif k == 3:
                toks.append(str(rng.getrandbits(rng.randint(1, 200))))
            else:
                toks.append("-%d.%d" % (rng.getrandbits(20), rng.getrandbits(64)))
        vals.append(("[" + rng.choice([", ", ",", " , "]).join(toks) + "]").encode())
    i = np.arange(len(vals))
    log = RecordLog(vals, i % 4, i // 4)
    check(log, JsonArray(torch.float64, max_len=64), 32)


BAD = [b"[1,]", b"[1.]", b"[01]", b"[+1]", b"[1 2]", b"[-]", b"[1]x", b"", b"[1, 2", b"[1e]",
       b"[null]", b'["a"]', b"[[1]]", b"3.5", b"[-NaN]", b"{}", b"[1,,2]", b"[,1]"]


@pytest.mark.parametrize("bad", BAD)
def test_bad_payload_raises_like_reference(bad):
    good = [json.dumps([1.0, 2.0]).encode()] * 5
    vals = good[:3] + [bad] + good[3:]
    i = np.arange(len(vals))
    log = RecordLog(vals, i % 2, i // 2)
    sc = JsonArray(torch.float32, max_len=4)
    # the reference raises from _process on the bad record ...
    if bad not in (b"[[1]]", b"3.5"):       # these two the reference happens to accept (other shapes)
        with pytest.raises(Exception):
            run_oracle(log, sc, 2)
    # ... the GPU path delivers the batch before it, then raises PayloadError
    with pytest.raises(PayloadError) as ei:
        run_gpu(log, sc, 2)
    assert ei.value.offset == 1 and ei.value.partition == 1


def test_int_overflow_and_truncation():
    sc = JsonArray(torch.int32, max_len=4)
    ok = RecordLog([b"[1.5, -2.9, 2147483647, -2147483648]", b"[1e5, 0, -0, 7]"], [0, 0], [0, 1])
    check(ok, sc, 2)
    for bad in (b"[2147483648]", b"[-2147483649]", b"[1e30]", b"[NaN]", b"[9223372036854775808]",
                b"[2147483647.5]"):
        log = RecordLog([bad], [0], [0])
        with pytest.raises(Exception):
            run_oracle(log, sc, 1)
        with pytest.raises(PayloadError):
            run_gpu(log, sc, 1)


def test_too_long_and_ragged():
    with pytest.raises(PayloadError):
        run_gpu(RecordLog([b"[1,2,3,4,5]"], [0], [0]), JsonArray(torch.float32, max_len=4), 1)
    with pytest.raises(PayloadError):
        run_gpu(RecordLog([b"[1,2,3]", b"[1,2]"], [0, 0], [0, 1]), JsonArray(torch.float32, fixed_len=3), 2)


def test_raw_wrong_size():
    with pytest.raises(PayloadError):
        run_gpu(RecordLog([b"x" * 64, b"y" * 63], [0, 0], [0, 1]), RawBytes((64,)), 2)


def test_live_stream_with_pauses_and_batch_iteration():
    """A consumer whose poll() comes back empty now and then (a live topic) must not end the
    iteration before its idle timeout, and partly filled slots must still be delivered."""
    import torch
    from torch.utils.data import DataLoader
    from helpers import run_oracle, to_cpu
    from torchkafka_b200 import KafkaDataset
    from torchkafka_b200.testing import stub_kafka as sk

    log = syn.make_log(2, 900, seed=21, layout="chunked", chunk=50)
    sc = JsonArray(torch.float32, max_len=256, min_size=32)

    class Hiccup(sk.PollOnlyConsumer):
        calls = 0

        def poll(self, timeout_ms=0, max_records=None, update_offsets=True):
            Hiccup.calls += 1
            if Hiccup.calls % 3 == 0:
                return {}                               # nothing arrived within the timeout
            return super().poll(timeout_ms, min(max_records or 64, 64), update_offsets)

    class DS(KafkaDataset):
        payload_schema = sc
        device_batch_size = 128

        @classmethod
        def new_consumer(cls, *a, **k):
            k.pop("enable_auto_commit", None)
            return Hiccup(*a, log=log, consumer_timeout_ms=300, **k)

    ds = DS("topic")
    got = [to_cpu(b) for b in DataLoader(ds, batch_size=None)]      # plain iteration, whole batches
    want = run_oracle(log, sc, 128)
    assert len(got) == len(want)
    for g, (w, _) in zip(got, want):
        assert torch.equal(g.values, w[0]) and torch.equal(g.lengths, w[1])
    ds.close()


def _fused_case(name):
    """(values, schema, batch size) of one scenario for the single-pass / three-kernel comparison."""
    import random
    rng = random.Random(hash(name) & 0xFFFF)
    if name == "cfg2_with_bad_records":
        vals = list(syn.make_log(2, 3000, seed=41)._byte_values)
        vals[17] = b"[1, 2,, 3]"                 # malformed: both must flag it
        vals[18] = b" [ 1 ,\t2.5e3 , -0 ,1e400 ] "
        vals[19] = b""
        vals[20] = b"[]"
        vals[21] = b"\xef\xbb\xbf[3.5]"
        return vals, JsonArray(torch.float32, max_len=256, min_size=32), 128
    if name == "tiny_records":                   # more records per tile than the per-tile tables hold
        vals = [rng.choice([b"[1]", b"[]", b"[1,2]", b"[7, 8, 9]", b"[x]", b"1"]) for _ in range(30000)]
        return vals, JsonArray(torch.int64, max_len=8, min_size=0), 64
    if name == "huge_records":                   # records longer than the shared-memory window
        vals = list(syn.make_log(2, 400, seed=42)._byte_values)
        big = json.dumps([rng.uniform(-1e3, 1e3) for _ in range(9000)]).encode()
        vals[50] = big
        vals[51] = b"[" + b" " * 70000 + b"1.5,2.5 ]"
        vals[399] = big[:-1] + b", 1e3]"
        return vals, JsonArray(torch.float64, max_len=16384, min_size=2), 32
    if name == "tokens":
        return list(syn.make_log(3, 700, seed=43)._byte_values), TokenList(torch.int32, max_len=2048), 32
    if name == "fixed_len_ragged":
        vals = list(syn.make_log(1, 5000, seed=44)._byte_values)
        vals[1234] = b"[1.0, 2.0]"
        return vals, JsonArray(torch.float32, fixed_len=8), 4
    raise KeyError(name)


@pytest.mark.parametrize("case", ["cfg2_with_bad_records", "tiny_records", "huge_records", "tokens", "fixed_len_ragged"])
def test_single_pass_kernel_agrees_with_the_three_kernel_chain(case):
    """tk_json_ingest (one pass over the payload, tiles in shared memory) and the chain
    tk_json_scan -> tk_compact -> tk_json_decode are independent decompositions of the same
    contract: identical rows, lengths, masks, per-record status and index, batch ends, longest
    records and commit matrix -- including tiles that take the single-pass kernel's scalar path."""
    from torchkafka_b200.engine import SlotEngine
    vals, sc, B = _fused_case(case)
    n = len(vals)
    log = RecordLog(vals, (np.arange(n) // 7 % 16).astype(np.int32), np.arange(n) // 16 + (np.arange(n) % 7))
    outs = []
    for fused in (False, True):
        eng = SlotEngine(sc, B, slot_bytes=16 << 20, slot_records=n + 8, n_slots=2, device=0, n_partitions=16)
        eng.set_json_fused(fused)
        m, nb = eng.fill_packed(0, log.values, log.starts.astype(np.uint32), log.partitions, log.offsets)
        eng.submit(0, m, nb)
        out = eng.process(0, m, None, None)
        res = eng.result(0)
        ln, st, kx = eng.record_debug(0, m)
        rows = B + res.n_kept
        outs.append(([t[B:rows].cpu() for t in out.tensors], res, ln.copy(), st.copy(), kx.copy()))
        eng.close()
    (ta, ra, la, sa, ka), (tb, rb, lb, sb, kb) = outs
    assert (ra.n_records, ra.n_kept, ra.n_closed, ra.carry_out) == (rb.n_records, rb.n_kept, rb.n_closed, rb.carry_out)
    assert ra.n_errors == rb.n_errors and ra.first_error_rec == rb.first_error_rec
    assert np.array_equal(la, lb) and np.array_equal(sa, sb) and np.array_equal(ka, kb)
    assert np.array_equal(ra.commit, rb.commit)
    assert np.array_equal(ra.batch_end, rb.batch_end) and np.array_equal(ra.batch_lmax, rb.batch_lmax)
    # rows of records without an error are identical bit for bit (a bad record's row is unspecified)
    good = torch.from_numpy(sa[ka >= 0] == 0)
    assert int(good.sum()) > 0
    for x, y in zip(ta, tb):
        assert x.shape == y.shape
        assert bit_equal(x[good], y[good])


@pytest.mark.parametrize("dist", ["unit", "tiny", "huge", "mixed"])
@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_numbers_of_every_magnitude_and_both_parse_kernel_variants(dist, dtype):
    """json.dumps writes "0.00123..." (21-22 bytes) below 0.01 and exponents below 1e-4 and from
    1e16: the parse kernel takes them in its wide variant inline / through the out-of-line long lane
    (tk_fl_scan_long).  Both variants of the kernel and the engine's own choice must give the
    oracle's bits."""
    from torchkafka_b200.engine import SlotEngine
    rng = random.Random({"unit": 1, "tiny": 2, "huge": 3, "mixed": 4}[dist])
    draw = {"unit": lambda: rng.uniform(-1, 1), "tiny": lambda: rng.gauss(0, 1e-3),
            "huge": lambda: rng.uniform(-1, 1) * 10 ** rng.randrange(14, 300),
            "mixed": lambda: rng.choice([rng.uniform(-1e3, 1e3), rng.gauss(0, 1e-6), rng.randrange(-10 ** 7, 10 ** 7),
                                         rng.random() * 1e-320, 0.0, -0.0, 1e22, 5e-324])}[dist]
    vals = [json.dumps([draw() for _ in range(rng.randint(16, 120))]).encode() for _ in range(1500)]
    sc, B, n = JsonArray(dtype, max_len=128, min_size=20), 64, len(vals)
    log = RecordLog(vals, (np.arange(n) // 7 % 16).astype(np.int32), np.arange(n) // 16 + (np.arange(n) % 7))
    outs = []
    for wide in (False, True, None):
        eng = SlotEngine(sc, B, slot_bytes=16 << 20, slot_records=n + 8, n_slots=2, device=0, n_partitions=16)
        if wide is not None:
            eng.json_wide_fixed = wide
            eng.set_json_wide(wide)
        m, nb = eng.fill_packed(0, log.values, log.starts.astype(np.uint32), log.partitions, log.offsets)
        eng.submit(0, m, nb)
        out = eng.process(0, m, None, None)
        res = eng.result(0)
        assert res.n_errors == 0
        if wide is None and dist != "mixed":
            assert eng.json_wide, "small / large magnitudes: the engine must pick the wide kernel"
        outs.append([t[B:B + res.n_kept].cpu() for t in out.tensors])
        eng.close()
    kept = [json.loads(v) for v in vals if v.count(b",") + 1 >= 20]
    want = torch.zeros((len(kept), 128), dtype=dtype)
    for i, row in enumerate(kept):
        want[i, :len(row)] = torch.tensor(row, dtype=dtype)
    for tensors in outs:
        assert bit_equal(tensors[0], want)
        assert tensors[1].tolist() == [len(r) for r in kept]


def test_images_table_fallback_when_fma_is_not_exact():
    """mean/std for which the one-FMA normalisation differs from the exact chain in one of the
    256 inputs: the library must notice (device-side exhaustive check) and use the table kernel;
    either way the result equals torch's `((x/255) - mean) / std -> bf16` bit for bit."""
    log = syn.make_log(4, 20, seed=51, h=40, w=36, c=3)
    check(log, ImageU8(40, 36, 3, mean=(0.341, 0.348, 0.804), std=(0.064, 0.085, 0.494)), 8,
          slot_bytes=7 * 40 * 36 * 3)
    check(log, ImageU8(40, 36, 3), 8, slot_bytes=7 * 40 * 36 * 3)          # ImageNet constants: FMA path
    # odd shapes: single channel, width that breaks the 16-byte store alignment
    log1 = syn.make_log(4, 9, seed=52, h=13, w=7, c=1)
    check(log1, ImageU8(13, 7, 1, mean=(0.5,), std=(0.25,)), 4)


# ---------------------------------------------------------------- fixed-layout structs (§8f-4)
def struct_log(schema, n, seed, partitions=4, bad_at=None):
    rng = np.random.default_rng(seed)
    vals = []
    for i in range(n):
        parts = []
        for _, dtype, shape in schema.fields:
            count = int(np.prod(shape)) if shape else 1
            nb = count * torch.empty((), dtype=dtype).element_size()
            parts.append(rng.integers(0, 256, nb, dtype=np.uint8).tobytes())
        v = b"".join(parts)
        if bad_at is not None and i == bad_at:
            v = v[:-1]
        vals.append(v)
    i = np.arange(n)
    return RecordLog(vals, i % partitions, i // partitions)


def test_struct_records_mixed_dtypes_and_odd_offsets():
    from torchkafka_b200.schema import StructBytes
    # offsets 0, 12, 13, 21, 29: word copies from unaligned sources, a 1-byte and a 3-byte field
    schema = StructBytes((("x", torch.float32, (3,)), ("flag", torch.uint8, ()), ("label", torch.int64, ()),
                          ("h", torch.float16, (2, 2)), ("rgb", torch.uint8, (3,)), ("w", torch.int32, (5,))))
    log = struct_log(schema, 1000, seed=41)
    gpu = check(log, schema, 64)
    assert type(gpu[0][0]).__name__ == "StructBatch" and gpu[0][0].h.shape == (64, 2, 2)
    check(log, schema, 7, slot_bytes=schema.record_bytes * 50)          # many slots, carried rows of every field


def test_struct_record_of_wrong_size_raises():
    from torchkafka_b200.schema import StructBytes
    schema = StructBytes((("a", torch.int32, (4,)), ("b", torch.float64, ())))
    log = struct_log(schema, 40, seed=42, bad_at=17)
    with pytest.raises(PayloadError) as ei:
        run_gpu(log, schema, 8)
    assert ei.value.status == 5 and (ei.value.partition, ei.value.offset) == (17 % 4, 17 // 4)


# ---------------------------------------------------------------- JSON objects with named fields (§8f-4)
OBJ_FIELDS = ("x", "y", "z", "w", "vx", "vy", "vz", "t")


def object_log(n, seed, partitions=4, **kw):
    vals = syn.object_values(n, seed, OBJ_FIELDS, **kw)
    i = np.arange(n)
    return RecordLog(vals, i % partitions, i // partitions)


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64, torch.int32, torch.int64])
def test_object_records_every_dtype(dtype):
    from torchkafka_b200.schema import JsonObject
    log = object_log(3000, seed=51)
    gpu = check(log, JsonObject(OBJ_FIELDS, dtype), 64)
    assert gpu[0][0].shape == (64, len(OBJ_FIELDS)) and gpu[0][0].dtype == dtype


def test_object_records_small_slots_subset_of_fields_and_reordered_columns():
    from torchkafka_b200.schema import JsonObject
    log = object_log(2000, seed=52)
    check(log, JsonObject(("t", "x", "vz"), torch.float32), 7, slot_bytes=6000)     # many slots, carried rows


def test_object_records_layouts_the_scanner_must_get_right():
    """White space, member order, duplicates, bools, ignored members of every kind, non-ASCII,
    records longer than a warp's window (scanned from global memory), empty strings."""
    from torchkafka_b200.schema import JsonObject
    rng = np.random.default_rng(53)
    vals = [
        b'{"a": 1.5, "b": -2, "c": 3e4}',
        b'\xef\xbb\xbf \r\n{ "c" : 3 ,\t"b" : 2 , "a" : 1 } \n',
        b'{"a": 1, "a": "x", "a": 2.25, "b": true, "c": false}',
        b'{"a": 1e999, "a": 7, "b": 0, "c": -0.0, "b": -0}',
        '{"s": "h\\u00e9llo \\"q\\" \\\\ 世界 \U0001F600", "a": 1.0000000596046448, "n": null, "b": 16777217, '
        '"l": [1, [2, {"k": []}], {}], "c": 1152921573326323713, "": {"a": "not this one"}}'.encode(),
        b'{"pad": "' + b"x" * 7000 + b'", "a": 1, "b": 2, "c": 3}',                 # longer than the window
        b'{"a": NaN, "b": -Infinity, "c": 8.41e21, "big": ' + b"9" * 4300 + b"}",
        b'{"a":0.1,"b":1e23,"c":4.9e-324}',
    ]
    vals += [json.dumps({"pad": "y" * int(rng.integers(0, 3000)), "c": float(rng.normal()), "a": int(rng.integers(-99, 99)),
                         "b": float(rng.uniform(-1, 1))}).encode() for _ in range(200)]
    i = np.arange(len(vals))
    log = RecordLog(vals, i % 3, i // 3)
    for dtype in (torch.float32, torch.float64):
        check(log, JsonObject(("a", "b", "c"), dtype), 5)


OBJ_BAD = [(b'{"a": 1, "b": 2}', 8), (b'{"a": 1, "b": 2, "c": null}', 4), (b'{"a": 1, "b": 2, "c": [3]}', 4),
           (b'{"a": 1, "b": 2, "c": "3"}', 4), (b'{"a": 1, "b": 2, "c": 3,}', 1), (b'{"a": 1, "b": 2, "c": 3} x', 1),
           (b'[1, 2, 3]', 4), (b'', 1), (b'{"a": 1, "b": 2, "c": 3, "s": "\\x"}', 1),
           (b'{"a": 1, "b": 2, "c": 3, "s": "\xc0\xaf"}', 1), (b'{"a": 1, "b": 2, "c": 01}', 1),
           (b'{"a": 1, "b": 2, "c": 3', 1), (b'{"a": 1, "b": 2, "c": ' + b"9" * 400 + b'}', 2)]


@pytest.mark.parametrize("bad,status", OBJ_BAD)
def test_object_bad_payload_raises_like_reference(bad, status):
    from torchkafka_b200.schema import JsonObject
    good = [b'{"a": 1, "b": 2, "c": 3}'] * 5
    vals = good[:3] + [bad] + good[3:]
    i = np.arange(len(vals))
    log = RecordLog(vals, i % 2, i // 2)
    sc = JsonObject(("a", "b", "c"), torch.float32)
    with pytest.raises(Exception):
        run_oracle(log, sc, 2)
    with pytest.raises(PayloadError) as ei:
        run_gpu(log, sc, 2)
    assert (ei.value.partition, ei.value.offset, ei.value.status) == (1, 1, status)


def test_object_int_targets_overflow_and_truncation():
    from torchkafka_b200.schema import JsonObject
    sc = JsonObject(("a", "b"), torch.int32)
    check(RecordLog([b'{"a": 1.5, "b": -2147483648}', b'{"b": true, "a": 2147483647}'], [0, 0], [0, 1]), sc, 2)
    for bad in (b'{"a": 2147483648, "b": 0}', b'{"a": NaN, "b": 0}', b'{"a": 1e30, "b": 0}'):
        log = RecordLog([bad], [0], [0])
        with pytest.raises(Exception):
            run_oracle(log, sc, 1)
        with pytest.raises(PayloadError) as ei:
            run_gpu(log, sc, 1)
        assert ei.value.status == 2


# ---------------------------------------------------------------- nested arrays of a fixed shape (§8f-4)
def nested_log(n, shape, seed, partitions=4, ints=False):
    rng = np.random.default_rng(seed)
    vals = []
    for _ in range(n):
        a = rng.integers(-10 ** 6, 10 ** 6, shape) if ints else rng.uniform(-1e3, 1e3, shape)
        vals.append(json.dumps(a.tolist()).encode())
    i = np.arange(n)
    return RecordLog(vals, i % partitions, i // partitions)


@pytest.mark.parametrize("shape,dtype", [((4, 3), torch.float32), ((2, 3, 4), torch.float64), ((16,), torch.int32),
                                         ((2, 2, 2, 2), torch.int64), ((64, 24), torch.float32)])
def test_nested_arrays_match_torch_tensor(shape, dtype):
    from torchkafka_b200.schema import JsonNested
    log = nested_log(1500 if np.prod(shape) < 100 else 200, shape, seed=61, ints=not dtype.is_floating_point)
    gpu = check(log, JsonNested(shape, dtype), 32)
    assert gpu[0][0].shape == (32,) + shape and gpu[0][0].dtype == dtype
    check(log, JsonNested(shape, dtype), 5, slot_bytes=40000)           # many slots, carried rows


@pytest.mark.parametrize("shape,dtype", [((5, 7), torch.float32), ((3, 2, 4), torch.int64), ((64, 24), torch.float64)])
def test_nested_flat_chain_agrees_with_the_sequential_scanner(shape, dtype):
    """Nested arrays run through the flat chain (one thread per leaf, brackets checked from the leaf
    index); the object kernel's sequential scanner (`tk_json_nested_decode`, the staged entry point)
    is an independent implementation: same rows, same records flagged."""
    from torchkafka_b200.engine import SlotEngine
    from torchkafka_b200.schema import JsonNested
    log = nested_log(300, shape, seed=62, ints=not dtype.is_floating_point)
    vals = list(log._byte_values)
    vals[7] = vals[7].replace(b"], [", b"] ,\n [", 2)                 # white space around the brackets
    vals[11] = vals[11][:-2] + b", 1]]"                               # one leaf too many
    vals[13] = vals[13].replace(b"], [", b", ", 1)                    # a bracket pair missing: ragged
    vals[17] = b"[" + vals[17] + b"]"                                 # one level too deep
    n = len(vals)
    log = RecordLog(vals, np.arange(n) % 4, np.arange(n) // 4)
    sc = JsonNested(shape, dtype)
    outs = []
    for staged in (True, False):
        eng = SlotEngine(sc, 16, slot_bytes=16 << 20, slot_records=n + 8, n_slots=2, device=0, n_partitions=4)
        m, nb = eng.fill_packed(0, log.values, log.starts.astype(np.uint32), log.partitions, log.offsets)
        eng.submit(0, m, nb)
        out = eng.process(0, m, None, None, timers=[] if staged else None)
        res = eng.result(0)
        _, st, kx = eng.record_debug(0, m)
        rows = out.tensors[0][16:].cpu()
        outs.append((rows, res, st.copy(), kx.copy()))
        eng.close()
    (ra, resa, sa, ka), (rb, resb, sb, kb) = outs
    bad = {11, 13, 17}
    assert set(np.flatnonzero(sa).tolist()) == bad == set(np.flatnonzero(sb).tolist())
    assert resa.n_errors == resb.n_errors == 3 and resa.first_error_rec == resb.first_error_rec == 11
    # the sequential scanner keeps every record (its errors surface in the decode), the flat chain
    # drops the ones its scan already rejects: compare the rows record by record
    for i in range(n):
        if i in bad:
            continue
        assert ka[i] >= 0 and kb[i] >= 0
        assert bit_equal(ra[ka[i]], rb[kb[i]]), i


@pytest.mark.parametrize("bad", [b"[[1, 2, 3], [4, 5]]", b"[[1, 2, 3]]", b"[1, 2, 3]", b"[[[1, 2, 3], [4, 5, 6]]]",
                                 b"[[1, 2, 3], [4, 5, 6]] x", b"[[1, 2, 3], [4, 5, null]]",
                                 b"[[1, 2, 3], [4, 5, " + b"9" * 400 + b"]]"])
def test_nested_arrays_of_another_shape_raise(bad):
    from torchkafka_b200.schema import JsonNested
    good = [b"[[1, 2, 3], [4, 5, 6]]"] * 5
    vals = good[:3] + [bad] + good[3:]
    i = np.arange(len(vals))
    log = RecordLog(vals, i % 2, i // 2)
    sc = JsonNested((2, 3), torch.float32)
    with pytest.raises(Exception):
        run_oracle(log, sc, 2)          # torch.tensor on the ragged one, or torch.stack on the odd shape
    with pytest.raises(PayloadError) as ei:
        run_gpu(log, sc, 2)
    assert (ei.value.partition, ei.value.offset) == (1, 1)


# ---------------------------------------------------------------- leaving the loop early, slow topics
@pytest.mark.parametrize("mode,packed", [("records", False), ("packed", True), ("raw", True), ("raw_dma", True)])
def test_break_then_iterate_again_continues_behind_the_last_delivered_batch(mode, packed):
    """The pipeline polls ahead of what it has delivered.  Leaving the loop must not lose that
    read-ahead: the consumer is sought back, and a second iteration delivers exactly the batches
    the oracle delivers after the one the first iteration stopped at."""
    from torch.utils.data import DataLoader
    from helpers import to_cpu
    log = syn.make_log(2, 5000, seed=77, partitions=8, layout="chunked", chunk=40)
    schema = JsonArray(torch.float32, max_len=256, min_size=32)
    ds = make_dataset(log, schema, slot_bytes=200_000, use_packed=packed, device_input=mode,
                      consumer_kwargs={"raw_batch_bytes": 8192})
    loader = DataLoader(ds, batch_size=64)
    got = []
    for k, batch in enumerate(auto_commit(loader)):
        got.append(to_cpu(batch))
        if k == 6:
            break                                   # batch 6 delivered, never committed
    n_commits = len(ds._consumer.committed_log)
    assert n_commits == 6
    for batch in auto_commit(loader):
        got.append(to_cpu(batch))
    want = run_oracle(log, schema, 64)
    assert len(got) == len(want)
    for k, (g, (w, _)) in enumerate(zip(got, want)):
        assert bit_equal(g[0], w[0]) and torch.equal(g[1], w[1]) and torch.equal(g[2], w[2]), f"batch {k}"
    # the group's committed offsets end where the oracle's do
    assert dict(ds._consumer.committed_log[-1]) == want[-1][1]
    # and the first commit of the second iteration covered the batch that was left uncommitted
    assert dict(ds._consumer.committed_log[n_commits]) == want[7][1]
    ds.close()


def test_trickling_topic_is_delivered_within_the_latency_bound():
    """A topic that always has a few more records never lets poll() come back empty and never fills
    a slot: with device_max_latency_ms the partly filled slot is shipped after the bound."""
    import time
    log = syn.make_log(1, 600, seed=78, partitions=4, layout="chunked", chunk=5)

    class Trickle(sk.PollOnlyConsumer):
        def poll(self, timeout_ms=0, max_records=None, update_offsets=True):
            time.sleep(0.002)
            return super().poll(timeout_ms, min(max_records or 3, 3), update_offsets)

    class DS(KafkaDataset):
        payload_schema = JsonArray(torch.float32, fixed_len=8)
        device_slot_bytes = 1 << 20
        device_max_latency_ms = 40

        @classmethod
        def new_consumer(cls, *a, **k):
            k.pop("enable_auto_commit", None)
            return Trickle(*a, log=log, **k)

    ds = DS("topic", group_id="g")
    t0 = time.perf_counter()
    first_at = None
    rows = []
    for batch in ds.device_batches(4):
        if first_at is None:
            first_at = time.perf_counter() - t0
        rows.append(batch.cpu())
    total = time.perf_counter() - t0
    stats = ds.last_pipeline_stats
    ds.close()
    want = run_oracle(log, JsonArray(torch.float32, fixed_len=8), 4)
    assert bit_equal(torch.cat(rows), torch.cat([w for w, _ in want]))
    # ~0.4 s of trickling in ~40 ms slices: several slots, and the first batch long before the end
    assert stats["slots"] >= 4, stats
    assert first_at < 0.5 * total, (first_at, total)


def test_short_records_under_a_wide_schema_stay_inside_the_output_budget():
    """TokenList(max_len=2048) over ten-token records: one row per record at full width would be
    gigabytes per slot; the slot is cut by the output-byte budget instead."""
    import random
    rng = random.Random(5)
    n = 60000
    vals = [json.dumps([rng.randrange(0, 128000) for _ in range(rng.randint(1, 10))]).encode() for _ in range(n)]
    log = RecordLog(vals, (np.arange(n) % 8).astype(np.int32), np.arange(n) // 8)
    schema = TokenList(torch.int32, max_len=2048)
    torch.cuda.synchronize()
    torch.cuda.empty_cache()
    torch.cuda.reset_peak_memory_stats()
    base = torch.cuda.memory_allocated()
    budget = 64 << 20
    ds = make_dataset(log, schema, slot_bytes=64 << 20, device_output_bytes=budget)
    rows, lens = [], []
    for batch in ds.device_batches(512):
        assert batch.input_ids.shape[1] <= 10          # the view is as wide as the batch's longest record
        rows.append(batch.input_ids.cpu())
        lens.append(batch.lengths.cpu())
    stats = ds.last_pipeline_stats
    ds.close()
    peak = torch.cuda.max_memory_allocated() - base
    # slots in flight (default 3 + producers) times the budget, never "records x max_len"
    assert peak <= 8 * budget, peak
    assert stats["slots"] >= n * 10244 // budget       # the budget, not the byte size, cut the slots
    got_lens = torch.cat(lens)
    assert got_lens.tolist() == [v.count(b",") + 1 for v in vals]
    k = 0
    for r, ln in zip(rows, lens):
        for row, m in zip(r.tolist(), ln.tolist()):
            assert row[:m] == json.loads(vals[k]) and not any(row[m:])
            k += 1
    assert k == n


@pytest.mark.parametrize("cfg,schema,batch", [
    (1, JsonArray(torch.float32, fixed_len=8), 4),
    (2, JsonArray(torch.float32, max_len=256, min_size=32), 256),
    (2, JsonArray(torch.float64, max_len=256, min_size=100), 64),
    (3, TokenList(torch.int32, max_len=2048), 32),
])
def test_single_pass_kernel_through_the_public_api(cfg, schema, batch):
    """The same oracle comparison as the config tests above, with the dataset switched to the
    single-pass JSON kernel (small slots: batches straddle slots, carries, many tiles)."""
    log = syn.make_log(cfg, 4000 if cfg != 3 else 700, seed=60 + cfg, partitions=8, layout="chunked", chunk=30)
    gpu = run_gpu(log, schema, batch, slot_bytes=400_000, device_json_single_pass=True)
    assert_same(gpu, run_oracle(log, schema, batch), schema)
