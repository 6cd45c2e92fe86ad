"""tk_compact_noalloc: the compaction entry point for callers that cannot allocate.

Some users of the library run where cudaMalloc is off limits -- all HBM already belongs
to their own pool, or their plugin threads may not allocate because an allocation
synchronises the device.  For them the C-ABI has one entry point whose scratch is a fixed
array inside the library.  This test drives it through the C-ABI exactly as such a caller
would (device pointers in, device pointers out, no context object) and checks that it
gives what the regular entry point (tk_compact, context-owned scratch) gives, bit for bit
(the outputs are integers), on two calls in a row: the second call sees the scratch the
first one left behind and carries the open batch of the first.

The file name sorts last on purpose: a run that stops at the first failure has then
already executed every other test.
"""
import ctypes as C
import json

import numpy as np
import pytest
import torch

from torchkafka_b200 import _native
from torchkafka_b200.engine import SlotEngine
from torchkafka_b200.schema import JsonArray

pytestmark = pytest.mark.gpu

BATCH = 37
MIN_SIZE = 5


def _records(rng, n):
    """JSON arrays whose lengths straddle MIN_SIZE so that roughly 60 % are kept."""
    vals = [json.dumps([float(i)] * int(rng.integers(1, 12))).encode() for i in range(n)]
    starts = np.zeros(n + 1, np.uint32)
    starts[1:] = np.cumsum([len(v) for v in vals])
    data = np.frombuffer(b"".join(vals), dtype=np.uint8)
    return data, starts, (np.arange(n) % 8).astype(np.int32), (np.arange(n) // 8).astype(np.int64)


def _reference_compaction(eng, slot, run):
    """The existing path: scan + tk_compact on a context slot; returns (kidx, header)."""
    lib, ctx = eng.lib, eng.ctx
    st = C.c_void_p(eng.stream.cuda_stream)
    n, nbytes = eng.fill_packed(slot, *run)
    eng.submit(slot, n, nbytes)
    _native.check(lib.tk_slot_wait(ctx, slot, st))
    _native.check(lib.tk_json_scan(ctx, slot, 16, MIN_SIZE, st))
    _native.check(lib.tk_compact(ctx, slot, BATCH, st))
    _native.check(lib.tk_commit_offsets(ctx, slot, st))
    _native.check(lib.tk_slot_fetch_result(ctx, slot, st))
    res = eng.result(slot)
    _, _, kidx = eng.record_debug(slot, n)
    return kidx, res


def test_noalloc_compaction_matches_context_compaction_twice():
    lib = _native.load()
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(7)
    eng = SlotEngine(JsonArray(torch.float32, max_len=16, min_size=MIN_SIZE), BATCH,
                     slot_bytes=4 << 20, slot_records=20000, n_slots=2, device=0, n_partitions=8)
    try:
        carry = 0
        carries = []
        for call, n in enumerate((5000, 12345)):
            carries.append(carry)
            want_kidx, want = _reference_compaction(eng, call, _records(rng, n))
            assert want.carry_in == carry
            # the caller's own buffers: flags in, indices and header out
            keep = torch.from_numpy(np.where(want_kidx >= 0, 0, -1).astype(np.int32)).to(dev)
            kidx = torch.full((n,), 12345678, dtype=torch.int32, device=dev)
            hdr = torch.zeros(16, dtype=torch.int32, device=dev)
            stream = torch.cuda.current_stream(0)
            _native.check(lib.tk_compact_noalloc(keep.data_ptr(), n, BATCH, carry, kidx.data_ptr(),
                                                 hdr.data_ptr(), C.c_void_p(stream.cuda_stream)))
            torch.cuda.synchronize()
            assert np.array_equal(kidx.cpu().numpy(), want_kidx), f"call {call}: indices differ"
            # ... and both equal the definition, computed without any of the library's code: a kept
            # record's index is the number of kept records before it, a skipped one's is -1 - that
            flags = keep.cpu().numpy() >= 0
            before = np.cumsum(flags) - flags
            assert np.array_equal(kidx.cpu().numpy(), np.where(flags, before, -1 - before)), f"call {call}"
            n_kept = int(flags.sum())
            assert (want.n_kept, want.n_closed, want.carry_out) == (
                n_kept, (carry + n_kept) // BATCH, (carry + n_kept) % BATCH)
            got = hdr.cpu().numpy().astype(np.int64) & 0xFFFFFFFF
            assert (got[0], got[1], got[2], got[3], got[4], got[5]) == (
                want.n_records, want.n_kept, want.carry_in, want.carry_out, want.n_closed, BATCH)
            carry = int(want.carry_out)
        assert carries[1] != 0     # the second call really started from an open batch
    finally:
        eng.close()


def test_noalloc_rejects_oversized_runs():
    lib = _native.load()
    dev = torch.device("cuda", 0)
    buf = torch.zeros(16, dtype=torch.int32, device=dev)
    rc = lib.tk_compact_noalloc(buf.data_ptr(), (1 << 20) + 1, 8, 0, buf.data_ptr(), buf.data_ptr(), None)
    assert rc != 0
    assert b"at most" in lib.tk_last_error()
