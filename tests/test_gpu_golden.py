"""GPU path against the committed golden fixtures (outputs of the unmodified reference,
tests/golden/), independent of the live oracle."""
import os

import numpy as np
import pytest
import torch

from golden_cases import CASES, build_log
from helpers import run_gpu
from torchkafka_b200.schema import ImageU8, JsonArray, RawBytes, TokenList

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def schema_for(case):
    if case["kind"] == "json":
        dt = getattr(torch, case["dtype"])
        if not case.get("pad"):
            return JsonArray(dt, fixed_len=8)
        cls = TokenList if dt in (torch.int32, torch.int64) else JsonArray
        return cls(dt, max_len=2048 if cls is TokenList else 256, min_size=case.get("min_size", 0))
    if case["kind"] == "raw":
        return RawBytes((64,), torch.uint8)
    h, w, c = case["hwc"]
    return ImageU8(h, w, c)


@pytest.mark.parametrize("name", sorted(CASES))
@pytest.mark.parametrize("packed", [True, False])
def test_gpu_reproduces_reference_outputs(name, packed):
    case = CASES[name]
    g = np.load(os.path.join(GOLD, f"{name}.npz"))
    schema = schema_for(case)
    got = run_gpu(build_log(case), schema, case["batch_size"], case.get("drop_last", False),
                  use_packed=packed, slot_bytes=64 << 10 if name.startswith("cfg2") else None)
    assert len(got) == int(g["n_batches"])
    for k, (batch, commit) in enumerate(got):
        want = g[f"b{k}_values"]
        if case.get("pad"):
            vals, lengths, mask = batch
            assert vals.shape == want.shape, (k, vals.shape, want.shape)
            assert np.array_equal(vals.numpy().view(np.uint32), want.view(np.uint32)), k
            assert np.array_equal(lengths.numpy(), g[f"b{k}_lengths"]), k
            assert torch.equal(mask, torch.arange(vals.shape[1])[None, :] < lengths[:, None])
        elif batch.dtype == torch.bfloat16:
            assert np.array_equal(batch.view(torch.int16).numpy(), want), k
        elif batch.dtype == torch.float32:
            assert np.array_equal(batch.numpy().view(np.uint32), want.view(np.uint32)), k
        else:
            assert np.array_equal(batch.numpy(), want), k
        assert commit == {p: int(o) for p, o in enumerate(g["commits"][k]) if o >= 0}, k
