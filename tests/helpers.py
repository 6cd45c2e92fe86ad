"""Shared helpers for the parity tests: run the same RecordLog through the GPU path
(public API, C-ABI underneath) and through the oracle (oracle/reference_path.py)."""
import os
import sys

import torch
from torch.utils.data import DataLoader

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from torchkafka_b200 import KafkaDataset, auto_commit  # noqa: E402
from torchkafka_b200.schema import ImageU8, JsonArray, JsonNested, JsonObject, RawBytes, StructBytes  # noqa: E402
from torchkafka_b200.testing import stub_kafka as sk  # noqa: E402
from oracle import reference_path as ref  # noqa: E402


def make_dataset(log, schema, slot_bytes=None, use_packed=True, consumer_kwargs=None, **attrs):
    _Consumer = sk.KafkaConsumer if use_packed else sk.PollOnlyConsumer
    extra = dict(consumer_kwargs or {})

    class DS(KafkaDataset):
        payload_schema = schema

        @classmethod
        def new_consumer(cls, *a, **k):
            k.pop("enable_auto_commit", None)
            return _Consumer(*a, log=log, **extra, **k)

    if slot_bytes is not None:
        DS.device_slot_bytes = slot_bytes
    for k, v in attrs.items():
        setattr(DS, k, v)
    return DS("topic", group_id="g")


def run_gpu(log, schema, batch_size, drop_last=False, **kw):
    """-> list of (batch moved to CPU, {partition: offset} committed after it)."""
    ds = make_dataset(log, schema, **kw)
    out = []
    loader = DataLoader(ds, batch_size=batch_size, drop_last=drop_last)
    gen = auto_commit(loader)
    pending = None
    for batch in gen:
        if pending is not None:
            # the commit of the previous batch ran when this one was requested
            out.append((pending, dict(ds._consumer.committed_log[-1])))
        pending = to_cpu(batch)
    if pending is not None:
        out.append((pending, dict(ds._consumer.committed_log[-1])))
    assert len(ds._consumer.committed_log) == len(out)
    ds.close()
    return out


def to_cpu(batch):
    if isinstance(batch, torch.Tensor):
        return batch.cpu()
    return type(batch)(*[t.cpu() for t in batch])


def oracle_process(schema):
    if isinstance(schema, JsonArray):
        return ref.process_json(schema.dtype, schema.min_size)
    if isinstance(schema, JsonObject):
        return ref.process_object(schema.fields, schema.dtype)
    if isinstance(schema, JsonNested):
        return ref.process_json(schema.dtype, 0)           # torch.tensor(json.loads(v)) as it is
    if isinstance(schema, ImageU8):
        return ref.process_image(schema.height, schema.width, schema.channels, schema.mean, schema.std)
    if isinstance(schema, RawBytes):
        return ref.process_raw(schema.shape, schema.dtype)
    if isinstance(schema, StructBytes):
        return ref.process_struct(schema.fields)
    raise TypeError(schema)


def run_oracle(log, schema, batch_size, drop_last=False):
    consumer = sk.KafkaConsumer("topic", log=log)
    variable = isinstance(schema, JsonArray) and schema.fixed_len is None
    collate = ref.pad_collate if variable else ref.default_collate
    return list(ref.run(consumer, oracle_process(schema), batch_size, drop_last, collate))


def assert_same(gpu, cpu, schema):
    assert len(gpu) == len(cpu), f"{len(gpu)} GPU batches vs {len(cpu)} oracle batches"
    variable = isinstance(schema, JsonArray) and schema.fixed_len is None
    for k, ((gb, gc), (cb, cc)) in enumerate(zip(gpu, cpu)):
        if variable:
            vals, lengths, mask = cb
            assert gb[0].shape == vals.shape, (k, gb[0].shape, vals.shape)
            assert bit_equal(gb[0], vals), f"batch {k}: values differ"
            assert torch.equal(gb[1], lengths), f"batch {k}: lengths differ"
            assert torch.equal(gb[2], mask), f"batch {k}: mask differs"
        elif isinstance(schema, StructBytes):
            assert len(gb) == len(cb) == len(schema.fields)
            for (name, _, _), g, c in zip(schema.fields, gb, cb):
                assert bit_equal(g, c), f"batch {k}: field {name} differs"
        else:
            assert gb.shape == cb.shape, (k, gb.shape, cb.shape)
            assert gb.dtype == cb.dtype
            assert bit_equal(gb, cb), f"batch {k}: values differ"
        assert gc == cc, f"batch {k}: committed offsets differ\n gpu={gc}\n ref={cc}"


def bit_equal(a, b):
    """Bit-exact comparison (distinguishes -0.0 from 0.0, treats equal NaN payloads as equal)."""
    if a.dtype != b.dtype or a.shape != b.shape:
        return False
    if a.dtype.is_floating_point:
        it = {2: torch.int16, 4: torch.int32, 8: torch.int64}[a.element_size()]
        return torch.equal(a.contiguous().view(it), b.contiguous().view(it))
    return torch.equal(a, b)
