"""ORACLE -- TEST INFRASTRUCTURE ONLY.  Never imported by torchkafka_b200/.

CPU restatement of the reference hot path, used as the checker in tests/,
``__graft_entry__.smoke()`` and as the timed CPU arm of ``bench.py``:

    KafkaDataset.__iter__   /root/reference/src/kafka_dataset.py:147-171
    _process bodies         /root/reference/README.md:41-43, 72-79 (+ BASELINE.md §3)
    None-skip filter        /root/reference/src/kafka_dataset.py:161-162
    batching / tail batch   torch/utils/data/_utils/fetch.py:27-45
    default_collate         torch/utils/data/_utils/collate.py:246-275 (torch.stack)
    auto_commit timing      /root/reference/src/auto_commit.py:55-58
    commit()                /root/reference/src/kafka_dataset.py:120-145 -> consumer.commit()

The numeric work of the reference lives in third-party code that is present on
every machine this runs on -- CPython's ``json`` C scanner and ``torch.tensor``
-- so the restatement *calls the same functions* rather than re-deriving them.
kafka-python (>=2.0.2, ``setup.py:9``) is absent everywhere; its documented
position/commit semantics are restated by ``torchkafka_b200.testing.stub_kafka``.

Parity pin: the reference ships no tests or golden vectors, so
this oracle is pinned against *outputs of the reference itself*, generated in
the build container by ``tests/golden/make_golden.py`` (which imports the
unmodified ``/root/reference/src`` over the stub) and committed under
``tests/golden/``; ``tests/test_oracle_golden.py`` replays them.

The pad collate (``pad_collate``) has no counterpart in the reference
(``default_collate`` raises on ragged samples); it is the
oracle-side definition from SURVEY.md §8c.
"""
from __future__ import annotations

import json
from typing import Callable, Iterable, Iterator, List, Optional, Tuple

import torch
from torch.utils.data._utils.collate import default_collate

IMAGENET_MEAN = (0.485, 0.456, 0.406)
IMAGENET_STD = (0.229, 0.224, 0.225)


# --------------------------------------------------------------------------- _process bodies
def process_json(dtype, min_size: int = 0):
    """README.md:72-79 body with a tensor result: json.loads -> len filter -> torch.tensor."""

    def _process(record):
        elements = json.loads(record.value)
        if len(elements) < min_size:
            return None
        return torch.tensor(elements, dtype=dtype)

    return _process


def process_object(fields, dtype):
    """JSON object with named numeric fields: the README.md:72-79 body with a
    dict payload -- json.loads, then the named members in schema order as one tensor."""
    fields = tuple(fields)

    def _process(record):
        d = json.loads(record.value)
        return torch.tensor([d[k] for k in fields], dtype=dtype)

    return _process


def process_raw(shape, dtype=torch.uint8):
    def _process(record):
        return torch.frombuffer(bytearray(record.value), dtype=dtype).view(*shape)

    return _process


def process_struct(fields):
    """Fixed-layout struct: one torch.frombuffer view per field, returned as a
    tuple; default_collate stacks every position of the tuple (collate.py tuple branch -> :275).
    A record of the wrong size raises, as frombuffer / view would."""
    layout, off = [], 0
    for _name, dtype, shape in fields:
        count = 1
        for k in shape:
            count *= k
        layout.append((dtype, count, off, tuple(shape)))
        off += count * torch.empty((), dtype=dtype).element_size()
    total = off

    def _process(record):
        buf = bytearray(record.value)
        if len(buf) != total:
            raise ValueError(f"record of {len(buf)} bytes, struct of {total}")
        out = []
        for dtype, count, o, shape in layout:
            # frombuffer needs the offset aligned to the element size: copy the field out first
            out.append(torch.frombuffer(bytearray(buf[o:o + count * torch.empty((), dtype=dtype).element_size()]),
                                        dtype=dtype, count=count).view(*shape) if shape else
                       torch.frombuffer(bytearray(buf[o:o + torch.empty((), dtype=dtype).element_size()]),
                                        dtype=dtype, count=1).view(()))
        return tuple(out)

    return _process


def process_image(h=224, w=224, c=3, mean=IMAGENET_MEAN, std=IMAGENET_STD):
    """HWC u8 -> CHW float /255 -> (x-mean)/std in f32 -> bf16 (RNE)."""
    mean_t = torch.tensor(mean, dtype=torch.float32).view(c, 1, 1)
    std_t = torch.tensor(std, dtype=torch.float32).view(c, 1, 1)

    def _process(record):
        x = torch.frombuffer(bytearray(record.value), dtype=torch.uint8).view(h, w, c)
        x = x.permute(2, 0, 1).float().div(255)
        return ((x - mean_t) / std_t).to(torch.bfloat16)

    return _process


# --------------------------------------------------------------------------- collates
def pad_collate(samples: List[torch.Tensor]):
    """pad_sequence(batch_first=True, padding_value=0) + lengths + mask."""
    lengths = torch.tensor([s.shape[0] for s in samples], dtype=torch.int32)
    padded = torch.nn.utils.rnn.pad_sequence(samples, batch_first=True, padding_value=0)
    mask = torch.arange(padded.shape[1])[None, :] < lengths[:, None]
    return padded, lengths, mask


# --------------------------------------------------------------------------- the loop
def iter_samples(consumer, process: Callable) -> Iterator:
    """kafka_dataset.py:156-162 -- drive the consumer, apply _process, drop None."""
    for record in consumer:
        sample = process(record)
        if sample is not None:
            yield sample


def iter_batches(samples: Iterator, batch_size: int, drop_last: bool, collate: Callable) -> Iterator:
    """fetch.py:27-45 -- pull batch_size kept samples; short tail unless drop_last."""
    while True:
        data = []
        for _ in range(batch_size):
            try:
                data.append(next(samples))
            except StopIteration:
                break
        if len(data) == 0 or (drop_last and len(data) < batch_size):
            return
        yield collate(data)


def run(consumer, process: Callable, batch_size: int, drop_last: bool = False,
        collate: Callable = default_collate) -> Iterator[Tuple[object, dict]]:
    """auto_commit.py:55-58 -- yield each batch, then commit the consumer position.

    Yields ``(batch, committed)`` where ``committed`` is the ``{partition: offset}``
    snapshot the commit that follows this batch writes (taken right after the
    batch was produced: no record is consumed between the yield and the commit).
    """
    for batch in iter_batches(iter_samples(consumer, process), batch_size, drop_last, collate):
        consumer.commit()
        yield batch, dict(consumer.committed_log[-1])


def positions_after(partitions, offsets, upto: int) -> dict:
    """per partition max(offset)+1 over records [0, upto]."""
    pos = {}
    for p, o in zip(partitions[:upto + 1], offsets[:upto + 1]):
        p = int(p)
        o = int(o) + 1
        if pos.get(p, -1) < o:
            pos[p] = o
    return pos
