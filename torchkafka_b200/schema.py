"""Payload schemas: the additive, optional declaration that routes a
``KafkaDataset`` subclass onto the B200 ingest path.

The reference has no such notion: every record goes through the user's Python
``_process`` (reference ``src/kafka_dataset.py:159``, bodies in
``README.md:41-43, 72-79``).  A schema states, declaratively, what that
``_process`` body would have computed, for the three payload families
BASELINE.json names, so that CUDA kernels can compute it instead:

===============  ===========================================================
schema           equivalent reference ``_process`` + collate
===============  ===========================================================
``JsonArray``    ``x = json.loads(record.value)``; ``None`` if
                 ``len(x) < min_size`` else ``torch.tensor(x, dtype=dtype)``;
                 batches are ``torch.stack`` (``fixed_len``) or pad + lengths +
                 mask (variable length; the reference's ``default_collate``
                 cannot batch ragged samples)
``TokenList``    ``JsonArray`` with an integer dtype, batches named
                 ``input_ids / lengths / attention_mask``
``JsonObject``   ``d = json.loads(record.value)``;
                 ``torch.tensor([d[k] for k in fields], dtype=dtype)``; batches are
                 ``torch.stack`` ([B, len(fields)])
``JsonNested``   ``torch.tensor(json.loads(record.value), dtype=dtype)`` of a declared
                 shape (arrays nested ``len(shape)`` deep); batches are ``torch.stack``
                 ([B, *shape])
``RawBytes``     ``torch.frombuffer(record.value, dtype=dtype).view(shape)``;
                 batches are ``torch.stack``
``StructBytes``  fixed-layout struct: field k is ``torch.frombuffer(value, dtype_k,
                 count_k, offset_k).view(shape_k)``; ``_process`` returns the tuple,
                 ``default_collate`` stacks every field
``ImageU8``      HWC uint8 -> ``permute(2,0,1).float().div(255)`` ->
                 ``(x - mean) / std`` in float32 -> ``bfloat16``; batches are
                 ``torch.stack`` ([B, C, H, W])
===============  ===========================================================

A dataset *without* ``payload_schema`` behaves exactly like the reference.
"""
from __future__ import annotations

import collections
import dataclasses
from typing import Optional, Sequence, Tuple

import torch

PaddedBatch = collections.namedtuple("PaddedBatch", ["values", "lengths", "mask"])
TokenBatch = collections.namedtuple("TokenBatch", ["input_ids", "lengths", "attention_mask"])

# element-kind codes shared with the C-ABI (include/torchkafka_b200.h, tk_dtype_t)
TK_F32, TK_F64, TK_I32, TK_I64 = 0, 1, 2, 3
_JSON_DTYPES = {
    torch.float32: TK_F32,
    torch.float64: TK_F64,
    torch.int32: TK_I32,
    torch.int64: TK_I64,
}


class PayloadError(ValueError):
    """A record's payload is outside its declared schema.

    The reference would raise from inside ``_process`` (``json.JSONDecodeError``,
    or ``RuntimeError``/``ValueError``/``TypeError`` from ``torch.tensor``); the
    GPU path has no exceptions on the device, so every record carries a status
    word and the first bad record of a batch is reported here with its Kafka
    coordinates.
    """

    def __init__(self, message, partition=None, offset=None, status=None):
        super().__init__(message)
        self.partition = partition
        self.offset = offset
        self.status = status


@dataclasses.dataclass(frozen=True)
class JsonArray:
    """A flat JSON array of numbers.

    dtype      torch.float32 / float64 / int32 / int64 (element conversion
               follows ``torch.tensor(json.loads(v), dtype=dtype)``).
    max_len    upper bound on the element count; longer records are an error.
    min_size   records with fewer elements are skipped (the ``None`` filter of
               ``README.md:76-77``).
    fixed_len  when set, every kept record must have exactly this many elements
               and a batch is a plain ``[B, fixed_len]`` tensor (``torch.stack``
               semantics); otherwise batches are ``PaddedBatch``.
    pad_to_max_len  return ``[B, max_len]`` instead of ``[B, longest in batch]``.
    """

    dtype: torch.dtype = torch.float32
    max_len: int = 256
    min_size: int = 0
    fixed_len: Optional[int] = None
    pad_to_max_len: bool = False

    def __post_init__(self):
        if self.dtype not in _JSON_DTYPES:
            raise TypeError(f"JsonArray does not support dtype {self.dtype}")
        if self.fixed_len is not None:
            object.__setattr__(self, "max_len", int(self.fixed_len))
        if self.max_len <= 0:
            raise ValueError("max_len must be positive")

    @property
    def kind(self) -> int:
        return _JSON_DTYPES[self.dtype]

    def wrap(self, values, lengths, mask):
        if self.fixed_len is not None:
            return values
        return PaddedBatch(values, lengths, mask)


@dataclasses.dataclass(frozen=True)
class TokenList(JsonArray):
    """JSON list of integer token ids -> ``TokenBatch``."""

    dtype: torch.dtype = torch.int32
    max_len: int = 2048

    def wrap(self, values, lengths, mask):
        if self.fixed_len is not None:
            return values
        return TokenBatch(values, lengths, mask)


@dataclasses.dataclass(frozen=True)
class JsonObject:
    """A JSON object whose named members are numbers: ``fields`` are the member names, in the
    order of the columns of the ``[B, len(fields)]`` batch.

    The whole record must be valid JSON as ``json.loads`` sees it; members not named in
    ``fields`` may hold anything and are ignored.  A named member must be a number (or
    ``true`` / ``false``, which ``torch.tensor`` stores as 1 / 0); of duplicate members the last
    one counts; a missing one is an error (``KeyError`` in the reference).  Reported as errors
    rather than decoded: member names written with backslash escapes, values nested deeper
    than 32 levels, UTF-16/32 payloads.
    """

    fields: Sequence[str] = ("x",)
    dtype: torch.dtype = torch.float32

    def __post_init__(self):
        if self.dtype not in _JSON_DTYPES:
            raise TypeError(f"JsonObject does not support dtype {self.dtype}")
        names = tuple(str(f) for f in self.fields)
        if not 1 <= len(names) <= 64:
            raise ValueError("1 to 64 fields supported")
        if len(set(names)) != len(names):
            raise ValueError("field names must be distinct")
        for f in names:
            if '"' in f or "\\" in f or any(ord(ch) < 0x20 for ch in f):
                raise ValueError(f"field name {f!r} cannot appear unescaped in JSON")
        if sum(len(f.encode("utf-8", "surrogatepass")) for f in names) > 2048:
            raise ValueError("field names take more than 2048 bytes")
        object.__setattr__(self, "fields", names)

    @property
    def kind(self) -> int:
        return _JSON_DTYPES[self.dtype]

    def name_table(self):
        """(names back to back as bytes, offsets) for tk_ctx_set_object_fields."""
        enc = [f.encode("utf-8", "surrogatepass") for f in self.fields]
        offs = [0]
        for b in enc:
            offs.append(offs[-1] + len(b))
        return b"".join(enc), offs


@dataclasses.dataclass(frozen=True)
class JsonNested:
    """Nested JSON arrays of a fixed shape, numbers at the leaves: ``[[1, 2, 3], [4, 5, 6]]`` is a
    record of ``shape=(2, 3)``.  A record of another depth or extent is an error (``torch.tensor``
    raises on ragged lists, ``torch.stack`` on samples of different shapes)."""

    shape: Tuple[int, ...] = (1,)
    dtype: torch.dtype = torch.float32

    def __post_init__(self):
        if self.dtype not in _JSON_DTYPES:
            raise TypeError(f"JsonNested does not support dtype {self.dtype}")
        shape = tuple(int(k) for k in (self.shape if isinstance(self.shape, (tuple, list)) else (self.shape,)))
        if not 1 <= len(shape) <= 4 or any(k < 1 for k in shape):
            raise ValueError("1 to 4 dimensions, each of at least 1")
        n = 1
        for k in shape:
            n *= k
        if n > 1 << 24:
            raise ValueError("more than 2^24 elements per record")
        object.__setattr__(self, "shape", shape)

    @property
    def kind(self) -> int:
        return _JSON_DTYPES[self.dtype]

    @property
    def elements(self) -> int:
        n = 1
        for k in self.shape:
            n *= k
        return n


@dataclasses.dataclass(frozen=True)
class RawBytes:
    """Fixed-layout binary payload: ``prod(shape) * itemsize`` bytes per record."""

    shape: Tuple[int, ...] = (64,)
    dtype: torch.dtype = torch.uint8

    def __post_init__(self):
        object.__setattr__(self, "shape", tuple(int(s) for s in self.shape))

    @property
    def record_bytes(self) -> int:
        n = 1
        for s in self.shape:
            n *= s
        return n * torch.empty((), dtype=self.dtype).element_size()


@dataclasses.dataclass(frozen=True)
class StructBytes:
    """Fixed-layout binary record made of typed fields laid out back to back (a packed C struct,
    a numpy structured dtype without padding): ``fields`` is a sequence of
    ``(name, dtype, shape)``.  Batches are namedtuples with one ``[B, *shape]`` tensor per field.
    Byte order is the machine's (little endian), as with ``torch.frombuffer``."""

    fields: Sequence[Tuple[str, torch.dtype, Tuple[int, ...]]] = (("x", torch.float32, (1,)),)

    def __post_init__(self):
        norm = tuple((str(n), d, tuple(int(k) for k in (sh if isinstance(sh, (tuple, list)) else (sh,))))
                     for n, d, sh in self.fields)
        if not 1 <= len(norm) <= 8:
            raise ValueError("1 to 8 fields supported")
        object.__setattr__(self, "fields", norm)
        object.__setattr__(self, "_batch_type", collections.namedtuple("StructBatch", [n for n, _, _ in norm]))

    @property
    def field_bytes(self) -> Tuple[int, ...]:
        out = []
        for _, d, sh in self.fields:
            n = 1
            for k in sh:
                n *= k
            out.append(n * torch.empty((), dtype=d).element_size())
        return tuple(out)

    @property
    def field_offsets(self) -> Tuple[int, ...]:
        off, out = 0, []
        for b in self.field_bytes:
            out.append(off)
            off += b
        return tuple(out)

    @property
    def record_bytes(self) -> int:
        return sum(self.field_bytes)

    def wrap(self, *tensors):
        return self._batch_type(*tensors)


@dataclasses.dataclass(frozen=True)
class ImageU8:
    """HWC uint8 image -> normalised bfloat16 CHW (``[B, C, H, W]`` batches)."""

    height: int = 224
    width: int = 224
    channels: int = 3
    mean: Sequence[float] = (0.485, 0.456, 0.406)
    std: Sequence[float] = (0.229, 0.224, 0.225)

    def __post_init__(self):
        object.__setattr__(self, "mean", tuple(float(m) for m in self.mean))
        object.__setattr__(self, "std", tuple(float(s) for s in self.std))
        if len(self.mean) != self.channels or len(self.std) != self.channels:
            raise ValueError("mean/std must have one entry per channel")
        if not 1 <= self.channels <= 4:
            raise ValueError("1 to 4 channels supported")

    @property
    def record_bytes(self) -> int:
        return self.height * self.width * self.channels
