"""TopicPartition / OffsetAndMetadata constructors: kafka-python's when it is importable (or
injected into ``sys.modules``, as the tests do), else the stub's.  What ``KafkaConsumer.commit``
takes: ``{TopicPartition: OffsetAndMetadata}``."""
from __future__ import annotations

import sys


def _offset_struct():
    """OffsetAndMetadata constructor of kafka-python if present, else of the stub."""
    mod = sys.modules.get("kafka.structs")
    cls = getattr(mod, "OffsetAndMetadata", None) if mod is not None else None
    if cls is None:
        try:
            from kafka.structs import OffsetAndMetadata as cls  # type: ignore
        except Exception:
            from .testing.stub_kafka import OffsetAndMetadata as cls
    nfields = len(getattr(cls, "_fields", ("offset", "metadata")))
    if nfields >= 3:        # kafka-python >= 2.1 added leader_epoch
        return lambda off: cls(off, "", -1)
    return lambda off: cls(off, "")


def _offset_make():
    """(make, rest) for building many OffsetAndMetadata at once without a Python-level call per
    item: ``map(make, zip(offsets, *rest))`` -- ``make`` is ``tuple.__new__`` bound to the
    namedtuple class (the class's own ``__new__`` and ``_make`` are Python functions, ~10x slower),
    ``rest`` the constant columns behind the offset (metadata, and leader_epoch for
    kafka-python >= 2.1)."""
    import functools
    import itertools
    probe = _offset_struct()(0)
    cls = type(probe)
    if not (isinstance(probe, tuple) and hasattr(cls, "_fields")):   # not a namedtuple: the constructor
        ctor = _offset_struct()
        return (lambda t: ctor(t[0])), ()
    return functools.partial(tuple.__new__, cls), tuple(itertools.repeat(v) for v in tuple(probe)[1:])


def _tp_struct():
    mod = sys.modules.get("kafka.structs")
    cls = getattr(mod, "TopicPartition", None) if mod is not None else None
    if cls is None:
        try:
            from kafka.structs import TopicPartition as cls  # type: ignore
        except Exception:
            from .testing.stub_kafka import TopicPartition as cls
    return cls
