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


def _tp_struct():
    mod = sys.modules.get("kafka.structs")
    cls = getattr(mod, "TopicPartition", None) if mod is not None else None
    if cls is None:
        try:
            from kafka.structs import TopicPartition as cls  # type: ignore
        except Exception:
            from .testing.stub_kafka import TopicPartition as cls
    return cls
