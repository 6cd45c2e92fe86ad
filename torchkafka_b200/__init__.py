"""torchkafka_b200 -- B200-native ingest path behind torch-kafka's public API.

Exports mirror the reference package (``/root/reference/src/__init__.py:17-18``)
plus the additive schema types.  Importing this package never loads CUDA; the
C-ABI library is loaded on first use of a schema dataset and its absence is an
error there (no CPU fallback).
"""
from .kafka_dataset import KafkaDataset
from .auto_commit import auto_commit
from .schema import (ImageU8, JsonArray, JsonNested, JsonObject, PaddedBatch, PayloadError, RawBytes, StructBytes, TokenBatch,
                     TokenList)

__all__ = [
    "KafkaDataset", "auto_commit", "JsonArray", "JsonObject", "JsonNested", "TokenList", "RawBytes", "StructBytes", "ImageU8",
    "PaddedBatch", "TokenBatch", "PayloadError",
]
__version__ = "0.1.0"
