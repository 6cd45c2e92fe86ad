"""ctypes binding of libtorchkafka_b200.so (C ABI: include/torchkafka_b200.h).

The library is built in-tree by ``__graft_entry__.build()`` /
``make -C torchkafka_b200/csrc``.  There is deliberately no fallback: if the
shared object is missing or a call fails, an exception is raised.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libtorchkafka_b200.so")


class NativeError(RuntimeError):
    pass


class TkRec(C.Structure):
    _fields_ = [("start", C.c_uint32), ("partition", C.c_int32), ("offset", C.c_int64)]


class TkResult(C.Structure):
    _fields_ = [
        ("n_records", C.c_uint32), ("n_kept", C.c_uint32), ("carry_in", C.c_uint32),
        ("carry_out", C.c_uint32), ("n_closed", C.c_uint32), ("batch_size", C.c_uint32),
        ("n_errors", C.c_uint32), ("first_error_rec", C.c_uint32),
        ("first_error_status", C.c_uint32), ("reserved", C.c_uint32 * 7),
    ]


class TkPlane(C.Structure):
    _fields_ = [("base", C.c_void_p), ("row_bytes", C.c_uint64)]


class TkJob(C.Structure):
    _fields_ = [
        ("schema", C.c_int32), ("dtype", C.c_int32),
        ("max_len", C.c_uint32), ("min_size", C.c_uint32), ("fixed_len", C.c_uint32),
        ("height", C.c_uint32), ("width", C.c_uint32), ("channels", C.c_uint32),
        ("record_bytes", C.c_uint64), ("batch_size", C.c_uint32), ("prev_slot", C.c_int32),
        ("n_planes", C.c_int32), ("resident", C.c_int32),
        ("prev", TkPlane * 8), ("cur", TkPlane * 8),
        ("field_offset", C.c_uint32 * 8), ("field_bytes", C.c_uint32 * 8),
    ]


class TkBlob(C.Structure):
    _fields_ = [("start", C.c_uint32), ("length", C.c_uint32), ("partition", C.c_int32),
                ("reserved", C.c_uint32), ("fetch_offset", C.c_int64), ("offset_delta", C.c_int64)]


class TkRbSummary(C.Structure):
    _fields_ = [("n_records", C.c_uint32), ("n_batches", C.c_uint32), ("n_bytes", C.c_uint64),
                ("status", C.c_uint32), ("reserved", C.c_uint32)]


REC_FLAGS = 0x70000000

STATUS_NAMES = {
    0: "ok",
    1: "not a JSON array of numbers (json.JSONDecodeError in the reference)",
    2: "value does not fit the target dtype (overflow)",
    3: "more elements than the schema's max_len",
    4: "valid JSON outside the numeric-array schema (true/false/null, strings, nesting)",
    5: "raw record has the wrong byte count",
    6: "record length or shape differs from the schema's (torch.tensor / torch.stack would raise)",
    7: "record batch failed its CRC-32C or its record framing (CorruptRecordException in kafka-python)",
    8: "JSON object without one of the schema's fields (KeyError in the reference)",
    9: "the (decompressed) records of this fetch do not fit the staging slot; raise device_slot_bytes",
}

_p = C.POINTER
_SIGNATURES = {
    "tk_last_error": (C.c_char_p, []),
    "tk_abi_version": (C.c_int, []),
    "tk_ctx_create": (C.c_int, [C.c_int, C.c_uint64, C.c_uint32, C.c_int, C.c_uint32, C.c_uint32,
                                _p(C.c_void_p)]),
    "tk_ctx_destroy": (C.c_int, [C.c_void_p]),
    "tk_ctx_reset": (C.c_int, [C.c_void_p, C.c_void_p]),
    "tk_slot_host_ptrs": (C.c_int, [C.c_void_p, C.c_int, _p(C.c_void_p), _p(C.c_void_p)]),
    "tk_slot_device_ptrs": (C.c_int, [C.c_void_p, C.c_int, _p(C.c_void_p), _p(C.c_void_p)]),
    "tk_slot_submit": (C.c_int, [C.c_void_p, C.c_int, C.c_uint32, C.c_uint64]),
    "tk_slot_assume_resident": (C.c_int, [C.c_void_p, C.c_int, C.c_uint32, C.c_uint64]),
    "tk_slot_wait": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p]),
    "tk_slot_pack": (C.c_int, [C.c_void_p, C.c_int, C.c_uint32, C.c_uint64, C.c_void_p, C.c_void_p,
                               C.c_void_p, C.c_void_p, C.c_uint32, C.c_int, _p(C.c_uint64)]),
    "tk_slot_fill": (C.c_int, [C.c_void_p, C.c_int, C.c_uint32, C.c_uint64, C.c_void_p, C.c_void_p,
                               C.c_void_p, C.c_void_p, C.c_uint32, C.c_int, _p(C.c_uint64)]),
    "tk_h2d_peak_gbs": (C.c_int, [C.c_int, C.c_uint64, C.c_int, _p(C.c_double)]),
    "tk_json_scan": (C.c_int, [C.c_void_p, C.c_int, C.c_uint32, C.c_uint32, C.c_void_p]),
    "tk_raw_scan": (C.c_int, [C.c_void_p, C.c_int, C.c_uint64, C.c_void_p]),
    "tk_compact": (C.c_int, [C.c_void_p, C.c_int, C.c_uint32, C.c_void_p]),
    "tk_commit_offsets": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p]),
    "tk_carry_rows": (C.c_int, [C.c_void_p, C.c_int, C.c_int, _p(TkPlane), _p(TkPlane), C.c_void_p]),
    "tk_json_decode": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_uint32, C.c_uint32, _p(TkPlane),
                                 _p(TkPlane), _p(TkPlane), C.c_void_p]),
    "tk_json_ingest": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32,
                                 _p(TkPlane), _p(TkPlane), _p(TkPlane), C.c_void_p]),
    "tk_ctx_set_json_fused": (C.c_int, [C.c_void_p, C.c_int]),
    "tk_ctx_set_json_wide": (C.c_int, [C.c_void_p, C.c_int]),
    "tk_collate_raw": (C.c_int, [C.c_void_p, C.c_int, C.c_uint64, _p(TkPlane), C.c_void_p]),
    "tk_collate_field": (C.c_int, [C.c_void_p, C.c_int, C.c_uint64, C.c_uint32, C.c_uint32, _p(TkPlane), C.c_void_p]),
    "tk_ctx_set_object_fields": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]),
    "tk_json_object_decode": (C.c_int, [C.c_void_p, C.c_int, C.c_int, _p(TkPlane), C.c_void_p]),
    "tk_json_nested_decode": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_uint32, C.c_void_p, _p(TkPlane), C.c_void_p]),
    "tk_json_nested_scan": (C.c_int, [C.c_void_p, C.c_int, C.c_uint32, C.c_void_p, C.c_void_p]),
    "tk_json_nested_decode_flat": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_uint32, C.c_void_p, _p(TkPlane), C.c_void_p]),
    "tk_image_set_norm": (C.c_int, [C.c_void_p, C.c_int, _p(C.c_float), _p(C.c_float), C.c_void_p]),
    "tk_collate_image_u8_bf16_nchw": (C.c_int, [C.c_void_p, C.c_int, C.c_uint32, C.c_uint32, C.c_uint32,
                                                _p(TkPlane), C.c_void_p]),
    "tk_process_slot": (C.c_int, [C.c_void_p, C.c_int, _p(TkJob), C.c_void_p]),
    "tk_compact_noalloc": (C.c_int, [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p,
                                     C.c_void_p, C.c_void_p]),
    "tk_slot_submit_recordbatches": (C.c_int, [C.c_void_p, C.c_int, C.c_uint64, _p(TkBlob), C.c_uint32,
                                             _p(C.c_uint32)]),
    "tk_slot_submit_recordbatches_ext": (C.c_int, [C.c_void_p, C.c_int, C.c_uint64, _p(TkBlob), C.c_void_p,
                                                 C.c_uint32, _p(C.c_uint32)]),
    "tk_host_register": (C.c_int, [C.c_void_p, C.c_uint64]),
    "tk_host_unregister": (C.c_int, [C.c_void_p]),
    "tk_pack_set_affinity": (C.c_int, [C.c_void_p, C.c_int]),
    "tk_slot_recordbatch_result": (C.c_int, [C.c_void_p, C.c_int, _p(TkRbSummary)]),
    "tk_slot_record": (C.c_int, [C.c_void_p, C.c_int, C.c_uint32, _p(TkRec)]),
    "tk_crc32c": (C.c_uint32, [C.c_void_p, C.c_uint64, C.c_uint32]),
    "tk_ctx_set_timing": (C.c_int, [C.c_void_p, C.c_int]),
    "tk_slot_decode_ms": (C.c_int, [C.c_void_p, C.c_int, _p(C.c_float)]),
    "tk_slot_fetch_result": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p]),
    "tk_slot_result": (C.c_int, [C.c_void_p, C.c_int, _p(_p(TkResult)), _p(_p(C.c_uint32)),
                                 _p(_p(C.c_uint32)), _p(_p(C.c_int64)), _p(C.c_uint32)]),
    "tk_slot_debug_records": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                        C.c_uint32]),
}

_lib = None


def exported_symbols():
    """Names every build of the library must export (checked by the CPU test suite)."""
    return sorted(_SIGNATURES)


def load():
    """Load the shared library (once).  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise NativeError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C torchkafka_b200/csrc`.  There is no CPU fallback for schema datasets."
        )
    lib = C.CDLL(LIB_PATH)
    for name, (restype, argtypes) in _SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if a symbol is missing
        fn.restype = restype
        fn.argtypes = argtypes
    if lib.tk_abi_version() != 1:
        raise NativeError("libtorchkafka_b200.so has an unexpected ABI version")
    _lib = lib
    return lib


def check(rc: int):
    if rc != 0:
        msg = load().tk_last_error()
        raise NativeError(msg.decode("utf-8", "replace") if msg else f"native call failed ({rc})")
