"""The C-ABI library loads on a machine without a GPU and exports every symbol that
include/torchkafka_b200.h declares (no compute call is made here)."""
import ctypes
import os
import re

from torchkafka_b200 import _native

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "torchkafka_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(tk_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_agree():
    assert declared_symbols() == _native.exported_symbols()


def test_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(_native.LIB_PATH)
    for name in declared_symbols():
        assert hasattr(lib, name), name


def test_library_loads_and_reports_version_and_errors_without_a_gpu():
    lib = _native.load()
    assert lib.tk_abi_version() == 1
    # argument validation happens before any CUDA call
    assert lib.tk_slot_host_ptrs(None, 0, None, None) != 0
    assert b"null context" in lib.tk_last_error()


def test_pack_helper_module_is_built():
    from torchkafka_b200 import _tkpack
    assert callable(_tkpack.harvest)
