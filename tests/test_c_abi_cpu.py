"""CPU: the C-ABI library loads without a GPU and exports every entry point include/aphrodite_b200.h declares; the
host-side argument checks of the entry points answer with a status + message (no compute is launched)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "aphrodite_b200.h")
LIB = os.path.join(ROOT, "aphrodite-fork_b200", "lib", "libaphrodite_b200.so")


def _declared():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(aph_[a-z0-9_]+)\s*\(", text)))


def _lib():
    if not os.path.exists(LIB):
        pytest.skip("libaphrodite_b200.so not built (python __graft_entry__.py)")
    return ctypes.CDLL(LIB)


def test_every_declared_symbol_is_exported():
    names = _declared()
    assert len(names) >= 10, names
    lib = _lib()
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, f"declared in the header but not exported: {missing}"


def test_abi_version_and_kv_dtype_strings():
    lib = _lib()
    assert lib.aph_b200_abi_version() >= 1
    f = lib.aph_kv_dtype_from_string
    assert f(b"auto") == 0 and f(b"fp8") == f(b"fp8_e4m3") == 1 and f(b"fp8_e5m2") == 2 and f(b"int8") < 0


def test_argument_checks_return_status_and_message():
    lib = _lib()
    lib.aph_b200_last_error.restype = ctypes.c_char_p
    # bad sizes are rejected on the host before any CUDA call
    st = lib.aph_swap_blocks(None, None, None, ctypes.c_int64(-1), ctypes.c_int64(0), 0, None)
    assert st != 0 and b"swap_blocks" in lib.aph_b200_last_error()
    st = lib.aph_copy_blocks(None, None, None, 1, ctypes.c_int64(70000), ctypes.c_int64(16), None)
    assert st != 0 and b"65535" in lib.aph_b200_last_error()
    # zero work is a successful no-op
    assert lib.aph_copy_blocks(None, None, None, 0, ctypes.c_int64(0), ctypes.c_int64(16), None) == 0
