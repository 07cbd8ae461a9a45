"""CPU: libb2c.so loads without a GPU and exports every symbol that
include/*.h declares; compute entry points fail cleanly (no crash) off-GPU."""
import ctypes as C
import os
import re

from b200conv3d import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    names = set()
    for h in ("b2c.h", "b2c_debug.h"):
        src = open(os.path.join(ROOT, "include", h)).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        names |= set(re.findall(r"\b(b2c_[a-z0-9_]+)\s*\(", src))
    return names


def test_every_declared_symbol_is_exported(product_path):
    lib = C.CDLL(product_path)
    syms = declared_symbols()
    assert len(syms) >= 25
    for s in sorted(syms):
        assert hasattr(lib, s), s


def test_binding_table_matches_header(product_path):
    hdr = {s for s in declared_symbols() if not s.startswith("b2c_debug")}
    assert hdr == set(_lib.EXPORTS)


def test_oracle_exports_same_abi(oracle_path):
    lib = C.CDLL(oracle_path)
    for s in _lib.EXPORTS:
        assert hasattr(lib, s), s


def test_version_and_strerror(product_path):
    lib = _lib._bind(product_path)
    assert lib.b2c_version() >= 1000
    assert lib.b2c_strerror(-2) == b"unsupported configuration"
    assert lib.b2c_strerror(0) == b"ok"


def test_invalid_arguments_are_rejected_without_a_gpu(product_path):
    lib = _lib._bind(product_path)
    assert lib.b2c_cast(None, None, 4, 0, 2, None) == -1
    d = _lib.ConvDesc()
    assert lib.b2c_conv3d_fwd(C.byref(d), None, None, None, None, None, 0, None) == -1
    assert lib.b2c_filter_packed_bytes(C.byref(d), 0) == 0
