"""The C-ABI library loads and exports every symbol include/b200km.h declares, and the ctypes
prototype table agrees with the header (no compute calls: runs without a GPU)."""
import ctypes
import os
import re

import pytest

from dask_ml_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b200km.h")


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    out = {}
    for m in re.finditer(r"\b(int|size_t|const char \*)\s*(b200km_\w+)\s*\(([^;]*?)\)\s*;", src, flags=re.S):
        args = [a.strip() for a in m.group(3).split(",")]
        if args == ["void"]:
            args = []
        out[m.group(2)] = (m.group(1), args)
    return out


def test_library_is_built():
    assert os.path.exists(_lib.LIB_PATH), "run __graft_entry__.build() / make -C dask_ml_b200/csrc"


def test_every_declared_symbol_is_exported():
    decl = _declared()
    assert len(decl) >= 29
    lib = ctypes.CDLL(_lib.LIB_PATH)
    missing = [n for n in decl if not hasattr(lib, n)]
    assert not missing, missing


def test_prototype_table_matches_header():
    decl = _declared()
    assert set(decl) == set(_lib.PROTOTYPES), set(decl) ^ set(_lib.PROTOTYPES)
    for name, (ret, args) in decl.items():
        restype, argtypes = _lib.PROTOTYPES[name]
        assert len(args) == len(argtypes), name
        for a, t in zip(args, argtypes):
            if "*" in a:
                assert t is ctypes.c_void_p, (name, a)
            elif a.startswith("int64_t"):
                assert t is ctypes.c_int64, (name, a)
            elif a.startswith("uint64_t"):
                assert t is ctypes.c_uint64, (name, a)
            elif a.startswith("uint32_t"):
                assert t is ctypes.c_uint32, (name, a)
            elif a.startswith("double"):
                assert t is ctypes.c_double, (name, a)
            elif a.startswith("float"):
                assert t is ctypes.c_float, (name, a)
            elif a.startswith("int "):
                assert t is ctypes.c_int, (name, a)
            else:
                raise AssertionError((name, a))


def test_host_side_queries_and_error_strings():
    lib = _lib.load()
    assert lib.b200km_version() == 100
    assert lib.b200km_pack_bytes(1024, 64) > 1024 * 64 * (4 + 8 + 4)
    assert lib.b200km_pack_bytes(0, 4) == 0
    assert lib.b200km_acc_len(8, 16) == 8 * 16 + 8 + 1
    assert b"invalid" in lib.b200km_error_string(-1)
    with pytest.raises(ValueError):
        _lib.check(-2, "x")
    # argument validation happens before any CUDA call
    assert lib.b200km_assign_f32(None, 10, 4, 4, None, 3, None, None, None, None, 0, None) == -1
    assert lib.b200km_finalize(None, 1, 1, None, None, None, None) == -1


def test_no_cpu_fallback():
    import torch

    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    from dask_ml_b200._engine import get_engine

    with pytest.raises(_lib.B200LibraryError):
        get_engine()
