"""CPU-side checks of the C-ABI boundary: the library loads and exports every symbol
include/volcanosv_b200.h declares; without a GPU, context creation fails loudly
(no CPU fallback)."""
import ctypes
import os
import re

import pytest

import volcanosv_b200 as V
from volcanosv_b200 import lib as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "volcanosv_b200.h")).read()
    return sorted(set(re.findall(r"\b(vsv_[a-z_0-9]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    assert os.path.exists(V.LIB_PATH), "run python -m volcanosv_b200.build"
    lib = ctypes.CDLL(V.LIB_PATH)
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), n
    assert sorted(L.SIGNATURES) == names, "ctypes table and header disagree"
    L.load_library()


def test_version_string():
    lib = L.load_library()
    assert b"volcanosv_b200" in lib.vsv_version()


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(V.VsvError):
        V.Context(0)


def test_sig_struct_layout():
    assert V.SIG_DTYPE.itemsize == 32
    assert V.SIG_DTYPE.names == ("rec", "type", "pos", "len", "q0", "q1", "first", "count")
