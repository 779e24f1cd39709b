"""CPU tests (-m "not gpu"): libfsgpu.so loads and exports every symbol include/fsgpu.h declares
(no compute calls without a GPU), and the product refuses to run without a CUDA device instead
of falling back to anything."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "fsgpu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(fsgpu_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(F):
    lib = ctypes.CDLL(F.LIB_PATH)
    syms = header_symbols()
    assert len(syms) >= 35
    missing = [s for s in syms if not hasattr(lib, s)]
    assert not missing, missing
    # the Python binding's list is the header's list
    assert sorted(F.SYMBOLS) == syms


def test_no_cpu_fallback(F):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(F.FsgpuError, match="no CUDA device"):
        F.FS(16, 16, 0.1, 0.1)


def test_product_does_not_import_oracle():
    """only tests/, __graft_entry__.smoke() and bench.py's CPU legs may touch oracle/"""
    pkg = os.path.join(ROOT, "fluidsolver_b200")
    for dp, _, fns in os.walk(pkg):
        for fn in fns:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                txt = open(os.path.join(dp, fn)).read()
                assert "liboracle" not in txt and "oracle_lib" not in txt and "libfsref" not in txt, fn
