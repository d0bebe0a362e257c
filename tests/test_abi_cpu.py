"""CPU tests of the drop-in boundary: libmsr.so builds, loads without a GPU, exports every symbol include/msr.h
declares, and refuses to compute without a device (no CPU fallback).  No compute calls."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "msr.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(msr_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported(lib_built):
    lib = ctypes.CDLL(lib_built)
    names = declared_symbols()
    assert len(names) >= 18
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/msr.h but not exported by libmsr.so"


def test_python_binding_covers_header(lib_built):
    from mllm_sparse_retrieval_b200 import _lib
    assert sorted(_lib.exported_symbols()) == declared_symbols()


def test_no_cpu_fallback(lib_built):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from mllm_sparse_retrieval_b200 import _lib
    lib = _lib.load()
    h = ctypes.c_void_p()
    rc = lib.msr_dense_index_create(0, 128, 0, ctypes.byref(h))
    assert rc == -2 and b"no CPU fallback" in lib.msr_last_error()
    from mllm_sparse_retrieval_b200.ops import SparseIndex
    import numpy as np
    with pytest.raises(Exception):
        SparseIndex(np.array([0, 1], np.uint64), np.array([0], np.uint32), np.array([1], np.uint16), 1, device="cpu")


def test_sources_never_import_the_oracle():
    """The product package must not reference oracle/ (a product path through the oracle voids parity)."""
    pkg = os.path.join(ROOT, "mllm_sparse_retrieval_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dp, f), errors="replace").read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f"{f} imports oracle"
                assert "liboracle" not in src
