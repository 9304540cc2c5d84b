"""The drop-in boundary: libzkb.so loads and exports every symbol include/zkb.h declares, the ctypes mirror
binds exactly that set, and without a GPU the library fails loudly instead of falling back to the CPU.
No compute calls here (CPU box)."""
import ctypes
import os
import re

import pytest

import zklaim_b200 as zk

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "zkb.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(zkb_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported():
    names = declared_symbols()
    assert len(names) >= 25
    lib = ctypes.CDLL(zk.LIB_PATH)
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing


def test_python_mirror_binds_the_header():
    assert sorted(zk.SYMBOLS) == declared_symbols()
    zk.load_library()


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(zk.ZkbError) as e:
        zk.Context(0)
    assert e.value.code == -1  # ZKB_ERR_CUDA
    assert "cuda" in str(e.value).lower()


def test_missing_library_is_an_error(tmp_path):
    with pytest.raises(zk.ZkbError):
        zk.load_library(str(tmp_path / "libzkb.so"))


def test_product_does_not_import_the_oracle():
    """oracle/ is test infrastructure: nothing under zklaim_b200/, bridge/ or include/ may reference it"""
    bad = []
    for top in ("zklaim_b200", "bridge", "include"):
        for d, _, files in os.walk(os.path.join(ROOT, top)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h", ".sh")) or f == "Makefile":
                    text = open(os.path.join(d, f), errors="ignore").read()
                    if re.search(r"(from|import)\s+oracle|oracle/|zkoracle|zkref", text):
                        bad.append(os.path.join(d, f))
    assert not bad, bad
