"""The C-ABI library loads on a CPU-only box and exports every symbol include/gk.h declares; the ctypes
signature table names exactly the same set (no compute calls here: there is no GPU)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "gk.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gk_[A-Za-z0-9_]+)\s*\(", src)))


def test_header_declares_entry_points():
    names = _declared()
    assert "gk_block_stats" in names and "gk_sbm_sweep" in names and "gk_eig_resample_F" in names
    assert len(names) >= 30


def test_library_exports_every_declared_symbol():
    from graphistician_b200 import build as gkbuild
    lib = ctypes.CDLL(gkbuild.build())
    missing = [n for n in _declared() if not hasattr(lib, n)]
    assert not missing, missing


def test_ctypes_table_matches_header():
    from graphistician_b200 import _lib
    assert sorted(_lib.SIGNATURES) == _declared()


def test_no_cpu_fallback_without_device():
    """gk_ctx_create must fail (not fall back) when no CUDA device is present."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA device present")
    from graphistician_b200 import _lib
    with pytest.raises(_lib.GKError):
        _lib.Context(0)


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "graphistician_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                txt = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), f
