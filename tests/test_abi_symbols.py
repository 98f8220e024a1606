"""The C-ABI library loads and exports every symbol include/pcc_b200.h declares (no compute calls)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "pcc_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(pcc_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported_and_bound(pcc):
    names = _declared()
    assert len(names) >= 30
    lib = ctypes.CDLL(pcc._lib.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), f"{n} declared in pcc_b200.h but not exported"
    assert set(names) == set(pcc._lib.SIGNATURES), set(names) ^ set(pcc._lib.SIGNATURES)
    bound = pcc._lib.load()
    assert bound.pcc_version() >= 100
    assert bound.pcc_hash_capacity(1000) == 2048 and bound.pcc_hash_capacity(0) == 1024


def test_conv_args_struct_layout(pcc):
    A = pcc._lib.ConvArgs
    # offsets must match the C struct (8-byte pointers, natural alignment)
    assert A.in_.offset == 0 and A.n_in.offset == 8 and A.cin.offset == 16 and A.cout.offset == 20
    assert A.weight.offset == 24 and A.nbr_stride.offset == 48 and A.K.offset == 64
    assert A.film.offset == 80 and A.out.offset == 112 and A.mode.offset == 136 and A.weight_t.offset == 144
    assert A.order.offset == 152 and A.n_groups.offset == 160 and A.tile_order.offset == 168 and ctypes.sizeof(A) == 176


def test_product_does_not_import_oracle():
    """The product path must never route through the oracle."""
    pkg = os.path.join(ROOT, "learned-compression-of-point-cloud-geometry-and-attributes_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f
                assert "liboracle" not in src and not re.search(r"\borc_", src), f


def test_missing_library_fails_loudly(pcc, monkeypatch):
    import pytest
    monkeypatch.setattr(pcc._lib, "_lib", None)
    monkeypatch.setattr(pcc._lib, "LIB_PATH", "/nonexistent/libpcc_b200.so")
    with pytest.raises(pcc.PccError):
        pcc._lib.load()
