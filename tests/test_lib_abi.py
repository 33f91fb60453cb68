"""CPU: the C-ABI library builds, loads and exports every symbol include/nemit.h declares
(no compute calls: there is no GPU here)."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def built():
    import __graft_entry__ as g
    g.build()
    from nerf_emitter_b200 import lib
    return lib


def test_header_symbols_are_exported(built):
    text = open(os.path.join(ROOT, "include", "nemit.h")).read()
    declared = set(re.findall(r"\b(nemit_[a-z_0-9]+)\s*\(", text))
    assert declared, "no declarations parsed"
    assert declared == set(built.SYMBOLS), declared ^ set(built.SYMBOLS)
    lib = built.load()
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.nemit_version() >= 100


def test_struct_sizes_match_header(built):
    import ctypes as C
    # 22 parameter pointers, 8 level outputs: a layout drift between nemit.h and lib.py shows up here
    assert C.sizeof(built.NemitParams) == 22 * C.sizeof(C.c_void_p)
    assert C.sizeof(built.NemitLevelsOut) == 8 * C.sizeof(C.c_void_p)
    assert C.sizeof(built.NemitGrid) == 8 + 4 * 16


def test_argument_errors_without_gpu(built):
    import ctypes as C
    lib = built.load()
    h = C.c_void_p()
    assert lib.nemit_create(None, 0, C.byref(h)) == -1
    assert b"null" in lib.nemit_last_error()
    cfg = built.NemitConfig()
    assert lib.nemit_create(C.byref(cfg), 0, C.byref(h)) == -1      # unsupported (zeroed) shapes
    assert lib.nemit_destroy(None) == 0


def test_product_path_has_no_oracle_import():
    pkg = os.path.join(ROOT, "nerf_emitter_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert "oracle" not in src.replace("# oracle", ""), fn
